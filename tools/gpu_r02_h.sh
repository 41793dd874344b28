# round 2, 1 GPU: the whole GPU suite after the agreement / generator-family changes, and how long the R R' generator takes
mkdir -p gpurun_out
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_g1.log 2>&1; tail -12 gpurun_out/r02_gpu_suite_g1.log
python - <<'PY'
import time, numpy as np
import conjugate_gradient_b200 as cg
with cg.Context(0) as c:
    for n in (4096, 8192, 16384, 32768):
        for fam, fl in (("gram", cg.GEN_GRAM), ("spectrum", cg.GEN_SPECTRUM | cg.GEN_KAPPA_LOG2(10))):
            c.generate(n, 1, fl)
            t0 = time.perf_counter(); c.generate(n, 1, fl); dt = time.perf_counter() - t0
            st = c.solve()
            print(f"n={n} family={fam}: generate {dt*1e3:.1f} ms" + (f" = {2.0*n**3/dt/1e12:.1f} TFLOP/s FP64" if fam == "gram" else f" = {8.0*n*n/dt/1e9:.0f} GB/s") + f"; solve {st['iterations']} iterations, {st['solve_seconds']*1e3:.1f} ms, residual {st['final_residual']:.2e}", flush=True)
PY
