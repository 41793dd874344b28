# round 2, text ingest with the spinning copy team: tests, then rate per (chunk, team)
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_ingest.py -x -q ) > gpurun_out/r02_ingest_tests.log 2>&1; tail -2 gpurun_out/r02_ingest_tests.log
run() { tag=$1; shift; env "$@" timeout 300 python tools/ingest_rate.py $N $CH > gpurun_out/r02_ingest_${tag}_n$N.json 2> gpurun_out/r02_ingest_${tag}_n$N.err; echo "$tag rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r02_ingest_${tag}_n$N.json").read())
print("$tag", [(r["chunk_mib"], r["seconds_best_of_3"]) for r in d["runs"]], "host parser", d["host_parser"]["seconds"], d["bit_identical_to_host_parser"])
PY
  awk 'NR%3==0' gpurun_out/r02_ingest_${tag}_n$N.err | head -6; }
N=8192 CH="8 4 16 2 8" run team_default CGB_X=0
N=8192 CH="8 4 16" run team8 CGB_INGEST_TEAM=8
N=8192 CH="8 16" run team12 CGB_INGEST_TEAM=12
N=8192 CH="8 4" run team4 CGB_INGEST_TEAM=4
N=16384 CH="8 16 4" run team_default CGB_X=0
N=16384 CH="8 16" run team8 CGB_INGEST_TEAM=8
