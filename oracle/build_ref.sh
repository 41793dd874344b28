#!/usr/bin/env bash
# oracle/build_ref.sh -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
#
# Compiles the REAL reference sources, from where they lie under
# /root/reference, into oracle/_ref/ (git-ignored, travels to the GPU box).
# No reference source is copied into the repo: each translation unit is piped
# from sed straight into gcc.  The only edits are the ones the reference itself
# requires to change problem size or precision:
#   - `#define ROWS 8192` / `#define COLS 8192` are unconditional macros
#     (serialConjugate.c:29-30, parallel_cg.c:30-31), so every n is its own build;
#   - the FP64 variant is the mechanical float->double, "%f%*c"->"%lf%*c",
#     MPI_FLOAT->MPI_DOUBLE rewrite.
# Products, per n and precision p in {f32,f64}:
#   serial_<p>_n<N>            the serial program        (CLI:  A.txt b.txt x0.txt)
#   libserial_<p>_n<N>.so      its conjugrad(A,b,x)      (main renamed away)
#   parallel_<p>_n<N>          the MPI program linked against oracle/mpi_shim
#   libparallel_<p>_n<N>.so    its conjugrad(A,b,x,m,rank,np) + the shim
#   libp2p_<p>_n<N>.so         the same from point-to-point_cg.c (MPI_Send / MPI_Recv through rank 0)
# Flags: gcc -O2 (no -march: no FMA contraction, matching a stock x86-64 build).
# Usage: oracle/build_ref.sh [n ...]      default: 2 4 512 2048 8192 16384
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ref="${CG_REFERENCE_DIR:-/root/reference}"
out="$here/_ref"
sizes=("$@"); [ ${#sizes[@]} -eq 0 ] && sizes=(2 4 512 2048 8192 16384)
if [ ! -f "$ref/serialConjugate.c" ] || [ ! -f "$ref/parallel_cg.c" ]; then
  echo "build_ref: $ref not present; keeping prebuilt oracle/_ref as is" >&2
  exit 0
fi
mkdir -p "$out"
CC="${CC:-gcc}"
CFLAGS="-O2 -w"
# xform <n> <f32|f64> <file>: the reference translation unit on stdout
xform() {
  local args=(-e "s/^#define ROWS 8192/#define ROWS $1/" -e "s/^#define COLS 8192/#define COLS $1/")
  if [ "$2" = f64 ]; then
    args+=(-e 's/float/double/g' -e 's/%f%\*c/%lf%*c/g' -e 's/MPI_FLOAT/MPI_DOUBLE/g')
  fi
  sed "${args[@]}" "$3"
}
$CC $CFLAGS -fPIC -c "$here/mpi_shim/mpi_shim.c" -I"$here/mpi_shim" -o "$out/mpi_shim.o"
for n in "${sizes[@]}"; do
  for p in f32 f64; do
    # serial program and library
    xform "$n" "$p" "$ref/serialConjugate.c" \
      | $CC $CFLAGS -x c - -o "$out/serial_${p}_n${n}" -lm
    xform "$n" "$p" "$ref/serialConjugate.c" \
      | $CC $CFLAGS -fPIC -shared -Wl,-Bsymbolic -Dmain=ref_main -x c - -o "$out/libserial_${p}_n${n}.so" -lm
    # MPI program and library over the shim
    xform "$n" "$p" "$ref/parallel_cg.c" \
      | $CC $CFLAGS -I"$here/mpi_shim" "$out/mpi_shim.o" -x c - -o "$out/parallel_${p}_n${n}" -lm -lpthread
    xform "$n" "$p" "$ref/parallel_cg.c" \
      | $CC $CFLAGS -fPIC -shared -Wl,-Bsymbolic -Dmain=ref_main -I"$here/mpi_shim" "$out/mpi_shim.o" -x c - \
          -o "$out/libparallel_${p}_n${n}.so" -lm -lpthread
    # the point-to-point twin (same conjugrad signature, collectives emulated through rank 0 with MPI_Send / MPI_Recv)
    if [ -f "$ref/point-to-point_cg.c" ]; then
      xform "$n" "$p" "$ref/point-to-point_cg.c" \
        | $CC $CFLAGS -fPIC -shared -Wl,-Bsymbolic -Dmain=ref_main -I"$here/mpi_shim" "$out/mpi_shim.o" -x c - \
            -o "$out/libp2p_${p}_n${n}.so" -lm -lpthread
    fi
  done
done
echo "build_ref: built reference binaries for n = ${sizes[*]} into $out"
