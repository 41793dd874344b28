#!/usr/bin/env bash
# Under torchrun: rank 0 runs its bench.py under ncu, the other ranks run it plainly.
# usage (as the torchrun "script"): tools/ncu_rank0.sh <out-prefix> <metrics> <bench.py args...>
# Only metric sets that ncu collects in ONE pass are usable: a replayed pass restores rank 0's device memory,
# peer window included, and the exchange-fused kernels would then wait for flags the peers never send again.
out=$1; metrics=$2; shift 2
if [ "${LOCAL_RANK:-0}" = "0" ]; then
  exec ncu --metrics "$metrics" --clock-control none -k regex:comm\|xrp\|update --launch-skip 30 -c 18 --csv --log-file "${out}.csv" \
       python bench.py "$@"
else
  exec python bench.py "$@"
fi
