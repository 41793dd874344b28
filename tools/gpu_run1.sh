for il in 0 1; do echo "interleave=$il"; CGB_L2_INTERLEAVE=$il timeout 600 python tools/sweep_l2.py 4096 5120 6144; done
