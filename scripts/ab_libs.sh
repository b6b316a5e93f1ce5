#!/bin/bash
# A/B of prebuilt libraries exp_libs/lib_<name>.so on one box: ab_libs.sh name1 name2 ...  (twice, alternating)
cp tianshou_fork_b200/csrc/libts_b200.so /tmp/lib_keep.so
for round in 1 2; do
  for name in "$@"; do
    cp exp_libs/lib_$name.so tianshou_fork_b200/csrc/libts_b200.so; echo "lib=$name round=$round"
    timeout 120 python scripts/gae_sweep.py --variants 4,5 --iters 500 2>/dev/null | cut -c1-110
    timeout 100 python scripts/gae_sweep.py --streams 16384 --variants 4 --iters 200 2>/dev/null | cut -c1-110
    TS_BENCH_ROTATE=3 timeout 200 python bench.py --steps 1000 --warmup 5 --skip-per --skip-e2e --skip-cpu --skip-updates 2>/dev/null | python -c "import json,sys; b=json.load(sys.stdin); print('bench us', b['roofline']['us_per_launch'], b['roofline']['frac'])"
  done
done
cp /tmp/lib_keep.so tianshou_fork_b200/csrc/libts_b200.so
