#!/bin/bash
# A/B of the shared-memory chunk order of the GAE kernels on one box: same process sequence twice.
cd tianshou_fork_b200/csrc
for round in 1 2; do
  for swz in 1 0; do
    rm -f gae.o libts_b200.so; make -s EXTRA=-DTS_SMEM_SWZ=$swz >/dev/null 2>&1
    cd ../..; echo "swz=$swz round=$round"; python scripts/gae_sweep.py --variants 4,5 --iters 1000 2>/dev/null; TS_BENCH_ROTATE=3 python bench.py --steps 1000 --warmup 5 --skip-per --skip-e2e --skip-cpu --skip-updates 2>/dev/null | python -c "import json,sys; b=json.load(sys.stdin); print('bench us', b['roofline']['us_per_launch'], b['roofline']['frac'])"; cd tianshou_fork_b200/csrc
  done
done
rm -f gae.o libts_b200.so; make -s >/dev/null 2>&1
