cp tianshou_fork_b200/csrc/libts_b200.so /tmp/lib_keep.so
for round in 1 2; do for name in old new; do cp exp_libs/lib_$name.so tianshou_fork_b200/csrc/libts_b200.so; echo "lib=$name"; timeout 200 python bench.py --steps 20 --warmup 3 --skip-cpu --skip-e2e --skip-updates 2>/dev/null | python -c "import json,sys; b=json.load(sys.stdin); p=b['per']; print('per us', p['us_per_step'], 'eager pair', p['eager_pair_us_per_step'])"; done; done
cp /tmp/lib_keep.so tianshou_fork_b200/csrc/libts_b200.so
timeout 300 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_segtree_per.py -m gpu -q -x 2>&1 | tail -2
