#!/bin/bash
# A/B of exp_libs/lib_{old,new}.so on the sharded PER leg at N GPUs: ab_per_n.sh N
N=$1
cp tianshou_fork_b200/csrc/libts_b200.so /tmp/lib_keep.so
for name in ${LIBS:-old new old new}; do
  cp exp_libs/lib_$name.so tianshou_fork_b200/csrc/libts_b200.so; echo "lib=$name"
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 3 --skip-cpu --skip-e2e --skip-updates 2>/dev/null | python -c "import json,sys; b=json.load(sys.stdin); p=b['per']; print('strong', p['strong']['us_per_step'], 'weak', p['weak']['us_per_step'], p['weak']['parity'], p['parity'])"
done
cp /tmp/lib_keep.so tianshou_fork_b200/csrc/libts_b200.so
