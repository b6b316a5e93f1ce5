#!/usr/bin/env python
"""Flag round trip between GPU 0 and GPU 1 over NVLink peer memory, with the primitives the sharded PER
exchange uses (run under torchrun, 2 ranks):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29541 scripts/nvlink_pingpong.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert world == 2
    torch.cuda.set_device(rank)
    dev = torch.device(f"cuda:{rank}")
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    slot = symm_mem.empty(8, dtype=torch.float64, device=dev)
    slot.zero_()
    hdl = symm_mem.rendezvous(slot, dist.group.WORLD)
    peers = int(hdl.buffer_ptrs_dev)
    out = torch.zeros(1, dtype=torch.int64, device=dev)
    res = {}
    iters = 2000
    for mode, name in ((0, "release flag, acquire poll with 64 ns back-off (the exchange's wait)"),
                       (1, "full shard record: 3 relaxed stores + system fence + release flag"),
                       (2, "release flag, tight acquire poll"),
                       (3, "shard record without the fence: 3 relaxed stores + release flag")):
        best = None
        for rep in range(3):
            slot.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            L.check(lib.ts_debug_pingpong(slot.data_ptr(), peers, 1 - rank, iters, int(rank == 0),
                                          mode, out.data_ptr(), L.stream()), "pingpong")
            torch.cuda.synchronize()
            ns = int(out.item())
            assert ns > 0, "peer timeout"
            us = ns / iters / 1e3
            best = us if best is None else min(best, us)
        res[name] = round(best, 3)
    if rank == 0:
        print(json.dumps({"nvlink_flag_round_trip_us": res, "iters": iters}), flush=True)
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
