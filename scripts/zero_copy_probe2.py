#!/usr/bin/env python
"""Probe 2: kernel reads MAPPED PINNED inputs directly (no H2D copies), writes device memory; the outputs
leave through the copy engine in chunks.  C2 shape, f64 outputs, synchronous call."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tianshou_fork_b200 import _lib as L  # noqa: E402


def main():
    E, T = 4096, 2048
    lib = L.lib()
    dev = torch.device("cuda:0")
    pinned = [torch.from_numpy(a).pin_memory() for a in bench.rollout(E, T, 0)]
    s_k, s_d = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    for out_dt, code in ((torch.float64, L.TS_F64), (torch.float32, L.TS_F32)):
        h_adv = torch.empty(E, T, dtype=out_dt, pin_memory=True)
        h_ret = torch.empty(E, T, dtype=out_dt, pin_memory=True)
        d_adv = torch.empty(E, T, dtype=out_dt, device=dev)
        d_ret = torch.empty(E, T, dtype=out_dt, device=dev)
        for chunks in (4, 8, 16, 32, 64):
            rows = E // chunks
            evs = [torch.cuda.Event() for _ in range(chunks)]

            def call():
                for c in range(chunks):
                    lo, hi = c * rows, (c + 1) * rows
                    L.check(lib.ts_gae_rows(*[t[lo:hi].data_ptr() for t in pinned], rows, T, T, T,
                                            0.99, 0.95, 0, 0, code, d_adv[lo:hi].data_ptr(),
                                            d_ret[lo:hi].data_ptr(), s_k.cuda_stream), "gae")
                    evs[c].record(s_k)
                    with torch.cuda.stream(s_d):
                        s_d.wait_event(evs[c])
                        h_adv[lo:hi].copy_(d_adv[lo:hi], non_blocking=True)
                        h_ret[lo:hi].copy_(d_ret[lo:hi], non_blocking=True)
                s_d.synchronize()

            for _ in range(3):
                call()
            t0 = time.perf_counter()
            n = 10
            for _ in range(n):
                call()
            ms = (time.perf_counter() - t0) / n * 1e3
            print(json.dumps({"out": str(out_dt)[6:], "chunks": chunks, "ms": round(ms, 3)}), flush=True)


if __name__ == "__main__":
    main()
