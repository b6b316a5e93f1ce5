#!/usr/bin/env python
"""GAE rows kernel over the shape sweep of SURVEY.md 8d (T x E grid, automatic kernel choice).

    python scripts/gae_shapes.py > gpurun_out/gae_shapes.jsonl
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tianshou_fork_b200 import _lib as L  # noqa: E402


def main():
    lib = L.lib()
    dev = torch.device("cuda:0")
    peak, _ = bench.measured_peaks()
    variants = [int(x) for x in os.environ.get("TS_VARIANTS", "0").split(",")]
    for T in (128, 512, 2048, 4096):
        for E in (64, 1024, 4096, 16384):
            d = [torch.from_numpy(x).to(dev) for x in bench.rollout(E, T, 0)]
            adv = torch.empty(E, T, device=dev)
            ret = torch.empty(E, T, device=dev)
            for v in variants:
                L.check(lib.ts_tune_set(b"gae_rows_variant", v), "tune")

                def step(s):
                    L.check(lib.ts_gae_rows(*[t.data_ptr() for t in d], E, T, T, T, 0.99, 0.95, 0,
                                            0, 0, adv.data_ptr(), ret.data_ptr(), s), "gae")
                side = torch.cuda.Stream(dev)
                with torch.cuda.stream(side):
                    for _ in range(3):
                        step(side.cuda_stream)
                side.synchronize()
                g = torch.cuda.CUDAGraph()
                calls = 8
                with torch.cuda.graph(g, stream=side):
                    for _ in range(calls):
                        step(side.cuda_stream)
                best = 1e9
                for _ in range(3):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(20):
                        g.replay()
                    e1.record()
                    torch.cuda.synchronize()
                    best = min(best, e0.elapsed_time(e1) * 1e3 / (20 * calls))
                gbs = E * T * 21 / best / 1e3
                print(json.dumps({"T": T, "E": E, "variant": v, "MB": E * T * 22 / 1e6,
                                  "us": round(best, 2), "alg_GBps": round(gbs, 1),
                                  "frac": round(gbs / peak, 3)}), flush=True)
    L.check(lib.ts_tune_set(b"gae_rows_variant", 0), "tune")


if __name__ == "__main__":
    main()
