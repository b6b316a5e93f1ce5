#!/usr/bin/env python
"""GPU-box tuning sweep of the gae_rows kernel variants on BASELINE.json configs[1].

    python scripts/gae_sweep.py [--streams 4096] [--horizon 2048] > gpurun_out/gae_sweep.json
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tianshou_fork_b200 import _lib as L  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--streams", type=int, default=4096)
    p.add_argument("--horizon", type=int, default=2048)
    p.add_argument("--variants", default="1,2,3,4,5,6,7,8,9")
    p.add_argument("--iters", type=int, default=500)
    p.add_argument("--knobs", default="", help="semicolon list of key=value,... sets to A/B, "
                   "e.g. 'gae_balance=0,pdl=0;gae_balance=1,pdl=0;gae_balance=1,pdl=1'")
    a = p.parse_args()
    E, T = a.streams, a.horizon
    lib = L.lib()
    dev = torch.device("cuda:0")
    # TS_SWEEP_ROTATE distinct input / output sets, visited round robin (no L2 reuse across launches)
    rot = max(1, int(os.environ.get("TS_SWEEP_ROTATE", "3")))
    sets = [([torch.from_numpy(x).to(dev) for x in bench.rollout(E, T, s)],
             torch.empty(E, T, device=dev), torch.empty(E, T, device=dev)) for s in range(rot)]
    d, adv, ret = sets[0]
    turn = [0]
    st = torch.cuda.current_stream()
    peak, _ = bench.measured_peaks()

    def step():
        di, ai, ri = sets[turn[0] % rot]
        turn[0] += 1
        L.check(lib.ts_gae_rows(*[t.data_ptr() for t in di], E, T, T, T, 0.99, 0.95, 0, 0, 0,
                                ai.data_ptr(), ri.data_ptr(), st.cuda_stream), "gae")

    ref = None
    out = []
    knob_sets = [k for k in a.knobs.split(";") if k] or [""]
    for v, knobs in [(int(x), k) for x in a.variants.split(",") for k in knob_sets]:
        L.check(lib.ts_tune_set(b"gae_rows_variant", v), "tune")
        for kv in [x for x in knobs.split(",") if x]:
            key, val = kv.split("=")
            L.check(lib.ts_tune_set(key.encode(), int(val)), "tune " + key)
        adv.fill_(float("nan"))
        turn[0] = 0
        for _ in range(3 * rot):
            step()
        torch.cuda.synchronize()
        if ref is None:
            ref = adv.clone()
        err = float((adv - ref).abs().max())
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(a.iters):
                step()
            e1.record(st)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / a.iters)
        gbs = E * T * 21 / (best * 1e-6) / 1e9
        out.append({"variant": v, "knobs": knobs, "us": best, "alg_GBps": gbs, "frac": gbs / peak,
                    "max_abs_diff_vs_first": err})
        print(json.dumps(out[-1]), flush=True)
    L.check(lib.ts_tune_set(b"gae_rows_variant", 0), "tune")


if __name__ == "__main__":
    main()
