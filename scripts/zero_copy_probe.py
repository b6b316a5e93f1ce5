#!/usr/bin/env python
"""Probe: the GAE rows kernel reading / writing MAPPED PINNED host memory directly (no staging copies)
against the chunked copy / kernel / copy pipeline of the drop-in call.  C2 shape, f64 outputs."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tianshou_fork_b200 import _lib as L  # noqa: E402


def main():
    E, T = 4096, 2048
    lib = L.lib()
    dev = torch.device("cuda:0")
    arrs = bench.rollout(E, T, 0)  # rew, v_s, v_next, term, trunc (NumPy)
    pinned = [torch.from_numpy(a).pin_memory() for a in arrs]
    devs = [p.to(dev) for p in pinned]
    st = torch.cuda.current_stream()
    out = {}
    for out_dt, code in ((torch.float64, L.TS_F64), (torch.float32, L.TS_F32)):
        h_adv = torch.empty(E, T, dtype=out_dt, pin_memory=True)
        h_ret = torch.empty(E, T, dtype=out_dt, pin_memory=True)
        d_adv = torch.empty(E, T, dtype=out_dt, device=dev)
        d_ret = torch.empty(E, T, dtype=out_dt, device=dev)
        lib.ts_gae_rows(*[t.data_ptr() for t in devs], E, T, T, T, 0.99, 0.95, 0, 0, code,
                        d_adv.data_ptr(), d_ret.data_ptr(), st.cuda_stream)
        torch.cuda.synchronize()
        want = d_adv.cpu()
        for variant in (0, 5, 2, 7):
            L.check(lib.ts_tune_set(b"gae_rows_variant", variant), "tune")
            for name, ins, outs in (("in+out mapped", pinned, (h_adv, h_ret)),
                                    ("in mapped, out device", pinned, (d_adv, d_ret)),
                                    ("in device, out mapped", devs, (h_adv, h_ret))):
                def call():
                    L.check(lib.ts_gae_rows(*[t.data_ptr() for t in ins], E, T, T, T, 0.99, 0.95,
                                            0, 0, code, outs[0].data_ptr(), outs[1].data_ptr(),
                                            st.cuda_stream), "gae")
                    st.synchronize()
                h_adv.zero_()
                for _ in range(3):
                    call()
                ok = bool(torch.equal(outs[0].cpu(), want))
                t0 = time.perf_counter()
                n = 10
                for _ in range(n):
                    call()
                ms = (time.perf_counter() - t0) / n * 1e3
                out[f"{str(out_dt)[6:]} v{variant} {name}"] = {"ms": round(ms, 3), "ok": ok}
                print(json.dumps({f"{str(out_dt)[6:]} v{variant} {name}": out[f"{str(out_dt)[6:]} v{variant} {name}"]}), flush=True)
    L.check(lib.ts_tune_set(b"gae_rows_variant", 0), "tune")


if __name__ == "__main__":
    main()
