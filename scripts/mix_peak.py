#!/usr/bin/env python
"""What HBM delivers for the GAE kernel's read : write mix (14 B in, 8 B out per transition =
1.75 : 1), measured with plain torch streaming kernels on buffers far larger than L2: copy (1 : 1,
the protocol of MEASURED_PEAKS.json), add (2 : 1), addcmul (3 : 1) and a pure read (sum).
Context for roofline.frac, not a product path."""
import json

import torch


def t(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n)
    return best * 1e-3


def main():
    out = {}
    for mb in (64, 256, 1024):
        n = mb * (1 << 20) // 4
        a, b, c = (torch.randn(n, device="cuda") for _ in range(3))
        d = torch.empty_like(a)
        out[f"{mb}MB_per_array"] = {
            "copy_1r1w_GBps": 8 * n / t(lambda: d.copy_(a)) / 1e9,
            "add_2r1w_GBps": 12 * n / t(lambda: torch.add(a, b, out=d)) / 1e9,
            "addcmul_3r1w_GBps": 16 * n / t(lambda: torch.addcmul(a, b, c, out=d)) / 1e9,
            "sum_1r0w_GBps": 4 * n / t(lambda: a.sum()) / 1e9,
        }
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
