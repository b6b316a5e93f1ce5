#!/usr/bin/env python
"""GPU-box microbenchmark of the PER kernels (BASELINE.json configs[3] shape): CUDA-graph replay
of sample-only, update-only and sample+update steps, for both tier group sizes."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tianshou_fork_b200 import _lib as L  # noqa: E402


def main():
    lib = L.lib()
    dev = torch.device("cuda:0")
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    size = bound = 1 << lg
    rng = np.random.default_rng(0)
    tree = torch.zeros(2 * bound, dtype=torch.float64, device=dev)
    stamp = torch.zeros(size, dtype=torch.int32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    prio = torch.ones(2, dtype=torch.float64, device=dev)
    leaves = torch.from_numpy((np.abs(rng.standard_normal(size)) + 1e-7) ** 0.6).to(dev)
    all_idx = torch.arange(size, device=dev)
    st0 = torch.cuda.current_stream().cuda_stream
    L.check(lib.ts_segtree_update(tree.data_ptr(), size, bound, all_idx.data_ptr(),
                                  leaves.data_ptr(), L.TS_F64, 1, size, stamp.data_ptr(),
                                  counter.data_ptr(), err.data_ptr(), st0), "fill")
    R = 8
    u = torch.from_numpy(rng.random((R, B))).to(dev)
    td = torch.from_numpy(rng.standard_normal((R, B)).astype(np.float32)).to(dev)
    idx = torch.from_numpy(rng.integers(0, size, (R, B))).to(dev)
    w = torch.empty(R, B, dtype=torch.float64, device=dev)

    def sample(r, s):
        L.check(lib.ts_per_sample(tree.data_ptr(), bound, u[r].data_ptr(), B, prio.data_ptr(),
                                  0.4, 1, idx[r].data_ptr(), w[r].data_ptr(), None,
                                  err.data_ptr(), s), "sample")

    def update(r, s):
        L.check(lib.ts_per_update(tree.data_ptr(), size, bound, idx[r].data_ptr(),
                                  td[r].data_ptr(), L.TS_F32, B, 0.6, prio.data_ptr(),
                                  stamp.data_ptr(), counter.data_ptr(), err.data_ptr(), s),
                "update")

    def sample_lazy(r, s):
        L.check(lib.ts_per_sample_lazy(tree.data_ptr(), bound, u[r].data_ptr(), B, prio.data_ptr(),
                                       0.4, 1, idx[r].data_ptr(), w[r].data_ptr(), err.data_ptr(),
                                       s), "sample_lazy")

    def update_lazy(r, s):
        L.check(lib.ts_per_update_lazy(tree.data_ptr(), size, bound, idx[r].data_ptr(),
                                       td[r].data_ptr(), L.TS_F32, B, 0.6, prio.data_ptr(),
                                       stamp.data_ptr(), counter.data_ptr(), err.data_ptr(), s),
                "update_lazy")

    def prefix_on(buf, r, s):
        L.check(lib.ts_segtree_prefix_idx(tree.data_ptr(), bound, buf[r].data_ptr(), L.TS_F64, B, 1,
                                          idx[r].data_ptr(), err.data_ptr(), s), "prefix")

    def timed(name, fn):
        side = torch.cuda.Stream(dev)
        with torch.cuda.stream(side):
            for r in range(R):
                fn(r, side.cuda_stream)
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for r in range(R):
                fn(r, side.cuda_stream)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / (50 * R))
        print(json.dumps({"what": name, "lg": lg, "batch": B, "us_per_call": best}), flush=True)

    import ctypes
    timed("sample+update", lambda r, s: (sample(r, s), update(r, s)))
    timed("sample_lazy+update_lazy", lambda r, s: (sample_lazy(r, s), update_lazy(r, s)))
    timed("update_lazy", update_lazy)
    timed("sample_lazy", sample_lazy)
    if os.environ.get("TS_PER_BENCH_SHORT"):
        L.check(lib.ts_segtree_finish_top(tree.data_ptr(), bound, st0), "finish")
        assert int(err.item()) == 0, int(err.item())
        return
    for group in (1, 0):
        L.check(lib.ts_tune_set(b"segtree_ranged", group), "tune")
        timed(f"update(ranged={group})", update)
        ph = (ctypes.c_uint64 * 8)()
        torch.cuda.synchronize()
        L.check(lib.ts_debug_update_phases(ph), "phases")
        t = list(ph)
        print(json.dumps({"phases_ns": {"zero": t[6] - t[0], "compact": t[7] - t[6],
                                        "pow+table": t[1] - t[7], "groups": t[2] - t[1],
                                        "ticket(last)": t[3] - t[2], "top": t[5] - t[3],
                                        "total": t[5] - t[0]}}), flush=True)
        ct = (ctypes.c_uint64 * 512)()
        L.check(lib.ts_debug_update_ctas(ct), "ctas")
        c = np.array(list(ct), dtype=np.int64).reshape(128, 4)
        t0 = c[:, 0].min()
        q = lambda a: [int(np.percentile(a, p)) for p in (0, 50, 90, 100)]
        print(json.dumps({"cta_ns(min,p50,p90,max)": {
            "start": q(c[:, 0] - t0), "scan_done": q(c[:, 1] - t0), "groups_done": q(c[:, 2] - t0),
            "ticket": q(c[:, 3] - t0), "scan_dur": q(c[:, 1] - c[:, 0]),
            "groups_dur": q(c[:, 2] - c[:, 1]), "slowest_cta": int(np.argmax(c[:, 3]))}}),
            flush=True)
    L.check(lib.ts_tune_set(b"segtree_ranged", 1), "tune")
    for cl in (8, 16):
        L.check(lib.ts_tune_set(b"per_cluster", cl), "tune")
        timed(f"sample(cluster={cl})", sample)
    L.check(lib.ts_tune_set(b"per_cluster", 0), "tune")
    timed("sample+update", lambda r, s: (sample(r, s), update(r, s)))
    for pdl in (0, 1):
        L.check(lib.ts_tune_set(b"pdl", pdl), "tune")
        timed(f"update(pdl={pdl})", update)
        timed(f"sample(pdl={pdl})", sample)
        timed(f"sample+update(pdl={pdl})", lambda r, s: (sample(r, s), update(r, s)))
    u2 = u.clone()
    timed("copy+prefix_idx", lambda r, s: (u2[r].copy_(u[r]), prefix_on(u2, r, s)))
    assert int(err.item()) == 0, int(err.item())


if __name__ == "__main__":
    main()
