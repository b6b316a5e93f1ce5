"""Host-buffer entry point for GAE: pinned host arrays in, pinned host arrays out.

The design point of this package is device-resident rollouts (no copies at all).  When a
caller does keep the rollout in host memory -- the reference's situation, where
``compute_episodic_return`` receives NumPy arrays (tianshou/policy/base.py:337-346) -- the
cost is the PCIe transfer (22 B per transition against a ~40 us kernel), so the rows are
split into chunks and H2D copy / kernel / D2H copy of successive chunks overlap on three
streams.  Env streams are independent, so chunking along E needs no carry.
"""
from __future__ import annotations

from typing import List, Tuple

import torch

from . import _lib as L


class GaeHostPipeline:
    """Reusable staging buffers + streams for ``[E, T]`` rollouts held in pinned host memory."""

    def __init__(self, E: int, T: int, chunks: int = 8, n_streams: int = 3,
                 device: torch.device | None = None) -> None:
        L.lib()
        self.E, self.T = E, T
        self.device = device or torch.device(f"cuda:{torch.cuda.current_device()}")
        chunks = max(1, min(chunks, E))
        base, rem = divmod(E, chunks)
        self.bounds: List[Tuple[int, int]] = []
        lo = 0
        for c in range(chunks):
            hi = lo + base + (1 if c < rem else 0)
            self.bounds.append((lo, hi))
            lo = hi
        rows = max(hi - lo for lo, hi in self.bounds)
        self.streams = [torch.cuda.Stream(self.device) for _ in range(n_streams)]
        f = lambda dt: [torch.empty(rows, T, dtype=dt, device=self.device)
                        for _ in range(n_streams)]
        self.d_rew, self.d_vs, self.d_vn = f(torch.float32), f(torch.float32), f(torch.float32)
        self.d_term, self.d_trunc = f(torch.uint8), f(torch.uint8)
        self.d_adv, self.d_ret = f(torch.float32), f(torch.float32)
        self.h2d_bytes = E * T * (4 * 3 + 2)
        self.d2h_bytes = E * T * 8
        self.launches = 0

    def run(self, rew: torch.Tensor, v_s: torch.Tensor, v_next: torch.Tensor,
            terminated: torch.Tensor, truncated: torch.Tensor, adv_out: torch.Tensor,
            ret_out: torch.Tensor, gamma: float, gae_lambda: float) -> None:
        """All arguments are pinned host tensors ``[E, T]`` (f32 / uint8).  Returns after the
        work is ENQUEUED; call ``synchronize()`` (or record an event) before reading outputs."""
        lib = L.lib()
        cur = torch.cuda.current_stream(self.device)
        start = torch.cuda.Event()
        start.record(cur)
        T = self.T
        for c, (lo, hi) in enumerate(self.bounds):
            s = c % len(self.streams)
            st = self.streams[s]
            n = hi - lo
            with torch.cuda.stream(st):
                if c < len(self.streams):
                    st.wait_event(start)
                self.d_rew[s][:n].copy_(rew[lo:hi], non_blocking=True)
                self.d_vs[s][:n].copy_(v_s[lo:hi], non_blocking=True)
                self.d_vn[s][:n].copy_(v_next[lo:hi], non_blocking=True)
                self.d_term[s][:n].copy_(terminated[lo:hi], non_blocking=True)
                self.d_trunc[s][:n].copy_(truncated[lo:hi], non_blocking=True)
                L.check(lib.ts_gae_rows(
                    self.d_rew[s].data_ptr(), self.d_vs[s].data_ptr(), self.d_vn[s].data_ptr(),
                    self.d_term[s].data_ptr(), self.d_trunc[s].data_ptr(), n, T, T, T,
                    float(gamma), float(gae_lambda), L.TS_F32, L.TS_F32, L.TS_F32,
                    self.d_adv[s].data_ptr(), self.d_ret[s].data_ptr(), st.cuda_stream),
                    "ts_gae_rows")
                self.launches += 1
                adv_out[lo:hi].copy_(self.d_adv[s][:n], non_blocking=True)
                ret_out[lo:hi].copy_(self.d_ret[s][:n], non_blocking=True)
        for st in self.streams:
            cur.wait_stream(st)

    def synchronize(self) -> None:
        torch.cuda.current_stream(self.device).synchronize()
