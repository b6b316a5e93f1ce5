"""Host-buffer entry point for GAE: pinned host arrays in, pinned host arrays out.

The design point of this package is device-resident rollouts (no copies at all).  When a
caller does keep the rollout in host memory -- the reference's situation, where
``compute_episodic_return`` receives NumPy arrays (tianshou/policy/base.py:337-346) -- the
cost is the PCIe transfer (22 B per transition against a ~40 us kernel), so the rows are
split into chunks and H2D copy / kernel / D2H copy of successive chunks overlap on three
streams.  Env streams are independent, so chunking along E needs no carry.
"""
from __future__ import annotations

import os
from typing import List, Tuple

import torch

from . import _lib as L


class GaeHostPipeline:
    """Reusable staging buffers + streams for ``[E, T]`` rollouts held in pinned host memory.

    Three streams, one per engine: every input copy is issued on the H2D stream in chunk order,
    every kernel on the compute stream, every output copy on the D2H stream, linked by events per
    staging buffer.  The H2D engine (the bottleneck direction: 14 of the 22 bytes per transition)
    therefore never idles behind a kernel or an output copy, and the link runs full duplex."""

    def __init__(self, E: int, T: int, chunks: int = 8, n_buffers: int = 3,
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
        nb = max(2, min(n_buffers, chunks))
        self.s_h2d, self.s_comp, self.s_d2h = (torch.cuda.Stream(self.device) for _ in range(3))
        f = lambda dt: [torch.empty(rows, T, dtype=dt, device=self.device) for _ in range(nb)]
        self.d_rew, self.d_vs, self.d_vn = f(torch.float32), f(torch.float32), f(torch.float32)
        self.d_term, self.d_trunc = f(torch.uint8), f(torch.uint8)
        self.d_adv, self.d_ret = f(torch.float32), f(torch.float32)
        ev = lambda: [torch.cuda.Event() for _ in range(nb)]
        self.ev_loaded, self.ev_kernel, self.ev_stored = ev(), ev(), ev()
        self.h2d_bytes = E * T * (4 * 3 + 2)
        self.d2h_bytes = E * T * 8
        self.launches = 0
        self.use_graph = True
        self.zero_copy = False
        # overlap=True: run() only enqueues on the three engine streams and does NOT order the
        # caller's stream behind them, so the input copies of call k+1 run under the output copies
        # of call k (the link is full duplex; H2D is the longer direction).  Outputs of every
        # enqueued call are complete after join() / synchronize().
        self.overlap = False
        self._next_buffer = 0
        self._graph: torch.cuda.CUDAGraph | None = None
        self._graph_key = None
        self._cap_stream = torch.cuda.Stream(self.device)

    def run(self, rew: torch.Tensor, v_s: torch.Tensor, v_next: torch.Tensor,
            terminated: torch.Tensor, truncated: torch.Tensor, adv_out: torch.Tensor,
            ret_out: torch.Tensor, gamma: float, gae_lambda: float) -> None:
        """All arguments are pinned host tensors ``[E, T]`` (f32 / uint8).  Returns after the
        work is ENQUEUED; call ``synchronize()`` (or record an event) before reading outputs.

        The copy / kernel / copy DAG of one pass (about a hundred nodes) is captured once per set
        of host buffers into a CUDA graph and replayed, so the host cost of a pass is one launch
        instead of ~12 API calls per chunk (which was the bottleneck: 3.2 ms eager vs 2.1 ms of
        link time)."""
        args = (rew, v_s, v_next, terminated, truncated, adv_out, ret_out)
        if self.zero_copy:
            # pinned host memory is mapped into the device address space (UVA): the persistent
            # kernel's bulk copies read the rollout straight over PCIe into shared memory and its
            # bulk stores write advantages / returns straight back -- no staging copies at all
            assert all(a.is_pinned() and a.is_contiguous() for a in args)
            E, T = self.E, self.T
            L.check(L.lib().ts_gae_rows(
                rew.data_ptr(), v_s.data_ptr(), v_next.data_ptr(), terminated.data_ptr(),
                truncated.data_ptr(), E, T, T, T, float(gamma), float(gae_lambda), L.TS_F32,
                L.TS_F32, L.TS_F32, adv_out.data_ptr(), ret_out.data_ptr(), L.stream()),
                "ts_gae_rows")
            self.launches += 1
            return
        if self.overlap:
            assert all(a.is_pinned() for a in args), "host buffers must be pinned"
            self._enqueue(*args, gamma, gae_lambda, join=False)
            return
        if not self.use_graph:
            self._enqueue(*args, gamma, gae_lambda)
            return
        key = tuple(a.data_ptr() for a in args) + (float(gamma), float(gae_lambda))
        if key != self._graph_key:
            assert all(a.is_pinned() for a in args), "host buffers must be pinned"
            self._enqueue(*args, gamma, gae_lambda)  # warm-up outside capture
            torch.cuda.current_stream(self.device).synchronize()
            n0 = self.launches
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self._cap_stream):
                self._enqueue(*args, gamma, gae_lambda)
            self._per_pass = self.launches - n0
            self._graph, self._graph_key = g, key
            self._keepalive = args
        self._graph.replay()
        self.launches += self._per_pass

    def _enqueue(self, rew, v_s, v_next, terminated, truncated, adv_out, ret_out, gamma,
                 gae_lambda, join: bool = True) -> None:
        lib = L.lib()
        cur = torch.cuda.current_stream(self.device)
        # whatever the caller enqueued before (e.g. the kernel that fills the inputs) comes first
        self.s_h2d.wait_stream(cur)
        if join:
            self.s_comp.wait_stream(cur)
            self.s_d2h.wait_stream(cur)
        cross_call = not join  # staging buffers may still be in use by the previous call
        T, nb = self.T, len(self.d_rew)
        first = self._next_buffer if cross_call else 0  # keep rotating across overlapped calls
        if cross_call:
            self._next_buffer = (first + len(self.bounds)) % nb
        for c, (lo, hi) in enumerate(self.bounds):
            b, n = (first + c) % nb, hi - lo
            with torch.cuda.stream(self.s_h2d):
                if c >= nb or cross_call:
                    self.s_h2d.wait_event(self.ev_kernel[b])  # inputs of chunk c - nb consumed
                self.d_rew[b][:n].copy_(rew[lo:hi], non_blocking=True)
                self.d_vs[b][:n].copy_(v_s[lo:hi], non_blocking=True)
                self.d_vn[b][:n].copy_(v_next[lo:hi], non_blocking=True)
                self.d_term[b][:n].copy_(terminated[lo:hi], non_blocking=True)
                self.d_trunc[b][:n].copy_(truncated[lo:hi], non_blocking=True)
                self.ev_loaded[b].record(self.s_h2d)
            with torch.cuda.stream(self.s_comp):
                self.s_comp.wait_event(self.ev_loaded[b])
                if c >= nb or cross_call:
                    self.s_comp.wait_event(self.ev_stored[b])  # outputs of chunk c - nb copied out
                L.check(lib.ts_gae_rows(
                    self.d_rew[b].data_ptr(), self.d_vs[b].data_ptr(), self.d_vn[b].data_ptr(),
                    self.d_term[b].data_ptr(), self.d_trunc[b].data_ptr(), n, T, T, T,
                    float(gamma), float(gae_lambda), L.TS_F32, L.TS_F32, L.TS_F32,
                    self.d_adv[b].data_ptr(), self.d_ret[b].data_ptr(),
                    self.s_comp.cuda_stream), "ts_gae_rows")
                self.launches += 1
                self.ev_kernel[b].record(self.s_comp)
            with torch.cuda.stream(self.s_d2h):
                self.s_d2h.wait_event(self.ev_kernel[b])
                adv_out[lo:hi].copy_(self.d_adv[b][:n], non_blocking=True)
                ret_out[lo:hi].copy_(self.d_ret[b][:n], non_blocking=True)
                self.ev_stored[b].record(self.s_d2h)
        if join:
            self.join()

    def join(self, stream: torch.cuda.Stream | None = None) -> None:
        """Order ``stream`` (default: the current one) behind everything enqueued so far."""
        cur = stream or torch.cuda.current_stream(self.device)
        for st in (self.s_h2d, self.s_comp, self.s_d2h):
            cur.wait_stream(st)

    def synchronize(self) -> None:
        self.join()
        torch.cuda.current_stream(self.device).synchronize()


# ---------------------------------------------------------------------------------------------
# The drop-in call with host arrays: BasePolicy.compute_episodic_return(NumPy in) -> NumPy out
# ---------------------------------------------------------------------------------------------
def _np_dtype(t: torch.dtype):
    import numpy as np
    return {torch.float32: np.float32, torch.float64: np.float64, torch.uint8: np.uint8}[t]


class _Source:
    """One ``[E, T]`` input of the rows kernel: already on the device (no copy), pinned host
    memory (DMA straight from it), or pageable host memory (bounced through a pinned buffer by
    a host copy that overlaps the previous chunk's DMA)."""

    def __init__(self, arr, E: int, T: int, dtype: torch.dtype, device: torch.device) -> None:
        import numpy as np
        self.dtype = dtype
        if isinstance(arr, torch.Tensor) and arr.is_cuda:
            t = arr.detach()
            if t.dtype == torch.bool:
                t = t.view(torch.uint8)
            if t.dtype != dtype:
                t = t.to(dtype)
            self.kind, self.t = "device", t.reshape(E, T).contiguous()
            return
        if isinstance(arr, torch.Tensor):
            t = arr.detach()
        else:
            a = np.ascontiguousarray(arr)
            if a.dtype == np.bool_:
                a = a.view(np.uint8)
            t = torch.from_numpy(a)
        if t.dtype == torch.bool:
            t = t.view(torch.uint8)
        if t.dtype != dtype:
            t = t.to(dtype)  # host-side widening (e.g. f64 rewards with f32 values stay f64 upstream)
        self.t = t.reshape(E, T)
        self.kind = "pinned" if self.t.is_pinned() else "pageable"

    @property
    def host_bytes_per_row(self) -> int:
        return 0 if self.kind == "device" else self.t.shape[1] * self.t.element_size()


class GaeHostCall:
    """``compute_episodic_return`` for callers that hold the critic values (and possibly the
    rollout) as host arrays and want NumPy back, as the reference's callers do
    (tianshou/policy/base.py:337-346, a2c.py:91-116).  The [E, T] rows are cut into chunks; the
    host-to-device copies of chunk c+1, the kernel of chunk c and the device-to-host copies of
    chunk c-1 run on three streams (one per engine), so a call costs about the longer of the two
    PCIe directions.  Outputs land in pinned host memory taken from torch's caching host
    allocator and are returned as NumPy views of it (no extra host copy; the block goes back to
    the allocator when the arrays are dropped)."""

    # Host bytes per chunk and direction.  Measured on the B200 box (profiles/r02_e2e.md): 8 / 12 /
    # 24 / 32 / 48 / 96 MB -> 4.13 / 3.82 / 3.60 / 3.63 / 3.76 / 3.99 ms per C2 call; smaller chunks
    # lose to the per-copy cost of the DMA engines (issuing the same pipeline from C++ did not help:
    # 2 / 4 / 8 / 16 MB -> 5.3 / 4.4 / 4.3 / 3.8 ms), larger ones to the un-overlapped first upload
    # and last download of a synchronous call.
    CHUNK_BYTES = int(os.environ.get("TS_HOST_CALL_CHUNK_MB", "24")) << 20

    def __init__(self, device: torch.device) -> None:
        L.lib()
        self.device = device
        self.s_h2d, self.s_comp, self.s_d2h = (torch.cuda.Stream(device) for _ in range(3))
        self._dev: dict = {}      # staging on the device, by (name, slot)
        self._bounce: dict = {}   # pinned bounce buffers for pageable inputs, by (name, slot)
        self.h2d_bytes = self.d2h_bytes = self.launches = 0

    def _dev_buf(self, name: str, slot: int, rows: int, T: int, dtype: torch.dtype) -> torch.Tensor:
        key = (name, slot)
        t = self._dev.get(key)
        if t is None or t.dtype != dtype or t.numel() < rows * T:
            t = torch.empty(rows * T, dtype=dtype, device=self.device)
            self._dev[key] = t
        return t[: rows * T].view(rows, T)

    def _bounce_buf(self, name: str, slot: int, rows: int, T: int, dtype: torch.dtype):
        key = (name, slot)
        t = self._bounce.get(key)
        if t is None or t.dtype != dtype or t.numel() < rows * T:
            t = torch.empty(rows * T, dtype=dtype).pin_memory()
            self._bounce[key] = t
        return t[: rows * T].view(rows, T)

    def run(self, E: int, T: int, rew, v_s, v_next, terminated, truncated, gamma: float,
            gae_lambda: float, rew_dtype: torch.dtype, val_dtype: torch.dtype,
            out_dtype: torch.dtype):
        """Returns ``(returns, advantage)`` as NumPy ``[E * T]`` arrays of ``out_dtype``."""
        dev = self.device
        src = {"rew": _Source(rew, E, T, rew_dtype, dev), "vs": _Source(v_s, E, T, val_dtype, dev),
               "vn": _Source(v_next, E, T, val_dtype, dev),
               "term": _Source(terminated, E, T, torch.uint8, dev),
               "trunc": _Source(truncated, E, T, torch.uint8, dev)}
        osz = 8 if out_dtype == torch.float64 else 4
        in_row = sum(s.host_bytes_per_row for s in src.values())
        out_row = 2 * T * osz
        rows = max(1, min(E, self.CHUNK_BYTES // max(in_row, out_row, 1)))
        chunks = (E + rows - 1) // rows
        rows = (E + chunks - 1) // chunks  # even chunks
        # The call is synchronous: the first chunk's upload and the last chunk's download overlap
        # with nothing, so both ends taper (1/4, 1/4, 1/2 of a chunk) to keep them short.
        taper = os.environ.get("TS_HOST_CALL_TAPER", "0") == "1"
        bounds, lo = [], 0
        for c in range(chunks):
            hi = min(E, lo + rows)
            n = hi - lo
            if taper and chunks >= 2 and n >= 8 and c in (0, chunks - 1):
                parts = [n // 4, n // 4, n - 2 * (n // 4)]
                if c == chunks - 1:
                    parts.reverse()
                for p_ in parts:
                    bounds.append((lo, lo + p_))
                    lo += p_
            else:
                bounds.append((lo, hi))
                lo = hi
        nb = 3
        h_ret = torch.empty(E, T, dtype=out_dtype, pin_memory=True)
        h_adv = torch.empty(E, T, dtype=out_dtype, pin_memory=True)
        cur = torch.cuda.current_stream(dev)
        for st in (self.s_h2d, self.s_comp, self.s_d2h):
            st.wait_stream(cur)  # device-resident inputs may still be in flight on the caller's stream
        ev_loaded = [torch.cuda.Event() for _ in range(nb)]
        ev_kernel = [torch.cuda.Event() for _ in range(nb)]
        ev_stored = [torch.cuda.Event() for _ in range(nb)]
        ev_bounced: dict = {}
        lib = L.lib()
        code = lambda d: L.TS_F64 if d == torch.float64 else L.TS_F32  # noqa: E731
        for c, (lo, hi) in enumerate(bounds):
            n, b = hi - lo, c % nb
            ptrs = {}
            with torch.cuda.stream(self.s_h2d):
                if c >= nb:
                    self.s_h2d.wait_event(ev_kernel[b])  # chunk c - nb has consumed this slot
                for name, s in src.items():
                    if s.kind == "device":
                        ptrs[name] = s.t[lo:hi].data_ptr()
                        continue
                    d = self._dev_buf(name, b, rows, T, s.dtype)[:n]
                    if s.kind == "pinned":
                        d.copy_(s.t[lo:hi], non_blocking=True)
                    else:
                        # pageable: host copy into a pinned bounce buffer (multi-threaded for
                        # chunks this size), then DMA; the bounce buffer of slot b is free once
                        # the DMA that last read it has finished
                        hb = self._bounce_buf(name, b, rows, T, s.dtype)[:n]
                        ev = ev_bounced.get((name, b))
                        if ev is not None:
                            ev.synchronize()
                        hb.copy_(s.t[lo:hi])
                        d.copy_(hb, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(self.s_h2d)
                        ev_bounced[(name, b)] = ev
                    ptrs[name] = d.data_ptr()
                    self.h2d_bytes += n * s.host_bytes_per_row
                ev_loaded[b].record(self.s_h2d)
            d_adv = self._dev_buf("adv", b, rows, T, out_dtype)[:n]
            d_ret = self._dev_buf("ret", b, rows, T, out_dtype)[:n]
            with torch.cuda.stream(self.s_comp):
                self.s_comp.wait_event(ev_loaded[b])
                if c >= nb:
                    self.s_comp.wait_event(ev_stored[b])  # outputs of chunk c - nb copied out
                L.check(lib.ts_gae_rows(
                    ptrs["rew"], ptrs["vs"], ptrs["vn"], ptrs["term"], ptrs["trunc"], n, T, T, T,
                    float(gamma), float(gae_lambda), code(rew_dtype), code(val_dtype),
                    code(out_dtype), d_adv.data_ptr(), d_ret.data_ptr(),
                    self.s_comp.cuda_stream), "ts_gae_rows")
                self.launches += 1
                ev_kernel[b].record(self.s_comp)
            with torch.cuda.stream(self.s_d2h):
                self.s_d2h.wait_event(ev_kernel[b])
                h_adv[lo:hi].copy_(d_adv, non_blocking=True)
                h_ret[lo:hi].copy_(d_ret, non_blocking=True)
                ev_stored[b].record(self.s_d2h)
            self.d2h_bytes += n * out_row
        self.s_d2h.synchronize()  # the caller reads host memory next
        cur.wait_stream(self.s_comp)
        return h_ret.view(-1).numpy(), h_adv.view(-1).numpy()


_host_calls: dict = {}


def gae_host_call(device: torch.device) -> GaeHostCall:
    key = device.index or 0
    hc = _host_calls.get(key)
    if hc is None:
        hc = _host_calls[key] = GaeHostCall(device)
    return hc
