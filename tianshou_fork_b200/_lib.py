"""ctypes binding of ``libts_b200.so`` (the C-ABI declared in ``include/ts_b200.h``).

The shared object is built in-tree by ``csrc/Makefile`` (``__graft_entry__.build()``).  There
is NO fallback: if the library is missing, fails to load, or a CUDA device is absent, every
product call raises.  Pointers are raw device addresses taken from torch tensors; the stream
is torch's current CUDA stream, so calls order with surrounding torch work and are legal
inside ``torch.cuda.graph`` capture.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libts_b200.so")

TS_F32, TS_F64 = 0, 1
TS_FLAG_INDEX_RANGE, TS_FLAG_PREFIX_RANGE, TS_FLAG_PEER_TIMEOUT = 1, 2, 4

_vp, _i64, _i32, _f64, _f32, _sz = (
    ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_double, ctypes.c_float,
    ctypes.c_size_t,
)

# name -> (restype, argtypes); must list every symbol of include/ts_b200.h
SIGNATURES = {
    "ts_version": (_i32, []),
    "ts_error_string": (ctypes.c_char_p, [_i32]),
    "ts_last_cuda_error": (_i32, []),
    "ts_tune_set": (_i32, [ctypes.c_char_p, _i64]),
    "ts_debug_update_phases": (_i32, [_vp]),
    "ts_debug_update_ctas": (_i32, [_vp]),
    "ts_gae_rows": (_i32, [_vp] * 5 + [_i64] * 4 + [_f64, _f64, _i32, _i32, _i32, _vp, _vp,
                                                    _vp]),
    "ts_gae_flat_workspace_bytes": (_sz, [_i64]),
    "ts_gae_flat": (_i32, [_vp] * 9 + [_i64] * 3 + [_f64, _f64, _i32, _i32, _i32, _vp, _vp, _vp,
                                                   _sz, _vp]),
    "ts_gae_flat_static_max_n": (_i64, []),
    "ts_gae_flat_static": (_i32, [_vp] * 9 + [_i64] * 3 + [_f64, _f64, _i32, _i32, _i32, _vp, _vp,
                                                          _vp]),
    "ts_prev_index": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "ts_next_index": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "ts_unfinished_index": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "ts_mark_unfinished": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "ts_nstep_chain": (_i32, [_vp, _i64, _vp, _i32, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _f64,
                              _i32, _vp, _vp, _vp, _vp]),
    "ts_nstep_combine": (_i32, [_vp, _i32, _i64, _i64, _vp, _vp, _vp, _vp]),
    "ts_segtree_update": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _i64, _vp, _vp, _vp,
                                 _vp]),
    "ts_segtree_reduce": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "ts_segtree_prefix_idx": (_i32, [_vp, _i64, _vp, _i32, _i64, _i32, _vp, _vp, _vp]),
    "ts_per_sample": (_i32, [_vp, _i64, _vp, _i64, _vp, _f64, _i32, _vp, _vp, _vp, _vp, _vp]),
    "ts_per_sample_sharded": (_i32, [_vp, _i64, _vp, _i64, _vp, _i32, _i32, _vp, _f64, _i32, _vp,
                                     _vp, _vp, _vp, _vp]),
    "ts_per_publish": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _vp]),
    "ts_per_update_publish": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _f64, _vp, _vp, _vp,
                                     _vp, _vp, _vp, _i32, _i32, _vp]),
    "ts_per_sample_p2p": (_i32, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _i32, _vp, _f64, _i32,
                                 _vp, _vp, _vp, _vp]),
    "ts_per_update_publish_owned": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _f64, _vp, _vp,
                                           _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp]),
    "ts_per_sample_p2p_owned": (_i32, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _i32, _vp, _f64,
                                       _i32, _vp, _vp, _vp, _vp, _vp]),
    "ts_debug_pingpong": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "ts_per_normalize": (_i32, [_vp, _i64, _vp, _vp]),
    "ts_per_weight": (_i32, [_vp, _i64, _vp, _i64, _vp, _f64, _i32, _vp, _vp, _vp]),
    "ts_per_update": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _f64, _vp, _vp, _vp, _vp,
                             _vp]),
    "ts_segtree_lazy_top_nodes": (_i32, [_i64]),
    "ts_per_update_lazy": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _f64, _vp, _vp, _vp, _vp,
                                  _vp]),
    "ts_per_sample_lazy": (_i32, [_vp, _i64, _vp, _i64, _vp, _f64, _i32, _vp, _vp, _vp, _vp]),
    "ts_segtree_finish_top": (_i32, [_vp, _i64, _vp]),
    "ts_per_init_weight": (_i32, [_vp, _i64, _i64, _vp, _i64, _f64, _vp, _i32, _vp, _vp, _vp,
                                  _vp]),
    "ts_ppo_loss_workspace_bytes": (_sz, [_i64]),
    "ts_ppo_loss": (_i32, [_vp] * 7 + [_i64, _f32, _f32, _i32, _i32, _f32, _vp, _vp, _vp, _vp,
                                      _sz, _vp]),
    "ts_gather_rows": (_i32, [_vp, _i64, _i64, _vp, _i64, _vp, _vp]),
    "ts_gather_fields": (_i32, [_i32, _vp, _vp, _vp, _i64, _vp, _i64, _vp]),
    "ts_scatter_fields": (_i32, [_i32, _vp, _vp, _vp, _i64, _vp, _i64, _vp]),
    "ts_add_step": (_i32, [_i32, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _vp, _vp, _i64] + [_vp] * 11
                    + [_vp]),
    "ts_gather_stack": (_i32, [_vp, _i64, _vp, _i64, _i32, _vp, _i64, _i64, _vp, _vp, _vp, _vp,
                               _vp, _vp]),
}

_lib: Optional[ctypes.CDLL] = None
_cuda_checked = False


class TsError(RuntimeError):
    """A C-ABI call returned a negative TS_ERR_* code."""


def build(verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a (``make -C csrc``); nvcc cross-compiles on CPU."""
    cmd = ["make", "-C", CSRC, "-j8"] + ([] if verbose else ["-s"])
    subprocess.check_call(cmd)
    return LIB_PATH


SHIM_DIR = os.path.join(CSRC, "shim_build")
SHIM_NAME = "ts_b200_shim"
_shim = None
_shim_tried = False


def build_shim(verbose: bool = False) -> str:
    """Compile csrc/shim.cpp (the C++ host shim over the C-ABI: tensors in, tensors out, the PPO
    loss as a C++ autograd function) in-tree with torch's extension builder.  Host code only --
    no kernels -- linked against libts_b200.so next to it."""
    from torch.utils import cpp_extension
    os.makedirs(SHIM_DIR, exist_ok=True)
    cpp_extension.load(
        name=SHIM_NAME, sources=[os.path.join(CSRC, "shim.cpp")], build_directory=SHIM_DIR,
        extra_cflags=["-O2", "-std=c++17"], extra_include_paths=["/usr/local/cuda/include"],
        # $$ORIGIN: ninja's escape for a literal `$`; the quotes keep it from the shell
        extra_ldflags=[f"-L{CSRC}", "-lts_b200", "-Wl,-rpath,'$$ORIGIN/..'", "-lc10_cuda",
                       "-L/usr/local/cuda/lib64", "-lcudart"],
        with_cuda=False, is_python_module=False, verbose=verbose)
    return os.path.join(SHIM_DIR, SHIM_NAME + ".so")


def shim():
    """The compiled shim module, or None when it has not been built (the Python layer then makes
    the same C-ABI calls through ctypes).  Never compiles anything at run time."""
    global _shim, _shim_tried
    if not _shim_tried:
        _shim_tried = True
        path = os.path.join(SHIM_DIR, SHIM_NAME + ".so")
        if os.path.exists(path) and os.environ.get("TS_NO_SHIM") != "1":
            import importlib.machinery
            import importlib.util
            load()  # libts_b200.so first (the shim links against it)
            try:
                loader = importlib.machinery.ExtensionFileLoader(SHIM_NAME, path)
                spec = importlib.util.spec_from_loader(SHIM_NAME, loader)
                mod = importlib.util.module_from_spec(spec)
                loader.exec_module(mod)
                _shim = mod
            except Exception as e:  # pragma: no cover - reported once, ctypes path is used
                import warnings
                warnings.warn(f"ts_b200 shim failed to load ({e}); using the ctypes binding")
    return _shim


def load() -> ctypes.CDLL:
    """Load the library (does not need a GPU) and bind every declared symbol."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                "g.build()'` (there is no CPU or PyTorch fallback for this path)")
        lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def lib() -> ctypes.CDLL:
    """The library, for compute calls: additionally requires a CUDA device."""
    global _cuda_checked
    if not _cuda_checked:
        if not torch.cuda.is_available():
            raise RuntimeError(
                "tianshou_fork_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        _cuda_checked = True
    return _lib if _lib is not None else load()


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().ts_error_string(code).decode()
        raise TsError(f"{what}: {msg} (code {code})")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device address of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "C-ABI arguments must be contiguous CUDA tensors"
    return t.data_ptr()


try:  # raw handles straight from the C bindings: torch.cuda.current_stream() walks through
    # device-index parsing and availability checks in Python, ~15 us per call
    _raw_stream = torch._C._cuda_getCurrentRawStream
    _cur_device = torch._C._cuda_getDevice
except AttributeError:  # pragma: no cover - older/newer torch without the private hooks
    _raw_stream = _cur_device = None


def stream() -> int:
    """``cudaStream_t`` of torch's current stream on the current device."""
    if _raw_stream is not None:
        return _raw_stream(_cur_device())
    return torch.cuda.current_stream().cuda_stream


def dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return TS_F32
    if t.dtype == torch.float64:
        return TS_F64
    raise TypeError(f"expected float32/float64, got {t.dtype}")


def as_u8(t: torch.Tensor) -> torch.Tensor:
    """bool tensors are passed as uint8 views (same layout)."""
    return t.view(torch.uint8) if t.dtype == torch.bool else t
