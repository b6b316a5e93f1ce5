"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/ts_b200.h declares, and the product path refuses to run without a GPU (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ts_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ts_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from tianshou_fork_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in ts_b200.h but not exported"
    # and the Python binding table covers exactly the header
    assert sorted(_lib.SIGNATURES) == declared


def test_version_and_error_strings_without_gpu():
    from tianshou_fork_b200 import _lib
    lib = _lib.load()
    assert lib.ts_version() == 100
    assert lib.ts_error_string(0) == b"ok"
    assert b"workspace" in lib.ts_error_string(-2)
    assert lib.ts_gae_flat_workspace_bytes(1) > 0
    assert lib.ts_ppo_loss_workspace_bytes(1) > 0


def test_argument_validation_needs_no_gpu():
    """Bad arguments are rejected on the host before any launch."""
    from tianshou_fork_b200 import _lib
    lib = _lib.load()
    assert lib.ts_gae_rows(None, None, None, None, None, 4, 8, 8, 8, 0.99, 0.95, 0, 0, 0, None,
                           None, None) == -1  # null pointers
    assert lib.ts_gae_rows(None, None, None, None, None, 4, 8, 4, 8, 0.99, 0.95, 0, 0, 0, None,
                           None, None) == -1  # stride < T
    assert lib.ts_gae_rows(None, None, None, None, None, 0, 8, 8, 8, 0.99, 0.95, 0, 0, 0, None,
                           None, None) == 0  # empty is a no-op
    assert lib.ts_segtree_prefix_idx(None, 0, None, 1, 4, 0, None, None, None) == -1
    assert lib.ts_nstep_chain(None, 4, None, 0, None, None, None, 1, 1, None, None, 0.99, 0,
                              None, None, None, None) == -1  # n_step < 1


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_path_fails_loudly_without_gpu():
    import tianshou_fork_b200 as ts
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.SegmentTree(8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.VectorReplayBuffer(16, 2)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "tianshou_fork_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "ts_oracle" not in src, f
