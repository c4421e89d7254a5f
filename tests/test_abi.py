"""The C-ABI library loads and exports every symbol include/jaco_b200.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(REPO, "include", "jaco_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(jaco_[a-z0-9_]+)\s*\(", hdr)))


def test_header_declares_expected_entry_points():
    d = _declared()
    for name in ("jaco_create", "jaco_destroy", "jaco_reset", "jaco_step", "jaco_step_host", "jaco_get_state",
                 "jaco_set_state", "jaco_debug_substep", "jaco_stats", "jaco_last_error"):
        assert name in d


def test_library_exports_every_declared_symbol():
    from mujoco_jaco_b200 import _lib
    _lib.build()
    L = ctypes.CDLL(_lib.SO_PATH)
    for name in _declared():
        assert hasattr(L, name), f"{name} declared in include/jaco_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == sorted(set(_lib.SYMBOLS) & set(_declared()))


def test_create_rejects_bad_arguments_without_touching_the_gpu():
    from mujoco_jaco_b200 import _lib
    _lib.build()
    L = ctypes.CDLL(_lib.SO_PATH)
    h = ctypes.c_void_p()
    assert L.jaco_create(None, 0, 4, 2, 0, 0, ctypes.c_uint64(0), 0, ctypes.byref(h)) == -1      # JACO_EINVAL
    buf = (ctypes.c_double * 8)()
    assert L.jaco_create(buf, 64, 4, 2, 0, 0, ctypes.c_uint64(0), 0, ctypes.byref(h)) == -1      # wrong blob size
    assert L.jaco_step(None, None, None, None, None, None, None, None, 0, None) == -1


def test_product_never_imports_the_oracle():
    for root, _, files in os.walk(os.path.join(REPO, "mujoco_jaco_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(root, f)).read()
                assert "oracle" not in src.lower() or f in ("compiler.py", "jaco_model.h", "jaco_sim.cuh"), f
                assert "import oracle" not in src and "from oracle" not in src, f


def test_vecenv_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mujoco_jaco_b200.env import JacoVecEnv
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        JacoVecEnv(4)
