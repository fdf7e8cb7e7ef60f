"""The C-ABI library loads without a GPU and exports exactly what include/*.h declares;
the product package never touches oracle/ and has no CPU fallback."""
import ctypes
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-gpu-test_b200")


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        names += re.findall(r"B200DS_API\s+[\w\s\*]+?\b(b200ds_\w+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_something():
    names = declared_symbols()
    assert len(names) >= 20 and "b200ds_adam_step" in names and "b200ds_tiny_train_steps_f32" in names


def test_library_exports_every_declared_symbol():
    from b200ds import _lib
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    dll = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(dll, n)]
    assert not missing, missing


def test_python_prototypes_cover_the_header():
    from b200ds import _lib
    assert sorted(_lib.PROTOTYPES) == declared_symbols()
    lib = _lib.load()
    assert lib.b200ds_version() >= 100
    assert isinstance(lib.b200ds_last_error(), bytes)


def test_argument_errors_do_not_need_a_gpu():
    from b200ds import _lib
    lib = _lib.load()
    # rejected before any CUDA call: bad shapes / null pointers
    assert lib.b200ds_tiny_train_steps_f32(None, None, None, None, 64, 64, None, None, None, 4, 0, 4, 1, 1,
                                           1e-3, 0.9, 0.999, 1e-8, 0, None, 0) == 1
    assert b"F must be in [1,32]" in lib.b200ds_last_error()
    assert lib.b200ds_bce_loss_grad(None, None, 0, None, None, None, None, 0, 0) == 1
    assert lib.b200ds_philox_normal(None, 10, 1, 3, None, 0) == 1        # offset not a multiple of 4
    with pytest.raises(_lib.B200dsError):
        _lib.check(lib.b200ds_adam_step(None, None, None, None, 8, 1e-3, 0.9, 0.999, 1e-8, 1, 1.0, 0, 0), "adam")
    assert lib.b200ds_mlp_backward_f32_scratch_bytes(4, 20, 64) >= 4 * 64 * 22
    assert lib.b200ds_bce_scratch_bytes(4) > 0


def test_product_code_never_touches_the_oracle():
    bad = []
    for path in glob.glob(os.path.join(PKG, "**", "*"), recursive=True) + [os.path.join(ROOT, "train.py")]:
        if os.path.isfile(path) and path.endswith((".py", ".cu", ".cuh", ".h")):
            src = open(path, errors="replace").read()
            if re.search(r"^\s*(from|import)\s+oracle\b", src, re.M) or "oracle/" in src:
                bad.append(path)
    assert not bad, bad


def test_missing_library_is_a_hard_error(monkeypatch, tmp_path):
    from b200ds import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.B200dsError, match="no CPU or PyTorch fallback"):
        _lib.load()
