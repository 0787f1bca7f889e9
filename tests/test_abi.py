"""The C-ABI library loads on a machine without a GPU and exports every symbol include/owl_dit.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "owl_dit.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(owl_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("owl_gemm_bf16", "owl_attn_fwd", "owl_attn_bwd", "owl_adaln_modulate_fwd", "owl_qk_norm_rope_fwd",
                 "owl_kv_append", "owl_abi_version"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from owl_audio_exps_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_matches_header():
    from owl_audio_exps_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()
    lib = _lib.load()
    assert lib.owl_abi_version() >= 1
    assert lib.owl_error_string(-2).decode()  # host-only calls: no GPU needed
    assert lib.owl_launch_count() == 0


def test_bad_arguments_are_rejected_without_a_gpu():
    """Argument validation happens before any CUDA call: error codes, never a crash."""
    from owl_audio_exps_b200 import _lib
    lib = _lib.load()
    assert lib.owl_silu_bf16(None, None, 16, None) == -1
    assert lib.owl_kv_append(None, 0, 0, None, 0, 0, 0, 1, 1, 1, 64, None) == -1
    e = _lib.GemmEpilogue()
    assert lib.owl_gemm_bf16(0, 0, 8, 8, None, 8, None, 8, ctypes.byref(e), None) == -1
    m = _lib.AttnMask()
    assert lib.owl_attn_fwd(None, 0, 0, None, 0, 0, None, 0, 0, None, 0, 0, None, 1, 1, 1, 1, 64, ctypes.byref(m), None) == -1


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from owl_audio_exps_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "owl_audio_exps_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_cpu_tensors_are_refused():
    import torch
    import owl_audio_exps_b200.nn as N
    from helpers import cfg_of, load_golden
    fx = load_golden("audio_T1")
    model = N.DiT(cfg_of(fx))
    with pytest.raises(RuntimeError, match="CUDA"):
        model(fx["x"].float(), fx["cond"].float())
