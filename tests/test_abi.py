"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/b200llm.h
declares, and fails loudly (no fallback) without a GPU.  No compute calls here."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200llm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200llm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from b200llm import _lib
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(_lib.lib, n), f"libb200llm.so does not export {n}"
        assert n in _lib.SYMBOLS, f"ctypes binding misses {n}"
    assert sorted(_lib.SYMBOLS) == names
    assert _lib.lib.b200llm_abi_version() == _lib.ABI_VERSION


def test_struct_sizes_match_header():
    from b200llm import _lib
    assert C.sizeof(_lib.Sampling) == 24
    assert C.sizeof(_lib.TensorDesc) == 32
    assert C.sizeof(_lib.Timing) == 24
    assert C.sizeof(_lib.Config) == 104


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly():
    from b200llm import _lib
    geo_cfg = _lib.Config(_lib.ABI_VERSION, 256, 2, 4, 2, 64, 512, 1024, 1e-5, 16, 2, 8, 4, 0, 1, 0, 1, 1, 0, 1, 1, None, None)
    h = C.c_void_p()
    rc = _lib.lib.b200llm_create(C.byref(geo_cfg), C.byref(h))
    assert rc == _lib.ENODEV
    assert b"no CPU fallback" in _lib.lib.b200llm_last_error()
    rc = _lib.lib.b200llm_op_fill_weights(None, 0, 0, 0.02, 0.0, None)
    assert rc == _lib.ENODEV
    import ollama
    with pytest.raises(ollama.ResponseError) as ei:
        ollama.generate(model="tiny", system="s", prompt="p")
    assert ei.value.status_code == 500


def test_config_validation_messages():
    from b200llm import _lib
    n = C.c_size_t()
    bad = _lib.Config(_lib.ABI_VERSION, 256, 2, 4, 2, 96, 512, 1024, 1e-5, 16, 2, 8, 4, 0, 1, 0, 1, 1, 0, 1, 1, None, None)
    assert _lib.lib.b200llm_workspace_bytes(C.byref(bad), C.byref(n)) == _lib.EINVAL
    assert b"head_dim" in _lib.lib.b200llm_last_error()
    ok = _lib.Config(_lib.ABI_VERSION, 256, 2, 4, 2, 64, 512, 1024, 1e-5, 16, 2, 8, 4, 0, 1, 0, 1, 1, 0, 1, 1, None, None)
    assert _lib.lib.b200llm_workspace_bytes(C.byref(ok), C.byref(n)) == 0 and n.value > 0
    assert _lib.lib.b200llm_kv_page_bytes(C.byref(ok), C.byref(n)) == 0
    assert n.value == 2 * 2 * 2 * 16 * 64 * 2
    wrong_abi = _lib.Config(99, 256, 2, 4, 2, 64, 512, 1024, 1e-5, 16, 2, 8, 4, 0, 1, 0, 1, 1, 0, 1, 1, None, None)
    assert _lib.lib.b200llm_workspace_bytes(C.byref(wrong_abi), C.byref(n)) == _lib.EINVAL


def test_gemv_grid_planning():
    from b200llm import _lib
    g = C.c_int32()
    for (N, K, bs, epi) in [(5120, 3072, 1, _lib.EPI_QKV_ROPE), (3072, 3072, 1, _lib.EPI_RESIDUAL),
                            (16384, 3072, 1, _lib.EPI_SWIGLU), (3072, 8192, 1, _lib.EPI_RESIDUAL),
                            (128256, 3072, 1, _lib.EPI_LOGITS), (4096, 11008, 4, _lib.EPI_RESIDUAL),
                            (22016, 4096, 2, _lib.EPI_SWIGLU)]:
        assert _lib.lib.b200llm_op_gemv_grid(N, K, bs, epi, C.byref(g)) == 0
        assert g.value % 148 == 0 or g.value <= N
    assert _lib.lib.b200llm_op_gemv_grid(64, 30, 1, 0, C.byref(g)) == _lib.EINVAL      # K % 8 != 0
    assert _lib.lib.b200llm_op_gemv_grid(64, 64, 5, 0, C.byref(g)) == _lib.EINVAL      # bs > 4
