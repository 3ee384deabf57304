"""CPU-side checks of the drop-in boundary: the shared library builds, loads and exports every
symbol include/pocket_tts_b200.h declares (no compute without a GPU)."""
import ctypes

import pytest


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge

    ge.build()
    from pocket_tts_b200 import _native

    return _native.lib()


def test_every_header_symbol_is_exported(lib):
    from pocket_tts_b200 import _native

    names = _native.exported_symbols_from_header()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), n


def test_version_and_error_string(lib):
    assert lib.ptts_version() == 100
    assert isinstance(lib.ptts_last_error(), bytes)


def test_null_handles_fail_cleanly(lib):
    from pocket_tts_b200 import _native as N

    assert lib.ptts_finalize(None) == N.PTTS_EINVAL
    assert b"null engine" in lib.ptts_last_error()
    assert lib.ptts_engine_destroy(None) == N.PTTS_OK
    assert lib.ptts_launch_count(None) == -1


def test_engine_create_without_gpu_reports_cuda_error(lib):
    import torch

    from pocket_tts_b200 import _native as N

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    cfg = N.PttsConfig()
    cfg.n_layers, cfg.d_model, cfg.n_heads, cfg.mimi_d, cfg.mimi_heads = 1, 1024, 16, 512, 8
    cfg.ldim, cfg.flow_dim, cfg.max_streams, cfg.kv_capacity = 32, 512, 1, 16
    h = ctypes.c_void_p()
    rc = lib.ptts_engine_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == N.PTTS_ECUDA and not h.value      # fails loudly: no CPU fallback
    assert b"CUDA error" in lib.ptts_last_error()
