"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/b200tts.h declares.  No compute calls (there is no GPU here)."""
import ctypes
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(native):
    lib = native.lib()
    declared = native.exported_symbols()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b200tts.h but not exported"
    assert lib.b2tts_version() >= 2
    assert lib.b2tts_last_error() is not None


def test_argument_validation_without_gpu(native):
    """Shape/pointer validation happens on the host before any launch."""
    lib = native.lib()
    assert lib.b2tts_store_kvcache(None, None, 0, 0, None, None, None, 4, 1024, None) == -1      # EINVAL
    assert b"null" in lib.b2tts_last_error()
    assert lib.b2tts_store_kvcache(None, None, 0, 0, None, None, None, 0, 1024, None) == 0       # empty batch
    assert lib.b2tts_rmsnorm(ctypes.c_void_p(16), 12, ctypes.c_void_p(16), ctypes.c_void_p(16), 1, 12, 1e-6, 0, None) == -2
    assert lib.b2tts_paged_attn_decode_workspace_bytes(0, 16, 128, 1024) == 0
    big = lib.b2tts_paged_attn_decode_workspace_bytes(1, 16, 128, 4608)
    assert big > 0 and big % (16 * 130 * 4) == 0


def test_no_cpu_fallback(native):
    import torch
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        native.rmsnorm(torch.zeros(2, 128, dtype=torch.bfloat16), torch.ones(128, dtype=torch.bfloat16), 1e-6)


def test_sass_uses_bulk_async_copy():
    """The attention kernel must stream K/V through the TMA unit (UBLKCP in SASS)."""
    so = os.path.join(ROOT, "nano-qwen3tts-vllm_b200", "lib", "libb200tts.so")
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "UBLKCP" in out.stdout
    assert "sm_100a" in subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout


def test_sass_has_tcgen05_tmem_and_tma_per_kernel():
    """SASS evidence per kernel (profiles/sass_counts.txt is the committed copy): the projection GEMM issues
    tcgen05.mma (UTCHMMA) with TMEM accumulators read by tcgen05.ld (LDTM) and operands moved by the TMA unit
    (UTMALDG tensor maps + UBLKCP bulk copies); the attention kernels stream K/V with UBLKCP; the tile prefill
    attention runs QK^T and PV on tcgen05."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import sass_counts
    try:
        c = sass_counts.counts()
    except (RuntimeError, OSError):
        pytest.skip("cuobjdump unavailable")
    gemm = [v for k, v in c.items() if k.startswith("b2::gemm_stream_kernel")]
    assert gemm and all(v["UTCHMMA"] > 0 and v["LDTM"] > 0 and v["UTMALDG"] > 0 and v["UBLKCP"] > 0 for v in gemm)
    for prefix in ("b2::paged_attn_kernel", "b2::fused_decode_attn_kernel"):
        ks = [v for k, v in c.items() if k.startswith(prefix)]
        assert ks and all(v["UBLKCP"] > 0 for v in ks), prefix
    assert not any(v["HMMA"] for v in c.values()), "no legacy mma.sync tensor-core path in this library"


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "nano-qwen3tts-vllm_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, os.path.join(dp, f)
