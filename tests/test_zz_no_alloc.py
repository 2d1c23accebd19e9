"""GPU: ``b2tts_paged_attn_decode_static`` - paged GQA decode attention with the split-KV scratch in a file-scope device
array of the library (no allocation inside the call; no counterpart in the reference)
- against the existing entry point ``b2tts_paged_attn_decode`` with a caller-provided workspace.  Same kernels,
same split count, same summation order: the outputs must be IDENTICAL, on two calls in a row that share the array
(different inputs, so stale contents of the first call would show), alone in a fresh process or after the other tests.

Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
PAGE, KVH, H, HD = 256, 8, 16, 128


def _case(b, max_len, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    width = (max_len + PAGE - 1) // PAGE
    npages = b * width + 3
    kc = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    vc = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    bt = torch.randperm(npages, generator=g, device="cuda")[: b * width].to(torch.int32).view(b, width).contiguous()
    lens = torch.randint(1, max_len + 1, (b,), generator=g, device="cuda", dtype=torch.int32)
    lens[0] = max_len
    if b > 2:
        lens[1], lens[2] = 1, 0                    # a one-token row and a CUDA-graph padding row
    q = torch.randn(b, H, HD, device="cuda", generator=g).to(BF16)
    return q, kc, vc, bt, lens, width * PAGE


def _static(native, q, kc, vc, bt, lens, max_ctx, out):
    """Straight through the C ABI (ctypes), nothing allocated on the way."""
    lib = native.lib()
    rc = lib.b2tts_paged_attn_decode_static(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(),
                                            bt.stride(0), lens.data_ptr(), out.data_ptr(), q.shape[0], H, KVH, HD, PAGE,
                                            ctypes.c_float(HD ** -0.5), max_ctx, torch.cuda.current_stream().cuda_stream)
    return rc, lib.b2tts_last_error().decode()


def test_static_scratch_entry_point_equals_workspace_entry_point_twice(native):
    lib = native.lib()
    assert lib.b2tts_paged_attn_decode_static_max_bytes() == 8 << 20
    cases = [_case(8, 2048, 1), _case(64, 700, 2)]              # both split the KV range: the scratch is really used
    for q, kc, vc, bt, lens, max_ctx in cases:
        assert 0 < native.paged_attn_decode_workspace_bytes(q.shape[0], H, HD, max_ctx) <= 8 << 20
    want = [native.paged_attn_decode(q, kc, vc, bt, lens, HD ** -0.5, max_ctx) for q, kc, vc, bt, lens, max_ctx in cases]
    outs = [torch.full_like(w, float("nan")) for w in want]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (q, kc, vc, bt, lens, max_ctx), out in zip(cases, outs):      # two calls in a row on the shared array
        rc, err = _static(native, q, kc, vc, bt, lens, max_ctx, out)
        assert rc == 0, err
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for w, o, (q, kc, vc, bt, lens, _) in zip(want, outs, cases):
        live = lens > 0
        assert torch.equal(o[live], w[live])
        assert not torch.isnan(o.float()).any()
    # and once more in the opposite order: the array now holds the other case's partials
    outs2 = [torch.full_like(w, float("nan")) for w in want]
    for i in (1, 0):
        q, kc, vc, bt, lens, max_ctx = cases[i]
        assert native.paged_attn_decode_static(q, kc, vc, bt, lens, HD ** -0.5, max_ctx, outs2[i]) is outs2[i]
    torch.cuda.synchronize()
    for w, o in zip(want, outs2):
        assert torch.equal(o, w)


def test_static_scratch_entry_point_small_and_oversized_shapes(native):
    # no split (short table): the scratch is not touched, same result
    q, kc, vc, bt, lens, max_ctx = _case(4, 200, 3)
    want = native.paged_attn_decode(q, kc, vc, bt, lens, HD ** -0.5, max_ctx)
    out = torch.empty_like(want)
    rc, err = _static(native, q, kc, vc, bt, lens, max_ctx, out)
    assert rc == 0, err
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    # beyond the array: a clear error, nothing launched
    b, big_ctx = 512, 16384
    assert native.paged_attn_decode_workspace_bytes(b, H, HD, big_ctx) > 8 << 20
    z = torch.zeros(b, H, HD, device="cuda", dtype=BF16)
    bt_big = torch.zeros(b, big_ctx // PAGE, device="cuda", dtype=torch.int32)
    ln = torch.ones(b, device="cuda", dtype=torch.int32)
    rc, err = _static(native, z, kc, vc, bt_big, ln, big_ctx, torch.empty_like(z))
    assert rc == -5 and "built-in array holds" in err            # B2TTS_EWORKSPACE
    assert _static(native, q, kc, vc, bt, lens, max_ctx, out)[0] == 0      # and the entry point still works afterwards
    torch.cuda.synchronize()
    assert torch.equal(out, want)
