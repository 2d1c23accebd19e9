"""ctypes binding of ``lib/libb200tts.so`` (C ABI: ``include/b200tts.h``).

This is the ONLY route from Python to the hot-path arithmetic.  There is no CPU fallback and no
alternative backend: if the library is missing or a call fails, a ``RuntimeError`` is raised
(it surfaces through the workers' existing ``except`` -> empty-result path,
workers/talker_worker.py:139-144 in the reference).

Every wrapper takes torch CUDA tensors, checks dtype/contiguity, and passes raw device pointers
plus ``torch.cuda.current_stream().cuda_stream`` so the launches are capturable by
``torch.cuda.graph``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

from . import _build

ROUND_EAGER = 0
ROUND_FUSED = 1

_lib = None


class NativeLibraryError(RuntimeError):
    pass


def _load():
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path):
        raise NativeLibraryError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no fallback path.")
    lib = ctypes.CDLL(path)
    P, I, L, F, Z = c_void_p, c_int, c_int64, c_float, c_size_t
    sig = {
        "b2tts_version": (I, []),
        "b2tts_last_error": (c_char_p, []),
        "b2tts_device_info": (I, [P, P, P, P]),
        "b2tts_launch_count": (ctypes.c_ulonglong, []),
        "b2tts_store_kvcache": (I, [P, P, L, L, P, P, P, I, I, P]),
        "b2tts_paged_attn_decode_workspace_bytes": (Z, [I, I, I, I]),
        "b2tts_paged_attn_decode": (I, [P, L, P, P, P, I, P, P, P, Z, I, I, I, I, I, F, I, P]),
        "b2tts_paged_attn_decode_static_max_bytes": (Z, []),
        "b2tts_paged_attn_decode_static": (I, [P, L, P, P, P, I, P, P, I, I, I, I, I, F, I, P]),
        "b2tts_varlen_attn_prefill": (I, [P, L, P, L, P, L, P, P, P, I, I, I, I, I, I, I, I, F, P, I, I, P]),
        "b2tts_rmsnorm": (I, [P, L, P, P, I, I, F, I, P]),
        "b2tts_add_rmsnorm": (I, [P, P, P, P, P, I, I, F, I, I, P]),
        "b2tts_qknorm_rope_kvwrite": (I, [P, L, P, L, P, L, P, P, P, P, F, P, L, P, P, F, P, P, P, I, I, I, I, I, P]),
        "b2tts_silu_and_mul": (I, [P, P, L, P, I, I, I, P]),
        "b2tts_topk_sample": (I, [P, L, I, P, P, P, P, I, I, I, F, I, P]),
        "b2tts_topk_sample_embed": (I, [P, L, I, P, P, P, I, I, I, P, I, P, L, P]),
        "b2tts_gather_rows": (I, [P, L, I, P, P, P, L, I, I, P]),
        "b2tts_codec_embed_sum": (I, [P, P, L, P, L, P, L, P, P, I, I, I, P]),
        "b2tts_decode_layout": (I, [P, P, I, I, I, I, I, P, P, P, I, P]),
        "b2tts_frame_advance": (I, [P, I, P, P, P, P, P, I, I, I, P]),
        "b2tts_gemm_bf16_supported": (I, [I, I, I]),
        "b2tts_gemm_packed_bytes": (Z, [I, I]),
        "b2tts_gemm_pack_weight": (I, [P, L, P, I, I, I, P]),
        "b2tts_gemm_bf16": (I, [P, L, P, P, P, L, P, L, I, I, I, I, I, I, I, P]),
        "b2tts_gemm_ssq_slots": (I, [I, I, I, I]),
        "b2tts_fused_decode_attn_workspace_bytes": (Z, [I, I, I, I]),
        "b2tts_fused_decode_attn": (I, [P, L, P, L, P, L, P, P, F, P, L, P, P, F, P, P, P, P, I, P, P, P, Z,
                                        I, I, I, I, I, F, I, I, I, P]),
        "b2tts_gemm_norm_bf16": (I, [P, L, P, P, P, L, P, L, I, I, I, I, I, I, I, P, P, I, F, P, P]),
        "b2tts_gemm_prefetch_bf16": (I, [P, L, P, P, P, L, P, L, I, I, I, I, I, I, I, P, P, I, F, P, P, Z, P]),
        "b2tts_gemm_chain_ssq_slots": (I, [I]),
        "b2tts_gemm_chain_workspace_bytes": (Z, [I, P, P, I]),
        "b2tts_gemm_chain": (I, [P, I, I, F, I, P, Z, I, P]),
        "b2tts_prompt_assemble": (I, [P, L, I, P, L, I, P, P, P, L, I, I, P]),
        "b2tts_float_to_pcm16": (I, [P, P, L, F, P]),
        "b2tts_decoder_stack_plan": (I, [I, I, I, I, P, P]),
        "b2tts_decoder_stack_workspace_bytes": (Z, [I, I, I, I]),
        "b2tts_decoder_stack_decode": (I, [P, L, P, L, P, I, P, I, I, I, I, I, I, F, I, P, P, F, P, P, I, P, I, F, P, Z, P]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(lib, name):
            raise NativeLibraryError(f"{path} does not export {name}; rebuild the library")
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def lib():
    return _load()


def exported_symbols() -> list[str]:
    """Names declared in include/b200tts.h (used by the CPU-side ABI test)."""
    import re
    with open(os.path.join(_build.INCLUDE, "b200tts.h")) as f:
        return sorted(set(re.findall(r"\b(b2tts_[a-z0-9_]+)\s*\(", f.read())))


def _check(rc: int, what: str):
    if rc != 0:
        msg = _load().b2tts_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libb200tts {what} failed (code {rc}): {msg}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _cuda(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"libb200tts: `{name}` must be a CUDA tensor (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"libb200tts: `{name}` must be {dtype}, got {t.dtype}")
    return t


BF16, I32, I64, F32 = torch.bfloat16, torch.int32, torch.int64, torch.float32


# ------------------------------------------------------------------------------------------ a1
def store_kvcache(key, value, k_cache, v_cache, slot_mapping):
    """key/value [N, kv_heads, hd] (inner dims contiguous, any row stride); caches
    [blocks, page, kv_heads, hd]; slot_mapping [N] int32."""
    _cuda(key, BF16, "key"); _cuda(value, BF16, "value"); _cuda(k_cache, BF16, "k_cache")
    _cuda(v_cache, BF16, "v_cache"); _cuda(slot_mapping, I32, "slot_mapping")
    n = key.shape[0]
    d = key.shape[-1] * key.shape[-2] if key.dim() == 3 else key.shape[-1]
    _check(_load().b2tts_store_kvcache(key.data_ptr(), value.data_ptr(), key.stride(0), value.stride(0),
                                       k_cache.data_ptr(), v_cache.data_ptr(), slot_mapping.data_ptr(),
                                       n, d, _stream()), "store_kvcache")


# ------------------------------------------------------------------------------------------ a2
def paged_attn_decode_workspace_bytes(batch, num_heads, head_dim, max_ctx) -> int:
    return int(_load().b2tts_paged_attn_decode_workspace_bytes(batch, num_heads, head_dim, max_ctx))


def paged_attn_decode(q, k_cache, v_cache, block_table, context_lens, scale, max_ctx, out=None, workspace=None):
    """q [B, heads, hd] (row stride free); returns out [B, heads, hd] bf16."""
    _cuda(q, BF16, "q"); _cuda(k_cache, BF16, "k_cache"); _cuda(v_cache, BF16, "v_cache")
    _cuda(block_table, I32, "block_table"); _cuda(context_lens, I32, "context_lens")
    b, h, hd = q.shape
    page, kvh = k_cache.shape[1], k_cache.shape[2]
    if out is None:
        out = torch.empty(b, h, hd, dtype=BF16, device=q.device)
    need = paged_attn_decode_workspace_bytes(b, h, hd, max_ctx)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        workspace = torch.empty(need, dtype=torch.uint8, device=q.device)
    wbytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    _check(_load().b2tts_paged_attn_decode(q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                                           block_table.data_ptr(), block_table.stride(0), context_lens.data_ptr(),
                                           out.data_ptr(), _ptr(workspace), wbytes, b, h, kvh, hd, page,
                                           float(scale), int(max_ctx), _stream()), "paged_attn_decode")
    return out


def paged_attn_decode_static(q, k_cache, v_cache, block_table, context_lens, scale, max_ctx, out):
    """``paged_attn_decode`` with the split-KV scratch in the library's own device array: the CALL allocates nothing,
    so ``out`` is required.  Shapes that need more than ``b2tts_paged_attn_decode_static_max_bytes()`` raise."""
    _cuda(q, BF16, "q"); _cuda(k_cache, BF16, "k_cache"); _cuda(v_cache, BF16, "v_cache")
    _cuda(block_table, I32, "block_table"); _cuda(context_lens, I32, "context_lens"); _cuda(out, BF16, "out")
    b, h, hd = q.shape
    page, kvh = k_cache.shape[1], k_cache.shape[2]
    _check(_load().b2tts_paged_attn_decode_static(q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                                                  block_table.data_ptr(), block_table.stride(0), context_lens.data_ptr(),
                                                  out.data_ptr(), b, h, kvh, hd, page, float(scale), int(max_ctx),
                                                  _stream()), "paged_attn_decode_static")
    return out


_PDL_ATTN = os.environ.get("B2TTS_PDL_ATTN", "1") != "0"       # experiments: launch the fused attention without PDL


def fused_decode_attn_workspace(batch, num_heads, head_dim, max_ctx, device) -> torch.Tensor:
    """Zero-filled scratch for ``fused_decode_attn`` (split partials + per-row tickets; the kernel leaves the
    tickets zero, so one allocation serves every later launch of up to ``batch`` rows)."""
    n = max(int(_load().b2tts_fused_decode_attn_workspace_bytes(b, num_heads, head_dim, max_ctx))
            for b in range(1, batch + 1))
    return torch.zeros(n + 16, dtype=torch.uint8, device=device)


def fused_decode_attn(q, k, v, q_weight, k_weight, eps, positions, inv_freq, slot_mapping, k_cache, v_cache,
                      block_table, context_lens, scale, max_ctx, workspace, *, modality=None, attention_scaling=1.0,
                      rounding=ROUND_EAGER, out=None, pdl=True):
    """The whole decode step of the attention block in one launch: q/k head norm + RoPE, KV write of the new
    token, paged GQA attention over ``context_lens`` tokens (the new one included), split-KV combine.
    q [B, heads, 128], k/v [B, kv_heads, 128] are views of the QKV projection; returns [B, heads, 128]."""
    _cuda(q, BF16, "q"); _cuda(k, BF16, "k"); _cuda(v, BF16, "v")
    _cuda(positions, I64, "positions"); _cuda(inv_freq, F32, "inv_freq"); _cuda(slot_mapping, I32, "slot_mapping")
    _cuda(block_table, I32, "block_table"); _cuda(context_lens, I32, "context_lens")
    b, h, hd = q.shape
    page, kvh = k_cache.shape[1], k_cache.shape[2]
    if out is None:
        out = torch.empty(b, h, hd, dtype=BF16, device=q.device)
    pos_stride = positions.stride(0) if positions.dim() == 2 else 0
    _check(_load().b2tts_fused_decode_attn(
        q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
        _ptr(q_weight), _ptr(k_weight), float(eps), positions.data_ptr(), pos_stride, inv_freq.data_ptr(),
        _ptr(modality), float(attention_scaling), slot_mapping.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
        block_table.data_ptr(), block_table.stride(0), context_lens.data_ptr(), out.data_ptr(),
        workspace.data_ptr(), workspace.numel() * workspace.element_size(), b, h, kvh, hd, page, float(scale),
        int(max_ctx), int(rounding), int(bool(pdl) and _PDL_ATTN), _stream()), "fused_decode_attn")
    return out


# ------------------------------------------------------------------------------------------ a3
def varlen_attn_prefill(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale,
                        block_table=None, out=None):
    """q [T, heads, hd]; k, v packed [Tk, kv_heads, hd] or (with block_table) the paged caches."""
    _cuda(q, BF16, "q"); _cuda(k, BF16, "k"); _cuda(v, BF16, "v")
    _cuda(cu_seqlens_q, I32, "cu_seqlens_q"); _cuda(cu_seqlens_k, I32, "cu_seqlens_k")
    t, h, hd = q.shape
    if out is None:
        out = torch.empty(t, h, hd, dtype=BF16, device=q.device)
    nseq = cu_seqlens_q.numel() - 1
    if block_table is not None:
        _cuda(block_table, I32, "block_table")
        page, kvh = k.shape[1], k.shape[2]
        ks = vs = kvh * hd
        bt_ptr, bt_stride = block_table.data_ptr(), block_table.stride(0)
        total_k = 0
    else:
        kvh, page = k.shape[1], 0
        ks, vs = k.stride(0), v.stride(0)
        bt_ptr, bt_stride = None, 0
        total_k = min(k.shape[0], v.shape[0])
    _check(_load().b2tts_varlen_attn_prefill(q.data_ptr(), q.stride(0), k.data_ptr(), ks, v.data_ptr(), vs,
                                             cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(), out.data_ptr(),
                                             nseq, t, total_k, int(max_seqlen_q), int(max_seqlen_k), h, kvh, hd, float(scale),
                                             bt_ptr, bt_stride, page, _stream()), "varlen_attn_prefill")
    return out


# ------------------------------------------------------------------------------------------ a4
def rmsnorm(x, weight, eps, rounding=ROUND_EAGER, out=None):
    _cuda(x, BF16, "x"); _cuda(weight, BF16, "weight")
    cols = x.shape[-1]
    x2 = x.reshape(-1, cols) if x.is_contiguous() else x
    if x2.dim() != 2 or x2.stride(1) != 1:
        x2 = x.contiguous().view(-1, cols)
    if out is None:
        out = torch.empty(x.shape, dtype=BF16, device=x.device)
    _check(_load().b2tts_rmsnorm(x2.data_ptr(), x2.stride(0), weight.data_ptr(), out.data_ptr(),
                                 x2.shape[0], cols, float(eps), rounding, _stream()), "rmsnorm")
    return out


def add_rmsnorm(x, residual, weight, eps, rounding=ROUND_EAGER, round_sum_first=False, out=None, res_out=None):
    _cuda(x, BF16, "x"); _cuda(residual, BF16, "residual"); _cuda(weight, BF16, "weight")
    cols = x.shape[-1]
    x2, r2 = x.contiguous().view(-1, cols), residual.contiguous().view(-1, cols)
    if out is None:
        out = torch.empty(x.shape, dtype=BF16, device=x.device)
    if res_out is None:
        res_out = torch.empty(x.shape, dtype=BF16, device=x.device)
    _check(_load().b2tts_add_rmsnorm(x2.data_ptr(), r2.data_ptr(), weight.data_ptr(), out.data_ptr(),
                                     res_out.data_ptr(), x2.shape[0], cols, float(eps), rounding,
                                     int(bool(round_sum_first)), _stream()), "add_rmsnorm")
    return out, res_out


# ------------------------------------------------------------------------------ a4+a5(+a1) fused
def qknorm_rope_kvwrite(q, k, v, q_weight, k_weight, eps, positions, inv_freq, *, modality=None,
                        attention_scaling=1.0, slot_mapping=None, k_cache=None, v_cache=None,
                        q_out=None, k_out=None, want_k_out=True, rounding=ROUND_EAGER):
    """q [N, heads, 128], k/v [N, kv_heads, 128] (row strides free).  Returns (q_out, k_out)."""
    _cuda(q, BF16, "q"); _cuda(k, BF16, "k")
    _cuda(positions, I64, "positions"); _cuda(inv_freq, F32, "inv_freq")
    n, h, hd = q.shape
    kvh = k.shape[1]
    if q_out is None:
        q_out = torch.empty(n, h, hd, dtype=BF16, device=q.device)
    if k_out is None and want_k_out:
        k_out = torch.empty(n, kvh, hd, dtype=BF16, device=q.device)
    pos_stride = positions.stride(0) if positions.dim() == 2 else 0
    if v is not None:
        _cuda(v, BF16, "v")
    if slot_mapping is not None:
        _cuda(slot_mapping, I32, "slot_mapping")
    if modality is not None:
        _cuda(modality, I32, "modality")
    _check(_load().b2tts_qknorm_rope_kvwrite(
        q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), _ptr(v), 0 if v is None else v.stride(0),
        q_out.data_ptr(), _ptr(k_out), _ptr(q_weight), _ptr(k_weight), float(eps),
        positions.data_ptr(), pos_stride, inv_freq.data_ptr(), _ptr(modality), float(attention_scaling),
        _ptr(slot_mapping), _ptr(k_cache), _ptr(v_cache), n, h, kvh, hd, rounding, _stream()), "qknorm_rope_kvwrite")
    return q_out, k_out


# ------------------------------------------------------------------------------------------ a6
def silu_and_mul(gate, up, rounding=ROUND_EAGER, out=None):
    """gate/up [rows, inter] with a common row stride (halves of a fused gate_up output).
    ``up=None`` computes silu(gate)."""
    _cuda(gate, BF16, "gate")
    inter = gate.shape[-1]
    g2 = gate if gate.dim() == 2 else gate.reshape(-1, inter)
    if g2.stride(1) != 1:
        g2 = g2.contiguous()
    u2 = None
    if up is not None:
        _cuda(up, BF16, "up")
        u2 = up if up.dim() == 2 else up.reshape(-1, inter)
        if u2.stride(1) != 1 or u2.stride(0) != g2.stride(0):
            g2, u2 = g2.contiguous(), u2.contiguous()
    if out is None:
        out = torch.empty(gate.shape, dtype=BF16, device=gate.device)
    _check(_load().b2tts_silu_and_mul(g2.data_ptr(), _ptr(u2), g2.stride(0), out.data_ptr(), g2.shape[0], inter,
                                      rounding, _stream()), "silu_and_mul")
    return out


# ------------------------------------------------------------------------------------------ a8
def topk_sample(logits, temperatures, noise, top_k=50, out=None, probs_out=None, embed_table=None, embed_out=None):
    """logits [B, V] bf16 or fp32; temperatures [B] fp32; noise [B, V] fp32 Exp(1) draws.
    ``embed_table`` [V, cols] bf16 + ``embed_out`` [B, cols] (row stride free): also gathers
    ``embed_table[index]`` into ``embed_out`` in the same launch (``b2tts_topk_sample_embed``)."""
    _cuda(logits, None, "logits"); _cuda(temperatures, F32, "temperatures"); _cuda(noise, F32, "noise")
    if logits.dtype not in (BF16, F32):
        raise RuntimeError(f"libb200tts: logits must be bf16 or fp32, got {logits.dtype}")
    b, v = logits.shape
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    if not noise.is_contiguous():
        noise = noise.contiguous()
    if out is None:
        out = torch.empty(b, dtype=I64, device=logits.device)
    if embed_table is not None:
        _cuda(embed_table, BF16, "embed_table"); _cuda(embed_out, BF16, "embed_out")
        if embed_table.dim() != 2 or embed_table.stride(1) != 1 or embed_table.stride(0) != embed_table.shape[1] \
                or embed_out.stride(-1) != 1:
            raise RuntimeError("libb200tts: embed_table must be a contiguous [V, cols] tensor")
        _check(_load().b2tts_topk_sample_embed(logits.data_ptr(), logits.stride(0), int(logits.dtype == F32),
                                               temperatures.data_ptr(), noise.data_ptr(), out.data_ptr(), b, v,
                                               int(min(top_k, v)), embed_table.data_ptr(), embed_table.shape[1],
                                               embed_out.data_ptr(), embed_out.stride(0), _stream()),
               "topk_sample_embed")
        return out
    _check(_load().b2tts_topk_sample(logits.data_ptr(), logits.stride(0), int(logits.dtype == F32),
                                     temperatures.data_ptr(), noise.data_ptr(), out.data_ptr(), _ptr(probs_out),
                                     b, v, int(min(top_k, v)), 1.0, 0, _stream()), "topk_sample")
    return out


# ------------------------------------------------------------------------------------------ a9
def gather_rows(tables, row_index, table_index=None, out=None):
    """tables [num_tables, rows, cols] (or [rows, cols]); out[i] = tables[table_index[i]][row_index[i]]."""
    _cuda(tables, BF16, "tables"); _cuda(row_index, I64, "row_index")
    if tables.dim() == 2:
        tables = tables.unsqueeze(0)
    nt, rows, cols = tables.shape
    n = row_index.numel()
    if out is None:
        out = torch.empty(n, cols, dtype=BF16, device=tables.device)
    if table_index is not None:
        _cuda(table_index, I32, "table_index")
    _check(_load().b2tts_gather_rows(tables.data_ptr(), tables.stride(0), rows, row_index.data_ptr(),
                                     _ptr(table_index), out.data_ptr(), out.stride(0), n, cols, _stream()),
           "gather_rows")
    return out


def codec_embed_sum(table0, tables_rest, codes, extra=None, out=None, extra_index=None):
    """table0 [V0, cols]; tables_rest [ncb-1, V, cols]; codes [n, ncb] int64; extra [n, cols] or [1, cols], or any
    [rows, cols] with ``extra_index`` [n] int64 picking the row of each stream."""
    _cuda(table0, BF16, "table0"); _cuda(codes, I64, "codes")
    n, ncb = codes.shape
    cols = table0.shape[-1]
    if out is None:
        out = torch.empty(n, cols, dtype=BF16, device=table0.device)
    es = 0
    if extra is not None:
        _cuda(extra, BF16, "extra")
        if extra_index is not None:
            _cuda(extra_index, I64, "extra_index")
            es = extra.stride(0)
        else:
            es = extra.stride(0) if extra.shape[0] == n else 0      # one row broadcasts
    _check(_load().b2tts_codec_embed_sum(table0.data_ptr(), _ptr(tables_rest),
                                         0 if tables_rest is None else tables_rest.stride(0),
                                         codes.data_ptr(), codes.stride(0), _ptr(extra), es,
                                         _ptr(extra_index) if extra is not None else None, out.data_ptr(),
                                         n, ncb, cols, _stream()), "codec_embed_sum")
    return out


def decode_layout(seq_lens, block_table, page_size, position_offset, advance, positions, slot_mapping, context_lens):
    """Device-side prepare_decode: see include/b200tts.h.  All tensors are device tensors, updated in place."""
    _cuda(seq_lens, I32, "seq_lens"); _cuda(block_table, I32, "block_table"); _cuda(positions, I64, "positions")
    _cuda(slot_mapping, I32, "slot_mapping"); _cuda(context_lens, I32, "context_lens")
    n = seq_lens.numel()
    _check(_load().b2tts_decode_layout(seq_lens.data_ptr(), block_table.data_ptr(), block_table.stride(0),
                                       block_table.shape[1], int(page_size), int(position_offset), int(advance),
                                       positions.data_ptr(), slot_mapping.data_ptr(), context_lens.data_ptr(), n,
                                       _stream()), "decode_layout")


def frame_advance(code0, eos_id, active, seq_lens, gen_step, frames_done, extra_row, trail_rows, max_frames):
    """Per-frame stream bookkeeping on the device (``b2tts_frame_advance``): EOS / frame-budget retirement,
    trailing-text row selection, frame counters.  ``eos_id`` < 0 ignores EOS."""
    _cuda(code0, I64, "code0"); _cuda(active, I32, "active"); _cuda(seq_lens, I32, "seq_lens")
    _cuda(gen_step, I32, "gen_step"); _cuda(frames_done, I32, "frames_done"); _cuda(extra_row, I64, "extra_row")
    _check(_load().b2tts_frame_advance(code0.data_ptr(), int(eos_id), active.data_ptr(), seq_lens.data_ptr(),
                                       gen_step.data_ptr(), frames_done.data_ptr(), extra_row.data_ptr(),
                                       int(trail_rows), int(max_frames), code0.numel(), _stream()), "frame_advance")


# ------------------------------------------------------------------------------------------ a7
GEMM_BIAS, GEMM_RESIDUAL, GEMM_SILU_MUL = 0, 1, 2
PACK_PLAIN, PACK_GATE_UP = 0, 1
GEMM_MAX_M = 1 << 16        # any batch: rows beyond 128 run as 128-row chunks of the same kernel (gemm.cu)
GEMM_CHAIN_MAX_M = 128


def launch_count() -> int:
    """Kernels this library has launched (or recorded under graph capture) in this process."""
    return int(_load().b2tts_launch_count())


def device_info():
    sm, smem, maj, mnr = c_int(), c_int(), c_int(), c_int()
    _check(_load().b2tts_device_info(ctypes.byref(sm), ctypes.byref(smem), ctypes.byref(maj), ctypes.byref(mnr)),
           "device_info")
    return {"sm_count": sm.value, "max_smem": smem.value, "cc": (maj.value, mnr.value)}


def gemm_supported(m: int, n: int, k: int) -> bool:
    return bool(_load().b2tts_gemm_bf16_supported(int(m), int(n), int(k)))


def gemm_pack_weight(weight, mode=PACK_PLAIN):
    """Re-tile a row-major bf16 weight [n, k] into the UMMA operand image the GEMM streams
    (``b2tts_gemm_pack_weight``).  Returns a uint8 tensor of n*k*2 bytes."""
    _cuda(weight, BF16, "weight")
    n, k = weight.shape
    if weight.stride(1) != 1:
        weight = weight.contiguous()
    nbytes = int(_load().b2tts_gemm_packed_bytes(n, k))
    if nbytes == 0:
        raise RuntimeError(f"libb200tts: gemm_pack_weight: n={n} must be a multiple of 128 and k={k} of 64")
    packed = torch.empty(nbytes, dtype=torch.uint8, device=weight.device)
    _check(_load().b2tts_gemm_pack_weight(weight.data_ptr(), weight.stride(0), packed.data_ptr(), n, k, int(mode),
                                          _stream()), "gemm_pack_weight")
    return packed


def gemm_ssq_slots(m: int, n: int, k: int, splits: int = 0) -> int:
    return int(_load().b2tts_gemm_ssq_slots(int(m), int(n), int(k), int(splits)))


def gemm_bf16(x, packed, n, bias=None, residual=None, epilogue=GEMM_BIAS, rounding=ROUND_EAGER, out=None,
              splits=0, pdl=False, norm_weight=None, ssq_in=None, eps=1e-6, ssq_out=None, prefetch=None):
    """y = epilogue(x[M,K] @ W[N,K]^T) with W pre-packed by ``gemm_pack_weight``; bf16 in/out, fp32
    accumulation in TMEM.  GEMM_SILU_MUL returns [M, N/2]; GEMM_RESIDUAL may write in place (out=residual).
    ``norm_weight`` + ``ssq_in`` [M, slots] fp32: x is normalised (RMSNorm) while it is staged.
    ``ssq_out`` [M, gemm_ssq_slots(n, k, splits)] fp32 receives the row sums of squares of the output.
    ``prefetch``: the packed image of the NEXT projection's weight - pulled into L2 while this GEMM runs."""
    _cuda(x, BF16, "x")
    m, k = x.shape
    if x.stride(1) != 1:
        x = x.contiguous()
    cols = n // 2 if epilogue == GEMM_SILU_MUL else n
    if out is None:
        out = torch.empty(m, cols, dtype=BF16, device=x.device)
    if residual is not None:
        _cuda(residual, BF16, "residual")
    res_stride = residual.stride(0) if residual is not None else 0
    if norm_weight is None and ssq_out is None and prefetch is None:
        _check(_load().b2tts_gemm_bf16(x.data_ptr(), x.stride(0), packed.data_ptr(), _ptr(bias), _ptr(residual),
                                       res_stride, out.data_ptr(), out.stride(0),
                                       m, n, k, int(epilogue), int(rounding), int(splits), int(bool(pdl)), _stream()),
               "gemm_bf16")
        return out
    slots_in = 0
    if norm_weight is not None:
        _cuda(norm_weight, BF16, "norm_weight"); _cuda(ssq_in, torch.float32, "ssq_in")
        if ssq_in.dim() != 2 or ssq_in.shape[0] < m or not ssq_in.is_contiguous():
            raise ValueError("ssq_in must be a contiguous [>=M, slots] fp32 tensor")
        slots_in = ssq_in.shape[1]
    if ssq_out is not None:
        _cuda(ssq_out, torch.float32, "ssq_out")
        need = gemm_ssq_slots(m, n, k, splits)
        if ssq_out.dim() != 2 or ssq_out.shape[0] < m or ssq_out.shape[1] != need or not ssq_out.is_contiguous():
            raise ValueError(f"ssq_out must be a contiguous [>=M, {need}] fp32 tensor")
    if prefetch is not None:
        _check(_load().b2tts_gemm_prefetch_bf16(
            x.data_ptr(), x.stride(0), packed.data_ptr(), _ptr(bias), _ptr(residual), res_stride, out.data_ptr(),
            out.stride(0), m, n, k, int(epilogue), int(rounding), int(splits), int(bool(pdl)), _ptr(norm_weight),
            _ptr(ssq_in), slots_in, float(eps), _ptr(ssq_out), prefetch.data_ptr(),
            prefetch.numel() * prefetch.element_size(), _stream()), "gemm_prefetch_bf16")
        return out
    _check(_load().b2tts_gemm_norm_bf16(x.data_ptr(), x.stride(0), packed.data_ptr(), _ptr(bias), _ptr(residual),
                                        res_stride, out.data_ptr(), out.stride(0), m, n, k, int(epilogue),
                                        int(rounding), int(splits), int(bool(pdl)), _ptr(norm_weight), _ptr(ssq_in),
                                        slots_in, float(eps), _ptr(ssq_out), _stream()), "gemm_norm_bf16")
    return out


# ------------------------------------------------------------------------------------------ a7 x 4: GEMM chain
class GemmOp(ctypes.Structure):
    """``b2tts_gemm_op`` (include/b200tts.h)."""
    _fields_ = [("x", c_void_p), ("x_row_stride", c_int64), ("w_packed", c_void_p),
                ("residual", c_void_p), ("residual_row_stride", c_int64),
                ("out", c_void_p), ("out_row_stride", c_int64),
                ("norm_weight", c_void_p), ("ssq_in", c_void_p), ("ssq_out", c_void_p),
                ("n", c_int), ("k", c_int), ("epilogue", c_int), ("ssq_in_slots", c_int)]


GEMM_CHAIN_MAX_OPS = 4


def gemm_chain_ssq_slots(n: int) -> int:
    return int(_load().b2tts_gemm_chain_ssq_slots(int(n)))


def gemm_chain_workspace(m: int, shapes, device) -> torch.Tensor:
    """Zero-filled workspace for chains of up to ``m`` rows whose ops have any of the (n, k) ``shapes``."""
    need = 0
    for i in range(0, len(shapes), GEMM_CHAIN_MAX_OPS):
        sub = shapes[i:i + GEMM_CHAIN_MAX_OPS]
        ns = (c_int * len(sub))(*[int(v[0]) for v in sub])
        ks = (c_int * len(sub))(*[int(v[1]) for v in sub])
        got = int(_load().b2tts_gemm_chain_workspace_bytes(int(m), ns, ks, len(sub)))
        if got == 0:
            raise RuntimeError(f"libb200tts gemm_chain_workspace: unsupported m={m} shapes={sub}")
        need = max(need, got)
    return torch.zeros(need + 16, dtype=torch.uint8, device=device)


def gemm_chain(ops: list[dict], m: int, workspace: torch.Tensor, eps: float = 1e-6, rounding: int = ROUND_EAGER,
               pdl: bool = True):
    """Up to 4 dependent GEMMs in one persistent launch (``b2tts_gemm_chain``).  Each op is a dict with
    ``x`` [m, k] bf16, ``packed`` (``gemm_pack_weight`` image), ``n``, ``out``, and optionally ``residual``,
    ``epilogue``, ``norm_weight`` + ``ssq_in`` [m, slots] fp32, ``ssq_out`` [m, n/128] fp32."""
    arr = (GemmOp * len(ops))()
    keep = []
    for i, o in enumerate(ops):
        x = o["x"]
        _cuda(x, BF16, "x")
        if x.stride(1) != 1:
            x = x.contiguous()
            keep.append(x)
        out = o["out"]
        _cuda(out, BF16, "out")
        res = o.get("residual")
        ssq_in, ssq_out = o.get("ssq_in"), o.get("ssq_out")
        n, k = int(o["n"]), int(x.shape[1])
        if ssq_out is not None and (ssq_out.dtype != F32 or not ssq_out.is_contiguous() or ssq_out.shape[1] != n // 128):
            raise ValueError(f"gemm_chain op {i}: ssq_out must be a contiguous [>=m, {n // 128}] fp32 tensor")
        if ssq_in is not None and (ssq_in.dtype != F32 or not ssq_in.is_contiguous()):
            raise ValueError(f"gemm_chain op {i}: ssq_in must be a contiguous [>=m, slots] fp32 tensor")
        a = arr[i]
        a.x, a.x_row_stride = x.data_ptr(), x.stride(0)
        a.w_packed = o["packed"].data_ptr()
        a.residual, a.residual_row_stride = _ptr(res), (res.stride(0) if res is not None else 0)
        a.out, a.out_row_stride = out.data_ptr(), out.stride(0)
        a.norm_weight = _ptr(o.get("norm_weight"))
        a.ssq_in, a.ssq_out = _ptr(ssq_in), _ptr(ssq_out)
        a.n, a.k, a.epilogue = n, k, int(o.get("epilogue", GEMM_BIAS))
        a.ssq_in_slots = int(ssq_in.shape[1]) if ssq_in is not None else 0
    _check(_load().b2tts_gemm_chain(ctypes.byref(arr), len(ops), int(m), float(eps), int(rounding),
                                    workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                                    int(bool(pdl)), _stream()), "gemm_chain")


# ------------------------------------------------------------------------------------------ a10 in one launch
class StackLayer(ctypes.Structure):
    """``b2tts_stack_layer`` (include/b200tts.h)."""
    _fields_ = [("qkv_w", c_void_p), ("o_w", c_void_p), ("gate_up_w", c_void_p), ("down_w", c_void_p),
                ("input_norm_w", c_void_p), ("post_norm_w", c_void_p), ("q_norm_w", c_void_p), ("k_norm_w", c_void_p),
                ("k_cache", c_void_p), ("v_cache", c_void_p)]


DECODER_STACK_MAX_M = 64
DECODER_STACK_MAX_LAYERS = 8


def decoder_stack_plan(hidden: int, q_size: int, kv_size: int, intermediate: int):
    """(o_splits, down_splits) of the persistent decoder-stack kernel for these widths, or None when the stack
    does not fit one job per SM (``b2tts_decoder_stack_plan``)."""
    so, sd = c_int(), c_int()
    ok = _load().b2tts_decoder_stack_plan(int(hidden), int(q_size), int(kv_size), int(intermediate),
                                          ctypes.byref(so), ctypes.byref(sd))
    return (so.value, sd.value) if ok else None


def decoder_stack_workspace(hidden: int, q_size: int, kv_size: int, intermediate: int, device) -> torch.Tensor:
    """Zero-filled, 256-byte aligned workspace of ``decoder_stack_decode``."""
    n = int(_load().b2tts_decoder_stack_workspace_bytes(int(hidden), int(q_size), int(kv_size), int(intermediate)))
    if n == 0:
        raise RuntimeError("libb200tts decoder_stack_workspace: unsupported widths")
    buf = torch.zeros(n + 256, dtype=torch.uint8, device=device)
    off = (-buf.data_ptr()) % 256
    return buf[off:off + n]


def decoder_stack_decode(x, layers: list[dict], final_norm_weight, positions, inv_freq, slot_mapping, block_table,
                         context_lens, *, num_heads, num_kv_heads, head_dim, intermediate, eps, scale, workspace,
                         attention_scaling=1.0, rounding=ROUND_EAGER, out=None, layer_array=None):
    """One decode step of a whole decoder stack in one persistent launch (``b2tts_decoder_stack_decode``).
    ``layers``: per layer a dict of tensors ``qkv_w, o_w, gate_up_w, down_w`` (packed images),
    ``input_norm_w, post_norm_w, q_norm_w, k_norm_w``, ``k_cache, v_cache``; ``layer_array`` a prebuilt
    ``StackLayer`` array (skips the per-call marshalling).  x [m, hidden] -> normalised hidden states [m, hidden]."""
    _cuda(x, BF16, "x"); _cuda(positions, I64, "positions"); _cuda(inv_freq, F32, "inv_freq")
    _cuda(slot_mapping, I32, "slot_mapping"); _cuda(block_table, I32, "block_table"); _cuda(context_lens, I32, "context_lens")
    m, hidden = x.shape
    if x.stride(1) != 1:
        x = x.contiguous()
    if out is None:
        out = torch.empty(m, hidden, dtype=BF16, device=x.device)
    arr = layer_array if layer_array is not None else stack_layer_array(layers)
    kc = layers[0]["k_cache"]
    _check(_load().b2tts_decoder_stack_decode(
        x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), ctypes.byref(arr), len(layers),
        final_norm_weight.data_ptr(), m, hidden, int(num_heads), int(num_kv_heads), int(head_dim), int(intermediate),
        float(eps), int(rounding), positions.data_ptr(), inv_freq.data_ptr(), float(attention_scaling),
        slot_mapping.data_ptr(), block_table.data_ptr(), block_table.stride(0), context_lens.data_ptr(),
        int(kc.shape[1]), float(scale), workspace.data_ptr(), workspace.numel() * workspace.element_size(), _stream()),
        "decoder_stack_decode")
    return out


def stack_layer_array(layers: list[dict]):
    arr = (StackLayer * len(layers))()
    for a, d in zip(arr, layers):
        for name, _ in StackLayer._fields_:
            setattr(a, name, d[name].data_ptr())
    return arr


# ------------------------------------------------------------------------------------------ f3: prompt assembly
def prompt_assemble(a_rows, b_rows, a_index, b_index, out):
    """out[r] = a_rows[a_index[r]] (+) b_rows[b_index[r]] (bf16 add; index < 0 = side absent) - ``b2tts_prompt_assemble``.
    a_rows [Na, H] / b_rows [Nb, H] bf16 (either may be empty), indices int32 [R], out [R, H] bf16 (row stride free)."""
    _cuda(out, BF16, "out"); _cuda(a_index, I32, "a_index"); _cuda(b_index, I32, "b_index")
    rows, cols = out.shape
    if a_index.numel() != rows or b_index.numel() != rows:
        raise ValueError("prompt_assemble: one index per output row")
    for t, name in ((a_rows, "a_rows"), (b_rows, "b_rows")):
        if t is not None and t.numel():
            _cuda(t, BF16, name)
            if t.shape[1] != cols or t.stride(1) != 1:
                raise ValueError(f"prompt_assemble: {name} must be [n, {cols}] with contiguous rows")
    na = 0 if a_rows is None else a_rows.shape[0]
    nb = 0 if b_rows is None else b_rows.shape[0]
    _check(_load().b2tts_prompt_assemble(
        a_rows.data_ptr() if na else None, a_rows.stride(0) if na else cols, na,
        b_rows.data_ptr() if nb else None, b_rows.stride(0) if nb else cols, nb,
        a_index.data_ptr(), b_index.data_ptr(), out.data_ptr(), out.stride(0), rows, cols, _stream()), "prompt_assemble")
    return out


# ------------------------------------------------------------------------------------------ f4 (tail): PCM16
def float_to_pcm16(wav: torch.Tensor, volume: float = 1.0, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp32 waveform (any shape, contiguous) -> int16 PCM on the device (``b2tts_float_to_pcm16``)."""
    _cuda(wav, F32, "wav")
    if not wav.is_contiguous():
        wav = wav.contiguous()
    if out is None:
        out = torch.empty(wav.shape, dtype=torch.int16, device=wav.device)
    _check(_load().b2tts_float_to_pcm16(wav.data_ptr(), out.data_ptr(), wav.numel(), float(volume), _stream()),
           "float_to_pcm16")
    return out
