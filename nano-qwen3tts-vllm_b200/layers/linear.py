"""Linear layers with the reference's class names and ``weight`` / ``bias`` / ``weight_loader``
attributes (layers/linear.py:12-153).  Replicas only (tp_size == 1): no process group is needed
and no collective follows any GEMM.

``linear()`` is the one GEMM entry of the hot path: y[M,N] = x[M,K] @ W[N,K]^T (+ b).  In decode
M is the number of streams, so the GEMM streams W once from HBM: the hand-written sm_100a kernel
(``b2tts_gemm_bf16``: weights as the 128-row tcgen05 operand, the batch as the N tile, accumulators
in TMEM).  Larger row counts (decode batches above 64 streams, the predictor's 2-token prefill) run the
same kernel over 64-row chunks of x with the same fused epilogues; the chunks of one weight tile are
launched next to each other and share it through L2.  Beyond ``GEMM_NATIVE_MAX_M`` rows (long prompt
prefills) the product is compute-bound and goes to the plain library GEMM
(``torch.nn.functional.linear`` = cuBLAS), as it does for shapes the kernel cannot tile
(n % 128 != 0 or k % 64 != 0), which no shipped model has."""
import os

import torch
from torch import nn

from .. import _native as nat

# rows up to which the projections run on the weight-streaming kernel; beyond it (long prefills) the GEMM is
# compute-bound and the plain library GEMM (cuBLAS, 1.43 PFLOP/s at M = 4096 against our 0.93) takes over
GEMM_NATIVE_MAX_M = int(os.environ.get("B2TTS_GEMM_NATIVE_MAX_M", "512"))
PDL = os.environ.get("B2TTS_PDL_GEMM", "1") != "0"     # programmatic dependent launch of the native GEMM (weight stream starts early)
# each GEMM pulls the NEXT projection's weights into L2 (b2tts_gemm_prefetch_bf16).  Off by default: measured at 64
# streams it does not shorten the frame (5873 us against 5885 us; talker pass 2052 against 2018) - the launches are
# bound by their dependent latencies, not by the weight stream.
PREFETCH = os.environ.get("B2TTS_GEMM_PREFETCH", "0") != "0"
_packed: dict = {}


def packed_weight(weight: torch.Tensor, mode: int = nat.PACK_PLAIN) -> torch.Tensor:
    """The UMMA operand image of ``weight`` [out, in] (``b2tts_gemm_pack_weight``), built once per weight
    (again if the weight was modified in place) and cached by storage address.  Packing allocates, so
    the first call for a weight must happen outside CUDA-graph capture (warm-up runs do that)."""
    key = (weight.data_ptr(), tuple(weight.shape), mode)
    try:
        ver = weight._version
    except RuntimeError:            # inference tensors carry no version counter
        ver = -1
    ent = _packed.get(key)
    if ent is None or ent[0] != ver:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("libb200tts: weight packed for the first time under graph capture; run a warm-up first")
        # the entry keeps `weight` alive: while it exists the allocator cannot hand the same address to another
        # weight, so an address-keyed hit is always the right image (release_packed() drops a dead model's entries)
        ent = (ver, nat.gemm_pack_weight(weight, mode), weight)
        _packed[key] = ent
    return ent[1]


def release_packed(module: nn.Module) -> int:
    """Drop the packed images of every weight that lives inside ``module``'s parameters (call when the model is
    torn down, after its CUDA graphs are gone).  Returns the number of entries released."""
    spans = [(p.data_ptr(), p.data_ptr() + p.numel() * p.element_size()) for p in module.parameters()]
    dead = [k for k in _packed if any(lo <= k[0] < hi for lo, hi in spans)]
    for k in dead:
        del _packed[k]
    return len(dead)


def native_ok(m: int, n: int, k: int) -> bool:
    return 1 <= m <= GEMM_NATIVE_MAX_M and nat.gemm_supported(m, n, k)


class RowNorm:
    """An RMSNorm that has not been applied yet: the consuming GEMM normalises its input rows while it
    stages them (``b2tts_gemm_norm_bf16``).  ``ssq`` [M, slots] fp32 holds partial sums of squares of the
    rows, left by the GEMM epilogue that produced them."""
    __slots__ = ("weight", "ssq", "eps", "rounding")

    def __init__(self, weight, ssq, eps, rounding):
        self.weight, self.ssq, self.eps, self.rounding = weight, ssq, eps, rounding


def _norm_kw(norm):
    return {} if norm is None else dict(norm_weight=norm.weight, ssq_in=norm.ssq, eps=norm.eps, rounding=norm.rounding)


def _ssq_buffer(m, n, k, device):
    return torch.empty(m, nat.gemm_ssq_slots(m, n, k), dtype=torch.float32, device=device)


def _pf(nxt):
    """``nxt`` = the weight the NEXT GEMM of the chain will stream (a tensor, or (tensor, pack mode)) -> its packed
    image, for the L2 prefetch."""
    if nxt is None or not PREFETCH:
        return None
    if isinstance(nxt, tuple):
        return packed_weight(nxt[0], nxt[1])
    return packed_weight(nxt)


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
           out: torch.Tensor | None = None, norm: RowNorm | None = None, want_ssq: bool = False, nxt=None):
    """y = x W^T (+ b).  ``norm``: x is un-normalised and the GEMM applies the RMSNorm on the fly (native
    path only).  ``want_ssq``: also return the [M, slots] partial row sums of squares of y.  ``nxt``: the weight
    of the next GEMM on the stream (prefetched into L2 meanwhile)."""
    if not x.is_cuda:
        raise RuntimeError("libb200tts: linear() needs CUDA tensors (there is no CPU path)")
    x2 = x.reshape(-1, x.shape[-1])
    m, n = x2.shape[0], weight.shape[0]
    if native_ok(m, n, x2.shape[1]):
        ssq = _ssq_buffer(m, n, x2.shape[1], x.device) if want_ssq else None
        y = nat.gemm_bf16(x2, packed_weight(weight), n, bias=bias, out=out, pdl=PDL, ssq_out=ssq, prefetch=_pf(nxt),
                          **_norm_kw(norm))
        y = y.view(*x.shape[:-1], n)
        return (y, ssq) if want_ssq else y
    if norm is not None or want_ssq:
        raise RuntimeError(f"libb200tts: fused-norm GEMM unsupported for m={m} n={n} k={x2.shape[1]}")
    y = torch.nn.functional.linear(x2, weight, bias)
    if out is not None:
        out.copy_(y)
        y = out
    return y.view(*x.shape[:-1], n)


def linear_residual(x: torch.Tensor, weight: torch.Tensor, residual: torch.Tensor,
                    out: torch.Tensor | None = None, want_ssq: bool = False, nxt=None):
    """bf16(residual + bf16(x W^T)) - the projection followed by the decoder layer's residual add
    (models/qwen3_tts_share.py:171,175 of the reference) in the GEMM epilogue.  ``out`` may be ``residual``.
    ``want_ssq`` also returns the partial row sums of squares of the result (for the next fused norm)."""
    m, n = x.shape[0], weight.shape[0]
    if native_ok(m, n, x.shape[1]):
        ssq = _ssq_buffer(m, n, x.shape[1], x.device) if want_ssq else None
        y = nat.gemm_bf16(x, packed_weight(weight), n, residual=residual, epilogue=nat.GEMM_RESIDUAL, out=out, pdl=PDL,
                          ssq_out=ssq, prefetch=_pf(nxt))
        return (y, ssq) if want_ssq else y
    if want_ssq:
        raise RuntimeError(f"libb200tts: ssq epilogue unsupported for m={m} n={n} k={x.shape[1]}")
    y = torch.nn.functional.linear(x, weight)
    if out is None:
        return residual + y
    return torch.add(residual, y, out=out)


def linear_silu_mul(x: torch.Tensor, gate_up_weight: torch.Tensor, rounding: int = nat.ROUND_EAGER,
                    norm: RowNorm | None = None, nxt=None) -> torch.Tensor:
    """silu(x Wg^T) * (x Wu^T) over the merged [gate; up] weight (qwen3_tts_share.py:131-135) with the
    activation in the GEMM epilogue: the [M, 2I] intermediate never reaches HBM."""
    m, n = x.shape[0], gate_up_weight.shape[0]
    if native_ok(m, n, x.shape[1]):
        kw = _norm_kw(norm)
        kw.setdefault("rounding", rounding)
        return nat.gemm_bf16(x, packed_weight(gate_up_weight, nat.PACK_GATE_UP), n, epilogue=nat.GEMM_SILU_MUL,
                             pdl=PDL, prefetch=_pf(nxt), **kw)
    if norm is not None:
        raise RuntimeError(f"libb200tts: fused-norm GEMM unsupported for m={m} n={n} k={x.shape[1]}")
    gu = torch.nn.functional.linear(x, gate_up_weight)
    return nat.silu_and_mul(gu[:, :n // 2], gu[:, n // 2:], rounding)


class LinearBase(nn.Module):
    def __init__(self, input_size: int, output_size: int, bias: bool = False, tp_dim: int | None = None):
        super().__init__()
        self.tp_dim, self.tp_rank, self.tp_size = tp_dim, 0, 1
        self.weight = nn.Parameter(torch.empty(output_size, input_size), requires_grad=False)
        self.weight.weight_loader = self.weight_loader
        if bias:
            self.bias = nn.Parameter(torch.empty(output_size), requires_grad=False)
            self.bias.weight_loader = self.weight_loader
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor):
        with torch.no_grad():
            param.copy_(loaded_weight)           # not param.data: bumps the version packed_weight() watches

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return linear(x, self.weight, self.bias)


class ReplicatedLinear(LinearBase):
    pass


class ColumnParallelLinear(LinearBase):
    def __init__(self, input_size: int, output_size: int, bias: bool = False):
        super().__init__(input_size, output_size, bias, 0)


class RowParallelLinear(LinearBase):
    def __init__(self, input_size: int, output_size: int, bias: bool = False):
        super().__init__(input_size, output_size, bias, 1)


class MergedColumnParallelLinear(ColumnParallelLinear):
    """Several column-parallel projections sharing one input, stored as one weight so that one GEMM
    serves them (gate|up, q|k|v).  ``weight_loader(param, w, shard_id)`` fills one shard."""

    def __init__(self, input_size: int, output_sizes: list[int], bias: bool = False):
        self.output_sizes = list(output_sizes)
        super().__init__(input_size, sum(output_sizes), bias)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor, loaded_shard_id: int = None):
        with torch.no_grad():
            if loaded_shard_id is None:
                param.copy_(loaded_weight)
                return
            off = sum(self.output_sizes[:loaded_shard_id])
            param.narrow(0, off, self.output_sizes[loaded_shard_id]).copy_(loaded_weight)


class QKVParallelLinear(MergedColumnParallelLinear):
    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int, total_num_kv_heads: int | None = None,
                 bias: bool = False):
        kv = total_num_kv_heads or total_num_heads
        super().__init__(hidden_size, [total_num_heads * head_size, kv * head_size, kv * head_size], bias)

    def weight_loader(self, param, loaded_weight, loaded_shard_id=None):
        if isinstance(loaded_shard_id, str):
            loaded_shard_id = "qkv".index(loaded_shard_id)
        super().weight_loader(param, loaded_weight, loaded_shard_id)
