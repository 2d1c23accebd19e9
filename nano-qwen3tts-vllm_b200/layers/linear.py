"""Linear layers with the reference's class names and ``weight`` / ``bias`` / ``weight_loader``
attributes (layers/linear.py:12-153).  Replicas only (tp_size == 1): no process group is needed
and no collective follows any GEMM.

``linear()`` is the one GEMM entry of the hot path: y[M,N] = x[M,K] @ W[N,K]^T (+ b).  In decode
M is the number of streams, so the GEMM streams W once from HBM; the hand-written sm_100a kernel
(``b2tts_gemm_bf16``: weights as the 128-row tcgen05 operand, the batch as the N tile, accumulators
in TMEM) serves M <= GEMM_NATIVE_MAX_M, and cuBLASLt (through torch) serves the large-M prefill
contractions - a plain library GEMM."""
import torch
from torch import nn

from .. import _native as nat

GEMM_NATIVE_MAX_M = 256
_ws = {}


def _workspace(dev, nbytes):
    w = _ws.get(dev)
    if w is None or w.numel() < nbytes:
        w = torch.empty(max(nbytes, 1 << 22), dtype=torch.uint8, device=dev)
        _ws[dev] = w
    return w


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
           out: torch.Tensor | None = None) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError("libb200tts: linear() needs CUDA tensors (there is no CPU path)")
    x2 = x.reshape(-1, x.shape[-1])
    m, n = x2.shape[0], weight.shape[0]
    if nat.has_gemm() and m <= GEMM_NATIVE_MAX_M and nat.gemm_supported(m, n, x2.shape[1]):
        y = nat.gemm_bf16(x2, weight, bias, out=out, workspace_fn=_workspace)
    else:
        y = torch.nn.functional.linear(x2, weight, bias)
        if out is not None:
            out.copy_(y)
            y = out
    return y.view(*x.shape[:-1], n)


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
        param.data.copy_(loaded_weight)

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
        if loaded_shard_id is None:
            param.data.copy_(loaded_weight)
            return
        off = sum(self.output_sizes[:loaded_shard_id])
        param.data.narrow(0, off, self.output_sizes[loaded_shard_id]).copy_(loaded_weight)


class QKVParallelLinear(MergedColumnParallelLinear):
    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int, total_num_kv_heads: int | None = None,
                 bias: bool = False):
        kv = total_num_kv_heads or total_num_heads
        super().__init__(hidden_size, [total_num_heads * head_size, kv * head_size, kv * head_size], bias)

    def weight_loader(self, param, loaded_weight, loaded_shard_id=None):
        if isinstance(loaded_shard_id, str):
            loaded_shard_id = "qkv".index(loaded_shard_id)
        super().weight_loader(param, loaded_weight, loaded_shard_id)
