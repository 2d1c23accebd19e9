"""Decoder-layer building blocks shared by the talker and the code predictor
(models/qwen3_tts_share.py:17-176 of the reference), re-cut for the B200 kernels:

  reference per layer (~37 launches)                     here (9 launches)
  ---------------------------------------------------    ----------------------------------------
  residual add, input RMSNorm                            add+RMSNorm (sum rounded to bf16 first)
  q_proj, k_proj, v_proj (3 GEMMs)                       one QKV GEMM over the merged weight
  q_norm, k_norm, ~20 eager RoPE ops, store_kvcache      one fused q/k-norm + RoPE + paged KV write
  flash_attn_with_kvcache | flash_attn_varlen_func       paged decode | varlen prefill attention
  o_proj                                                 o GEMM
  residual add, post-attention RMSNorm                   add+RMSNorm
  gate_proj, up_proj, silu*mul                           one gate_up GEMM, one SiLU*mul
  down_proj                                              down GEMM

Rounding points are the reference's (DESIGN.md "rounding points"); the state-dict keys of the
reference (``self_attn.{q,k,v,o}_proj.weight`` ...) are accepted by ``load_reference_weights`` and
merged at load time."""
from __future__ import annotations

import os

import torch
from torch import nn

from .. import _native as nat
from ..config import StackConfig
from ..layers.activation import SiluAndMul
from ..layers.attention import Attention
from ..layers.layernorm import Qwen3TTSRMSNorm
from ..layers.linear import (GEMM_NATIVE_MAX_M, MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear,
                             RowNorm, linear, linear_residual, linear_silu_mul, packed_weight)
from ..layers.rotary_embedding import get_rope
from ..utils.context import get_context


FUSED_DECODE_ATTENTION = True     # False: separate qknorm_rope_kvwrite + paged_attn_decode launches
GEMM_CHAIN = False                # decode: the 4 projections between two attention calls as ONE persistent launch
# experiments: RMSNorms applied by a separate launch instead of inside the consuming GEMM ("gate_up", "qkv", "gate_up,qkv")
UNFUSED_NORMS = os.environ.get("B2TTS_UNFUSED_NORMS", "")
# rows up to which the RMSNorms ride inside the consuming GEMM; above, the per-k-block hand-over of the in-place transform
# costs more than a separate launch.  Measured per frame with rmsnorm_kernel launched as a programmatic dependent (before
# that a separate norm drained the PDL chain and the fused path won up to 128 rows): 32 streams 4.33 fused / 4.64
# separate, 64: 5.15 / 5.51, 96: 6.17 / 5.79, 128: 6.12 / 5.70, 256: 8.68 / 8.20, 512: 12.4 / 10.95
FINAL_NORM_UNFUSED = os.environ.get("B2TTS_FINAL_NORM_UNFUSED", "0") != "0"      # experiments: the stack's final norm as its own launch
FUSED_NORM_MAX_ROWS = int(os.environ.get("B2TTS_FUSED_NORM_MAX_ROWS", "64"))
# ... and from which: below, the fixed cost of the fused path (rstd prologue + per-k-block hand-over) is not paid back
# (1 stream 3.83 fused / 3.75 separate, 4: 3.95 / 3.86, 8: 3.99 / 3.88, 12: 4.05 / 4.06, 16: 4.03 / 4.17)
FUSED_NORM_MIN_ROWS = int(os.environ.get("B2TTS_FUSED_NORM_MIN_ROWS", "9"))
# decode of a small stack (the code predictor) as ONE persistent dataflow launch (csrc/decoder_stack.cu).  Opt-in:
# bit-identical to the separate launches, but measured at 260 us per 5-layer pass against their 246 us at 64 rows
# (every hand-over between phases costs a gpu-scope release + an L2 poll, ~2.5 us, 7 per layer)
DATAFLOW_STACK = os.environ.get("B2TTS_DATAFLOW_STACK", "0") != "0"


class Qwen3TTSAttention(nn.Module):
    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER) -> None:
        super().__init__()
        self.num_heads = config.num_attention_heads
        self.num_kv_heads = config.num_key_value_heads
        self.head_dim = config.head_dim
        self.q_size = self.num_heads * self.head_dim
        self.kv_size = self.num_kv_heads * self.head_dim
        self.scaling = self.head_dim ** -0.5                              # qwen3_tts_share.py:37
        self.eps = config.rms_norm_eps
        self.rounding = rounding
        self.qkv_proj = QKVParallelLinear(config.hidden_size, self.head_dim, self.num_heads, self.num_kv_heads)
        self.o_proj = RowParallelLinear(self.q_size, config.hidden_size)
        self.q_norm = Qwen3TTSRMSNorm(self.head_dim, eps=self.eps, rounding=rounding)
        self.k_norm = Qwen3TTSRMSNorm(self.head_dim, eps=self.eps, rounding=rounding)
        self.rotary_emb = get_rope(self.head_dim, self.head_dim, config.max_position_embeddings, config.rope_theta,
                                   config.rope_scaling)
        self.attn = Attention(self.num_heads, self.head_dim, self.scaling, self.num_kv_heads)

    def forward(self, positions: torch.Tensor, hidden_states: torch.Tensor) -> torch.Tensor:
        o = self.attend(positions, hidden_states)
        return self.o_proj(o)

    def attend(self, positions: torch.Tensor, hidden_states: torch.Tensor, norm: RowNorm | None = None) -> torch.Tensor:
        """Everything up to (not including) o_proj: [n, heads * head_dim].  With ``norm`` the input is the
        un-normalised residual stream and the QKV GEMM applies the input RMSNorm while staging it."""
        return self.attend_qkv(positions, linear(hidden_states, self.qkv_proj.weight, norm=norm, nxt=self.o_proj.weight))

    def attend_qkv(self, positions: torch.Tensor, qkv: torch.Tensor) -> torch.Tensor:
        """``attend`` from the rows of the QKV projection (the GEMM chain of the previous layer produced them)."""
        ctx = get_context()
        n = qkv.shape[0]
        q = qkv[:, :self.q_size].view(n, self.num_heads, self.head_dim)
        k = qkv[:, self.q_size:self.q_size + self.kv_size].view(n, self.num_kv_heads, self.head_dim)
        v = qkv[:, self.q_size + self.kv_size:].view(n, self.num_kv_heads, self.head_dim)
        have_cache = self.attn.k_cache.numel() > 0
        packed_kv = ctx.is_prefill and ctx.block_tables is None           # prefill attends to the projections
        rope = self.rotary_emb
        if have_cache and not ctx.is_prefill and self.attn.fused_workspace is not None and FUSED_DECODE_ATTENTION:
            # decode: q/k-norm + RoPE + KV write + paged attention + split-KV combine in one launch
            max_ctx = ctx.max_context_len or ctx.block_tables.shape[1] * self.attn.k_cache.shape[1]
            o = nat.fused_decode_attn(
                q, k, v, self.q_norm.weight, self.k_norm.weight, self.eps, positions, rope.inv_freq, ctx.slot_mapping,
                self.attn.k_cache, self.attn.v_cache, ctx.block_tables, ctx.context_lens, self.scaling, max_ctx,
                self.attn.fused_workspace, modality=rope.modality if positions.dim() > 1 else None,
                attention_scaling=rope.attention_scaling, rounding=self.rounding)
            return o.view(n, self.q_size)
        q_rot, k_rot = nat.qknorm_rope_kvwrite(
            q, k, v, self.q_norm.weight, self.k_norm.weight, self.eps, positions, rope.inv_freq,
            modality=rope.modality if positions.dim() > 1 else None, attention_scaling=rope.attention_scaling,
            slot_mapping=ctx.slot_mapping if have_cache else None,
            k_cache=self.attn.k_cache if have_cache else None, v_cache=self.attn.v_cache if have_cache else None,
            want_k_out=packed_kv or not have_cache, rounding=self.rounding)
        o = self.attn.attend(q_rot, k_rot, v)
        return o.view(n, self.q_size)


class Qwen3TTSTalkerTextMLP(nn.Module):
    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER) -> None:
        super().__init__()
        self.gate_up_proj = MergedColumnParallelLinear(config.hidden_size, [config.intermediate_size] * 2)
        self.down_proj = RowParallelLinear(config.intermediate_size, config.hidden_size)
        self.act_fn = SiluAndMul(rounding)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.down_proj(self.act_fn(self.gate_up_proj(x)))


class Qwen3TTSDecoderLayer(nn.Module):
    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER) -> None:
        super().__init__()
        self.self_attn = Qwen3TTSAttention(config, rounding)
        self.mlp = Qwen3TTSTalkerTextMLP(config, rounding)
        self.input_layernorm = Qwen3TTSRMSNorm(config.hidden_size, eps=config.rms_norm_eps, rounding=rounding)
        self.post_attention_layernorm = Qwen3TTSRMSNorm(config.hidden_size, eps=config.rms_norm_eps, rounding=rounding)
        self.eps = config.rms_norm_eps
        self.rounding = rounding

    def forward(self, positions: torch.Tensor, hidden_states: torch.Tensor, residual: torch.Tensor | None):
        """Returns (block output, residual stream).  ``residual + hidden`` is rounded to bf16 before the
        norm, which is what the reference's separate add + rms_forward computes
        (qwen3_tts_share.py:163-176)."""
        if residual is None:
            residual = hidden_states
            hidden_states = self.input_layernorm(hidden_states)
        else:
            hidden_states, residual = nat.add_rmsnorm(hidden_states, residual, self.input_layernorm.weight, self.eps,
                                                      self.rounding, round_sum_first=True)
        hidden_states = self.self_attn(positions, hidden_states)
        hidden_states, residual = nat.add_rmsnorm(hidden_states, residual, self.post_attention_layernorm.weight,
                                                  self.eps, self.rounding, round_sum_first=True)
        hidden_states = self.mlp(hidden_states)
        return hidden_states, residual

    def forward_decode(self, positions: torch.Tensor, residual: torch.Tensor, ssq: torch.Tensor | None, in_place: bool,
                       nxt=None):
        """The same layer on the residual stream alone (decode and prefill, any row count), 5 launches:
        QKV GEMM (input RMSNorm applied while x is staged), q/k-norm + RoPE + KV write, attention,
        o_proj GEMM (+ residual add, row sums of squares out), gate_up GEMM (post-attention RMSNorm in,
        SiLU*mul out), down GEMM (+ residual add, row sums of squares out).  ``ssq`` = partial row sums of
        squares of ``residual`` from the GEMM that produced it (None: a separate RMSNorm launch).
        Rounding points are unchanged: bf16(residual + bf16(x W^T)), then the norm of the rounded sum.
        ``in_place`` False keeps the caller's tensor (the first layer's input embeddings) intact.
        ``nxt``: the weight the GEMM after this layer streams (the next layer's qkv_proj, or the LM head): every GEMM
        here pulls its successor's weights into L2 while it runs.  Returns (residual, ssq)."""
        wide = not FUSED_NORM_MIN_ROWS <= residual.shape[0] <= FUSED_NORM_MAX_ROWS
        if ssq is None or wide or "qkv" in UNFUSED_NORMS:
            h = nat.rmsnorm(residual, self.input_layernorm.weight, self.eps, self.rounding)
            o = self.self_attn.attend(positions, h)
        else:
            o = self.self_attn.attend(positions, residual,
                                      RowNorm(self.input_layernorm.weight, ssq, self.eps, self.rounding))
        # wide batches: the norms are separate launches, so nobody reads the row sums of squares - the GEMMs skip that epilogue
        residual = linear_residual(o, self.self_attn.o_proj.weight, residual, out=residual if in_place else None,
                                   want_ssq=not wide, nxt=(self.mlp.gate_up_proj.weight, nat.PACK_GATE_UP))
        residual, ssq = (residual, None) if wide else residual
        if wide or "gate_up" in UNFUSED_NORMS:
            h = nat.rmsnorm(residual, self.post_attention_layernorm.weight, self.eps, self.rounding)
            act = linear_silu_mul(h, self.mlp.gate_up_proj.weight, self.rounding, nxt=self.mlp.down_proj.weight)
        else:
            act = linear_silu_mul(residual, self.mlp.gate_up_proj.weight, self.rounding,
                                  norm=RowNorm(self.post_attention_layernorm.weight, ssq, self.eps, self.rounding),
                                  nxt=self.mlp.down_proj.weight)
        if wide:
            return linear_residual(act, self.mlp.down_proj.weight, residual, out=residual, nxt=nxt), None
        return linear_residual(act, self.mlp.down_proj.weight, residual, out=residual, want_ssq=True, nxt=nxt)


class DecoderStack(nn.Module):
    """``layers`` + final ``norm`` (Qwen3TTSTalkerModel / Qwen3TTSCodePredictorModel bodies)."""

    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER) -> None:
        super().__init__()
        self.config = config
        self.rounding = rounding
        self.layers = nn.ModuleList([Qwen3TTSDecoderLayer(config, rounding) for _ in range(config.num_hidden_layers)])
        self.norm = Qwen3TTSRMSNorm(config.hidden_size, eps=config.rms_norm_eps, rounding=rounding)

    def _chain_supported(self) -> bool:
        ok = getattr(self, "_chain_ok", None)
        if ok is None:
            c = self.config
            ok = self._chain_ok = all(nat.gemm_supported(1, n, k) for n, k in (
                (c.hidden_size, c.q_size), (2 * c.intermediate_size, c.hidden_size),
                (c.hidden_size, c.intermediate_size), (c.q_size + 2 * c.kv_size, c.hidden_size)))
        return ok

    def _native_gemm_shapes(self) -> bool:
        """All four projections have n % 128 == 0 and k % 64 == 0 (true for every shipped model size)."""
        ok = getattr(self, "_native_ok", None)
        if ok is None:
            c = self.config
            ok = self._native_ok = all(nat.gemm_supported(1, n, k) for n, k in (
                (c.hidden_size, c.q_size), (2 * c.intermediate_size, c.hidden_size),
                (c.hidden_size, c.intermediate_size), (c.q_size + 2 * c.kv_size, c.hidden_size)))
        return ok

    def _chain_workspace(self, device) -> torch.Tensor:
        ws = getattr(self, "_chain_ws", None)
        if ws is None or ws.device != torch.device(device):
            c = self.config
            shapes = [(c.hidden_size, c.q_size), (2 * c.intermediate_size, c.hidden_size),
                      (c.hidden_size, c.intermediate_size), (c.q_size + 2 * c.kv_size, c.hidden_size)]
            ws = self._chain_ws = nat.gemm_chain_workspace(nat.GEMM_CHAIN_MAX_M, shapes, device)
        return ws

    def run_layers_chained(self, input_embeds: torch.Tensor, positions: torch.Tensor,
                           input_ssq: torch.Tensor | None = None) -> torch.Tensor:
        """Decode (n <= GEMM_NATIVE_MAX_M rows): per layer ONE attention launch and ONE persistent GEMM-chain
        launch (o_proj + residual, gate_up with the post-attention norm and SiLU*mul, down_proj + residual, the
        next layer's qkv_proj with its input norm) - 2 launches per layer instead of 5, and the weight stream
        does not stop between the projections.  Same rounding points as ``forward_decode``."""
        c, eps, rnd = self.config, self.config.rms_norm_eps, self.rounding
        m, dev = input_embeds.shape[0], input_embeds.device
        ws = self._chain_workspace(dev)
        first = self.layers[0]
        if input_ssq is None:
            h = nat.rmsnorm(input_embeds, first.input_layernorm.weight, eps, rnd)
            qkv = linear(h, first.self_attn.qkv_proj.weight)
        else:
            qkv = linear(input_embeds, first.self_attn.qkv_proj.weight,
                         norm=RowNorm(first.input_layernorm.weight, input_ssq, eps, rnd))
        residual = input_embeds
        slots = nat.gemm_chain_ssq_slots(c.hidden_size)
        ssq1 = torch.empty(m, slots, dtype=torch.float32, device=dev)
        ssq2 = torch.empty(m, slots, dtype=torch.float32, device=dev)
        act = torch.empty(m, c.intermediate_size, dtype=input_embeds.dtype, device=dev)
        nlayers = len(self.layers)
        for i, layer in enumerate(self.layers):
            o = layer.self_attn.attend_qkv(positions, qkv)
            out = residual if i > 0 else torch.empty_like(residual)      # the caller's embeddings stay intact
            ops = [dict(x=o, packed=packed_weight(layer.self_attn.o_proj.weight), n=c.hidden_size, residual=residual,
                        out=out, epilogue=nat.GEMM_RESIDUAL, ssq_out=ssq1),
                   dict(x=out, packed=packed_weight(layer.mlp.gate_up_proj.weight, nat.PACK_GATE_UP),
                        n=2 * c.intermediate_size, out=act, epilogue=nat.GEMM_SILU_MUL,
                        norm_weight=layer.post_attention_layernorm.weight, ssq_in=ssq1),
                   dict(x=act, packed=packed_weight(layer.mlp.down_proj.weight), n=c.hidden_size, residual=out, out=out,
                        epilogue=nat.GEMM_RESIDUAL, ssq_out=ssq2)]
            if i + 1 < nlayers:
                nxt = self.layers[i + 1]
                qkv = torch.empty(m, c.q_size + 2 * c.kv_size, dtype=input_embeds.dtype, device=dev)
                ops.append(dict(x=out, packed=packed_weight(nxt.self_attn.qkv_proj.weight), n=qkv.shape[1], out=qkv,
                                norm_weight=nxt.input_layernorm.weight, ssq_in=ssq2))
            nat.gemm_chain(ops, m, ws, eps=eps, rounding=rnd)
            residual = out
        return nat.rmsnorm(residual, self.norm.weight, eps, rnd)

    def dataflow_plan(self):
        """(o_splits, down_splits) when this stack can run its decode pass as one persistent dataflow launch
        (every (projection, tile, split) job on its own SM, 1-D RoPE, <= 8 layers), else None."""
        plan = getattr(self, "_df_plan", False)
        if plan is False:
            c = self.config
            rope = self.layers[0].self_attn.rotary_emb
            plan = None
            if (len(self.layers) <= nat.DECODER_STACK_MAX_LAYERS and rope.modality is None and c.head_dim == 128
                    and c.num_attention_heads == 2 * c.num_key_value_heads):
                plan = nat.decoder_stack_plan(c.hidden_size, c.q_size, c.kv_size, c.intermediate_size)
            self._df_plan = plan
        return plan

    def run_layers_dataflow(self, input_embeds: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        """The whole decode pass (all layers + final norm, attention and KV write included) in one launch."""
        c, ctx = self.config, get_context()
        dev = input_embeds.device
        st = getattr(self, "_df_state", None)
        if st is None or st[0] != dev or st[1] != self.layers[0].self_attn.attn.k_cache.data_ptr():
            layers = []
            for layer in self.layers:
                a, m = layer.self_attn, layer.mlp
                layers.append(dict(
                    qkv_w=packed_weight(a.qkv_proj.weight), o_w=packed_weight(a.o_proj.weight),
                    gate_up_w=packed_weight(m.gate_up_proj.weight, nat.PACK_GATE_UP), down_w=packed_weight(m.down_proj.weight),
                    input_norm_w=layer.input_layernorm.weight, post_norm_w=layer.post_attention_layernorm.weight,
                    q_norm_w=a.q_norm.weight, k_norm_w=a.k_norm.weight, k_cache=a.attn.k_cache, v_cache=a.attn.v_cache))
            ws = nat.decoder_stack_workspace(c.hidden_size, c.q_size, c.kv_size, c.intermediate_size, dev)
            st = self._df_state = (dev, layers[0]["k_cache"].data_ptr(), layers, nat.stack_layer_array(layers), ws)
        _, _, layers, arr, ws = st
        rope = self.layers[0].self_attn.rotary_emb
        return nat.decoder_stack_decode(
            input_embeds, layers, self.norm.weight, positions, rope.inv_freq, ctx.slot_mapping, ctx.block_tables,
            ctx.context_lens, num_heads=c.num_attention_heads, num_kv_heads=c.num_key_value_heads, head_dim=c.head_dim,
            intermediate=c.intermediate_size, eps=c.rms_norm_eps, scale=self.layers[0].self_attn.scaling, workspace=ws,
            attention_scaling=rope.attention_scaling, rounding=self.rounding, layer_array=arr)

    def _run_layers_native(self, input_embeds, positions, input_ssq, tail_weight):
        residual, ssq = input_embeds, input_ssq
        nl = len(self.layers)
        for i, layer in enumerate(self.layers):
            nxt = self.layers[i + 1].self_attn.qkv_proj.weight if i + 1 < nl else tail_weight
            residual, ssq = layer.forward_decode(positions, residual, ssq, in_place=i > 0, nxt=nxt)
        return residual, ssq

    def run_layers_unnormed(self, input_embeds: torch.Tensor, positions: torch.Tensor,
                            input_ssq: torch.Tensor | None = None, tail_weight=None):
        """The layers WITHOUT the final norm: (residual stream, the final norm as a ``RowNorm`` for the GEMM that consumes
        it), or None when this batch does not take the fused-epilogue path (the caller then uses ``run_layers``)."""
        ctx = get_context()
        m = input_embeds.shape[0]
        if (DATAFLOW_STACK or GEMM_CHAIN or FINAL_NORM_UNFUSED or input_embeds.dim() != 2 or not max(1, FUSED_NORM_MIN_ROWS) <= m <= FUSED_NORM_MAX_ROWS
                or m > GEMM_NATIVE_MAX_M or not self._native_gemm_shapes()):
            return None
        residual, ssq = self._run_layers_native(input_embeds, positions, input_ssq, tail_weight)
        return residual, RowNorm(self.norm.weight, ssq, self.config.rms_norm_eps, self.rounding)

    def run_layers(self, input_embeds: torch.Tensor, positions: torch.Tensor,
                   input_ssq: torch.Tensor | None = None, tail_weight=None) -> torch.Tensor:
        """``tail_weight``: the weight streamed right after the stack (the LM head), prefetched by the last layer."""
        ctx = get_context()
        if (DATAFLOW_STACK and not ctx.is_prefill and ctx.block_tables is not None and positions.dim() == 1
                and 1 <= input_embeds.shape[0] <= nat.DECODER_STACK_MAX_M and input_embeds.dim() == 2
                and self.layers[0].self_attn.attn.k_cache.numel() > 0 and self.dataflow_plan() is not None):
            return self.run_layers_dataflow(input_embeds, positions)
        if GEMM_CHAIN and 1 <= input_embeds.shape[0] <= nat.GEMM_CHAIN_MAX_M and self._chain_supported():
            return self.run_layers_chained(input_embeds, positions, input_ssq)
        if 1 <= input_embeds.shape[0] <= GEMM_NATIVE_MAX_M and self._native_gemm_shapes():
            # decode and short prefills (<= GEMM_NATIVE_MAX_M rows): 5 launches per layer, norms / residual adds / SiLU*mul inside the GEMMs
            residual, _ = self._run_layers_native(input_embeds, positions, input_ssq, tail_weight)
            return nat.rmsnorm(residual, self.norm.weight, self.config.rms_norm_eps, self.rounding)
        hidden_states, residual = input_embeds, None
        for layer in self.layers:
            hidden_states, residual = layer(positions, hidden_states, residual)
        out, _ = nat.add_rmsnorm(hidden_states, residual, self.norm.weight, self.config.rms_norm_eps, self.rounding,
                                 round_sum_first=True)
        return out


_MERGED = {"q_proj": ("qkv_proj", 0), "k_proj": ("qkv_proj", 1), "v_proj": ("qkv_proj", 2),
           "gate_proj": ("gate_up_proj", 0), "up_proj": ("gate_up_proj", 1)}


def load_reference_weights(module: nn.Module, state_dict: dict, strict: bool = True,
                           prefilled: tuple[str, ...] = ()) -> list[str]:
    """Load a state dict that uses the reference's (converted) key names into ``module``, merging
    q/k/v and gate/up shards into the fused parameters.  Returns the keys that were not consumed.
    ``strict``: unexpected keys AND parameters (or shards of merged parameters) left unfilled raise;
    ``prefilled`` names parameters the caller has already loaded."""
    params = dict(module.named_parameters())
    unused = []
    filled: dict[str, set] = {}
    for name, w in state_dict.items():
        target, shard = name, None
        for src, (dst, sid) in _MERGED.items():
            if f".{src}." in name:
                target, shard = name.replace(f".{src}.", f".{dst}."), sid
                break
        p = params.get(target)
        if p is None:
            unused.append(name)
            continue
        loader = getattr(p, "weight_loader", None)
        filled.setdefault(target, set()).add(shard)
        if shard is not None:
            loader(p, w, shard)
        elif loader is not None:
            loader(p, w)
        else:
            with torch.no_grad():
                p.copy_(w)
    if strict and unused:
        raise KeyError(f"unexpected keys: {unused[:8]}{'...' if len(unused) > 8 else ''}")
    if strict:
        missing = [n for n in params if n not in filled and n not in prefilled]
        for n, shards in filled.items():
            want = 3 if n.endswith("qkv_proj.weight") else 2 if n.endswith("gate_up_proj.weight") else 0
            if want and None not in shards and len(shards) < want:
                missing.append(f"{n} (shards {sorted(shards)} of {want})")
        if missing:
            raise KeyError(f"missing keys: {missing[:8]}{'...' if len(missing) > 8 else ''}")
    return unused


@torch.no_grad()
def random_init_(module: nn.Module, seed: int = 1234, std: float = 0.02) -> None:
    """Random-init weights in place on the module's device (every BASELINE config uses random
    weights): matrices / embeddings N(0, std^2) (``initializer_range``, config.py:299,484), norm weights
    1 + N(0, std^2), biases 0 - the distribution of SURVEY.md section 8d."""
    gen = torch.Generator(device=next(module.parameters()).device).manual_seed(seed)
    for name, p in module.named_parameters():
        if name.endswith("bias"):
            p.zero_()
        elif "norm" in name and p.dim() == 1:
            p.copy_(1.0 + torch.randn(p.shape, generator=gen, device=p.device, dtype=torch.float32) * std)
        else:
            p.copy_(torch.randn(p.shape, generator=gen, device=p.device, dtype=torch.float32) * std)
