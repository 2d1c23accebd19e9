"""Multi-codebook code predictor (models/qwen3_tts_predictor.py:14-125 of the reference).

The 15 codec embeddings and 15 LM heads are single stacked tensors so that the per-sequence
Python loops of the reference (``codec_embedding[g_i-1](id_i)``, ``lm_head[g_i](h_i)``,
qwen3_tts_predictor.py:94-95,119-121) become one gather kernel with a per-row table index and one
GEMM per distinct generation step present in the batch (one, in practice: sequences of a burst
step together)."""
from __future__ import annotations

import torch
from torch import nn

from .. import _native as nat
from ..config import StackConfig
from ..layers.linear import ReplicatedLinear, linear, native_ok
from .qwen3_tts_share import DecoderStack, load_reference_weights


class Qwen3TTSCodePredictorModel(DecoderStack):
    def __init__(self, config: StackConfig, talker_config: StackConfig, rounding: int = nat.ROUND_EAGER):
        super().__init__(config, rounding)
        n = config.num_code_groups - 1
        # [n, vocab, talker_hidden]: embeddings live in the TALKER's width (predictor.py:24-27)
        self.codec_embedding = nn.Parameter(torch.empty(n, config.vocab_size, talker_config.hidden_size),
                                            requires_grad=False)

    def forward(self, input_embeds: torch.Tensor, positions: torch.Tensor, input_ssq=None, tail_weight=None) -> torch.Tensor:
        return self.run_layers(input_embeds, positions, input_ssq, tail_weight)


class Qwen3TTSCodePredictorForCausalLM(nn.Module):
    def __init__(self, config: StackConfig, talker_config: StackConfig, rounding: int = nat.ROUND_EAGER):
        super().__init__()
        self.config = config
        self.model = Qwen3TTSCodePredictorModel(config, talker_config, rounding)
        self.vocab_size = config.vocab_size
        n = config.num_code_groups - 1
        self.lm_head = nn.Parameter(torch.empty(n, config.vocab_size, config.hidden_size), requires_grad=False)
        if config.hidden_size != talker_config.hidden_size:
            self.small_to_mtp_projection = ReplicatedLinear(talker_config.hidden_size, config.hidden_size, bias=True)
        else:
            self.small_to_mtp_projection = None

    @staticmethod
    def convert_state_dict(state_dict: dict) -> dict:
        pre = "talker.code_predictor."
        return {k[len(pre):]: v for k, v in state_dict.items() if k.startswith(pre)}

    def load_state_dict(self, state_dict, strict=True):
        sd = self.convert_state_dict(state_dict)
        if not sd:                                   # already converted
            sd = dict(state_dict)
        rest, emb_seen, head_seen = {}, set(), set()
        with torch.no_grad():
            for k, v in sd.items():
                parts = k.split(".")
                if k.startswith("model.codec_embedding.") and parts[2].isdigit():
                    self.model.codec_embedding[int(parts[2])].copy_(v)
                    emb_seen.add(int(parts[2]))
                elif k.startswith("lm_head.") and parts[1].isdigit():
                    self.lm_head[int(parts[1])].copy_(v)
                    head_seen.add(int(parts[1]))
                else:
                    rest[k] = v
        n = self.lm_head.shape[0]
        if strict and (len(emb_seen) != n or len(head_seen) != n):
            raise KeyError(f"missing keys: codec_embedding {sorted(set(range(n)) - emb_seen)}, "
                           f"lm_head {sorted(set(range(n)) - head_seen)}")
        return load_reference_weights(self, rest, strict=strict, prefilled=("model.codec_embedding", "lm_head"))

    def get_input_embeddings(self, input_ids, input_embeds, generation_steps):
        """Prefill: the [hidden, embed(code0)] rows pass through.  Decode: row i is
        ``codec_embedding[generation_steps[i]-1][input_ids[i]]``."""
        if input_embeds is not None and input_embeds.shape[-2] > 1:
            return input_embeds
        steps = generation_steps if torch.is_tensor(generation_steps) else \
            torch.tensor(generation_steps, dtype=torch.int32, device=input_ids.device)
        return nat.gather_rows(self.model.codec_embedding, input_ids.view(-1), (steps - 1).to(torch.int32))

    def forward(self, inputs_embeds: torch.Tensor, positions: torch.Tensor, head: int | None = None) -> torch.Tensor:
        """``head``: the LM head the caller applies next (its weights are pulled into L2 by the last layer)."""
        proj, ssq = self.small_to_mtp_projection, None
        tail = self.lm_head[head] if head is not None else None
        if proj is not None:
            x2 = inputs_embeds.reshape(-1, inputs_embeds.shape[-1])
            if native_ok(x2.shape[0], proj.weight.shape[0], x2.shape[1]):
                # the projection's epilogue leaves the row sums of squares the first layer's norm needs
                inputs_embeds, ssq = linear(x2, proj.weight, proj.bias, want_ssq=True,
                                            nxt=self.model.layers[0].self_attn.qkv_proj.weight)
            else:
                inputs_embeds = proj(inputs_embeds)
        return self.model(inputs_embeds, positions, ssq, tail)

    def decode_logits(self, inputs_embeds: torch.Tensor, positions: torch.Tensor, head: int, nxt=None) -> torch.Tensor:
        """One decode pass + LM head ``head`` (one row per sequence): logits [n, vocab].  The stack's final RMSNorm rides
        inside the head GEMM (its input rows are normalised while they are staged) instead of being a launch of its own;
        batches that do not take the fused path get ``forward`` + ``linear``: same rounding points either way."""
        proj, ssq = self.small_to_mtp_projection, None
        x = inputs_embeds
        if proj is not None:
            x2 = x.reshape(-1, x.shape[-1])
            if not native_ok(x2.shape[0], proj.weight.shape[0], x2.shape[1]):
                return linear(self.forward(inputs_embeds, positions, head=head), self.lm_head[head], nxt=nxt)
            x, ssq = linear(x2, proj.weight, proj.bias, want_ssq=True, nxt=self.model.layers[0].self_attn.qkv_proj.weight)
        raw = self.model.run_layers_unnormed(x, positions, ssq, self.lm_head[head])
        if raw is None:
            return linear(self.model(x, positions, ssq, self.lm_head[head]), self.lm_head[head], nxt=nxt)
        return linear(raw[0], self.lm_head[head], norm=raw[1], nxt=nxt)

    def compute_logits(self, hidden_states: torch.Tensor, generation_steps: list[int]) -> torch.Tensor:
        """hidden_states [num_seqs * len, H] -> logits of the LAST position of every sequence through
        head ``generation_steps[i]`` (all sequences have equal length, predictor.py:118,124)."""
        nseq = len(generation_steps)
        per = hidden_states.shape[0] // nseq
        last = hidden_states.view(nseq, per, -1)[:, -1, :]
        uniq = sorted(set(generation_steps))
        if len(uniq) == 1:
            return linear(last, self.lm_head[uniq[0]])
        out = torch.empty(nseq, self.vocab_size, dtype=hidden_states.dtype, device=hidden_states.device)
        for g in uniq:
            idx = torch.tensor([i for i, s in enumerate(generation_steps) if s == g], device=out.device)
            out[idx] = linear(last[idx], self.lm_head[g])
        return out
