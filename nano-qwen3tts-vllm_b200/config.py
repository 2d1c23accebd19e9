"""Engine and model configuration for the hot path.

``Config`` keeps the field names of the reference's engine dataclass (config.py:41-66) so the
``TalkerLLM(model_path, enforce_eager=..., gpu_memory_utilization=...)`` call is unchanged; unlike the
reference it does NOT overwrite ``max_model_len`` with 1024 (config.py:64): the decode block-table
buffers are sized from the real value, so voice-clone prompts (~4k tokens, BASELINE config 4) run
on the graph path.  ``StackConfig`` is the subset of the HF ``config.json`` the decoder stacks read
(``talker_config`` and its nested ``code_predictor_config``, config.py:171-360,362-560).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field


@dataclass
class StackConfig:
    hidden_size: int = 1024
    intermediate_size: int = 3072
    num_hidden_layers: int = 28
    num_attention_heads: int = 16
    num_key_value_heads: int = 8
    head_dim: int = 128
    vocab_size: int = 3072
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1e6
    rope_scaling: dict | None = None          # {"mrope_section": [24,20,20], "interleaved": true}
    max_position_embeddings: int = 32768
    num_code_groups: int = 16
    text_vocab_size: int = 0
    text_hidden_size: int = 0

    @classmethod
    def from_dict(cls, d: dict) -> "StackConfig":
        known = {k: d[k] for k in cls.__dataclass_fields__ if k in d and (d[k] is not None or k == "rope_scaling")}
        c = cls(**known)
        if d.get("head_dim") is None:
            c.head_dim = c.hidden_size // c.num_attention_heads          # qwen3_tts_share.py:34
        return c

    @property
    def q_size(self) -> int:
        return self.num_attention_heads * self.head_dim

    @property
    def kv_size(self) -> int:
        return self.num_key_value_heads * self.head_dim


@dataclass
class ModelConfig:
    """Parsed ``config.json`` of a Qwen3-TTS checkpoint (Qwen3TTSConfig, config.py:564-609)."""
    talker: StackConfig
    predictor: StackConfig
    codec_eos_token_id: int = 2150            # hard-coded in the reference's frame loop, interface.py:1233

    @classmethod
    def from_dict(cls, d: dict) -> "ModelConfig":
        t = d.get("talker_config", d)
        p = t.get("code_predictor_config", {})
        talker = StackConfig.from_dict(t)
        pd = dict(p)
        pd.setdefault("num_code_groups", talker.num_code_groups)
        predictor = StackConfig.from_dict(pd)
        return cls(talker=talker, predictor=predictor)

    @classmethod
    def from_pretrained(cls, model_dir: str) -> "ModelConfig":
        with open(os.path.join(model_dir, "config.json")) as f:
            return cls.from_dict(json.load(f))

    def to_dict(self) -> dict:
        t = {k: getattr(self.talker, k) for k in StackConfig.__dataclass_fields__}
        t["code_predictor_config"] = {k: getattr(self.predictor, k) for k in StackConfig.__dataclass_fields__}
        return {"talker_config": t}


def qwen3_tts_config(size: str = "1.7B") -> ModelConfig:
    """The shipped model shapes; weights are random-init in every BASELINE
    config, so these numbers define the workload."""
    mrope = {"mrope_section": [24, 20, 20], "interleaved": True}
    if size in ("1.7B", "1.7b"):
        talker = StackConfig(hidden_size=2048, intermediate_size=6144, num_hidden_layers=28, vocab_size=3072,
                             rope_scaling=mrope, text_vocab_size=151936, text_hidden_size=2048)
    elif size in ("0.6B", "0.6b"):
        talker = StackConfig(hidden_size=1024, intermediate_size=3072, num_hidden_layers=28, vocab_size=3072,
                             rope_scaling=mrope, text_vocab_size=151936, text_hidden_size=2048)
    elif size == "tiny":                      # test-sized: same kernels, seconds on the CPU oracle
        talker = StackConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, vocab_size=3072,
                             rope_scaling=mrope)
        return ModelConfig(talker, StackConfig(hidden_size=128, intermediate_size=384, num_hidden_layers=2,
                                               vocab_size=2048))
    else:
        raise ValueError(f"unknown model size {size!r}")
    predictor = StackConfig(hidden_size=1024, intermediate_size=3072, num_hidden_layers=5, vocab_size=2048)
    return ModelConfig(talker, predictor)


@dataclass
class Config:
    """Engine configuration; field names follow the reference's ``Config`` (config.py:41-66)."""
    model: str = ""
    max_num_batched_tokens: int = 16384
    max_num_seqs: int = 512
    max_model_len: int = 4096
    gpu_memory_utilization: float = 0.9
    tensor_parallel_size: int = 1
    enforce_eager: bool = False
    hf_config: ModelConfig | None = None
    eos: int = -1
    kvcache_block_size: int = 256
    num_kvcache_blocks: int = -1
    distributed_port: int = 29500
    process_gpu_memory_fraction: float | None = None
    # additions (not in the reference)
    device: str = "cuda:0"
    rounding: int = 0                          # _native.ROUND_EAGER | ROUND_FUSED, see DESIGN.md
    graph_batch_sizes: list = field(default_factory=list)   # [] = reference's schedule

    def __post_init__(self):
        assert self.kvcache_block_size % 256 == 0                      # config.py:60
        assert self.tensor_parallel_size == 1, "replicas only: the hot path never shards a model (DESIGN.md)"
        assert self.max_num_batched_tokens >= self.max_model_len
        if self.hf_config is None and self.model:
            assert os.path.isdir(self.model), self.model               # config.py:59
            self.hf_config = ModelConfig.from_pretrained(self.model)
