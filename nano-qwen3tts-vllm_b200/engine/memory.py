"""KV-pool memory policy: how many 256-token pages each model gets.

Integer / float arithmetic only, no CUDA calls - the runner feeds it the live allocator numbers
(``ModelRunner.allocate_kv_cache``), the CPU tests feed it tables.  Restates, with the same operand order so
the results are bit-identical to the reference's Python floats (pinned by oracle/pin_against_reference.py
-> tests/golden/memory_policy.json):

  estimate_model_params   interface.py:57-77     parameter count of a talker / predictor stack
  kv_block_bytes          interface.py:80-84, engine/model_runner/base.py:139
  compute_memory_split    interface.py:87-197    one user number (gpu_memory_utilization) -> talker share,
                                                  predictor share ("the rest"), per-process fraction
  worker_utilization      workers/talker_worker.py:61-66, predictor_worker.py:57-62   the 0.85 cap per worker
  num_kv_blocks           engine/model_runner/base.py:131-151   pages from (total, free, allocated, peak, current)
"""
from __future__ import annotations

from ..config import ModelConfig, StackConfig

OVERHEAD_FACTOR = 1.5          # CUDA graphs, activations, fragmentation (interface.py:139)
WORKER_UTIL_CAP = 0.85         # talker_worker.py:66, predictor_worker.py:62


def estimate_model_params(cfg: StackConfig) -> int:
    """interface.py:57-77: q/k/v/o + gate/up/down + 2 norm vectors per layer, embedding + (untied) head."""
    h, i = cfg.hidden_size, cfg.intermediate_size
    head_dim = cfg.head_dim or (h // cfg.num_attention_heads)
    qkvo = h * (cfg.num_attention_heads * head_dim) + h * (cfg.num_key_value_heads * head_dim) * 2 \
        + (cfg.num_attention_heads * head_dim) * h
    per_layer = qkvo + h * i * 3 + h * 2
    return cfg.num_hidden_layers * per_layer + cfg.vocab_size * h + h * cfg.vocab_size


def kv_block_bytes(cfg: StackConfig, block_size: int = 256, dtype_bytes: int = 2) -> int:
    """interface.py:80-84 == base.py:139: 2 (K+V) * layers * block * kv_heads * head_dim * dtype."""
    head_dim = cfg.head_dim or (cfg.hidden_size // cfg.num_attention_heads)
    return 2 * cfg.num_hidden_layers * block_size * cfg.num_key_value_heads * head_dim * dtype_bytes


def compute_memory_split(model_config: ModelConfig, gpu_memory_utilization: float, total_vram: int,
                         kvcache_block_size: int = 256) -> dict:
    """interface.py:87-197.  ``total_vram`` = ``torch.cuda.get_device_properties(0).total_memory``.
    Returns talker_util, pred_util, process_gpu_memory_fraction (+ the intermediate byte counts)."""
    t, p = model_config.talker, model_config.predictor
    dtype_bytes = 2
    talker_weight_bytes = estimate_model_params(t) * dtype_bytes
    pred_weight_bytes = estimate_model_params(p) * dtype_bytes
    text_vocab = t.text_vocab_size or 151936                                  # getattr defaults, interface.py:131-132
    text_hidden = t.text_hidden_size or 2048
    talker_weight_bytes += text_vocab * text_hidden * dtype_bytes
    talker_weight_bytes += text_hidden * t.hidden_size * dtype_bytes
    talker_fixed = int(talker_weight_bytes * OVERHEAD_FACTOR)
    pred_fixed = int(pred_weight_bytes * OVERHEAD_FACTOR)
    kv_block_talker = kv_block_bytes(t, kvcache_block_size, dtype_bytes)
    kv_block_pred = kv_block_bytes(p, kvcache_block_size, dtype_bytes)
    process_fraction = gpu_memory_utilization
    effective_total = total_vram * process_fraction
    kv_budget = effective_total - talker_fixed - pred_fixed
    out = {"pred_util": 1.0, "process_gpu_memory_fraction": process_fraction, "talker_fixed": talker_fixed,
           "pred_fixed": pred_fixed, "kv_block_talker": kv_block_talker, "kv_block_pred": kv_block_pred}
    if kv_budget <= 0:                                                        # interface.py:151-162
        out["talker_util"] = max(0.05, (talker_fixed + kv_block_talker) / effective_total)
        out["blocks_each"] = 0
        return out
    n_blocks = kv_budget / (kv_block_talker + kv_block_pred)                  # equal page counts, :165
    talker_kv_bytes = n_blocks * kv_block_talker
    talker_util = (talker_fixed + talker_kv_bytes) / effective_total
    out["talker_util"] = max(0.05, min(talker_util, 0.95))                    # :172
    out["blocks_each"] = int(n_blocks)
    return out


def worker_utilization(split: dict, talker: bool) -> float:
    """What each worker passes as ``gpu_memory_utilization`` (talker_worker.py:61-66, predictor_worker.py:57-62)."""
    return min(split["talker_util" if talker else "pred_util"], WORKER_UTIL_CAP)


def num_kv_blocks(total: int, free: int, allocated: int, peak: int, current: int, gpu_memory_utilization: float,
                  process_gpu_memory_fraction: float | None, block_bytes: int) -> tuple[int, int]:
    """engine/model_runner/base.py:131-151.  ``total, free`` = ``torch.cuda.mem_get_info()``, ``allocated`` =
    ``torch.cuda.memory_allocated()``, ``peak/current`` = ``memory_stats()["allocated_bytes.all.{peak,current}"]``.
    Returns ``(num_kvcache_blocks, raw_blocks)``: at least 1 page; ``raw_blocks <= 0`` is the reference's warning
    case (:148-153)."""
    if process_gpu_memory_fraction is not None:
        effective_total = total * process_gpu_memory_fraction                  # :142-146 (several processes per GPU)
        raw = int(effective_total * gpu_memory_utilization - allocated - peak + current) // block_bytes
    else:
        used = total - free                                                    # :148-149
        raw = int(total * gpu_memory_utilization - used - peak + current) // block_bytes
    return max(1, raw), raw
