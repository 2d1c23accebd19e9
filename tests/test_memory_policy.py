"""a11: the KV-pool memory policy restated in engine/memory.py against the numbers the
reference's OWN code produced for the same inputs (tests/golden/memory_policy.json, written by
oracle/pin_against_reference.py from interface.py:57-197 and engine/model_runner/base.py:131-164)."""
import json
import os

import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "memory_policy.json")


@pytest.fixture(scope="module")
def table():
    with open(GOLDEN) as f:
        return json.load(f)


def test_compute_memory_split_matches_reference(table):
    from nano_qwen3tts_vllm_b200.config import qwen3_tts_config
    from nano_qwen3tts_vllm_b200.engine import memory as mem
    assert len(table["split"]) >= 40
    for row in table["split"]:
        got = mem.compute_memory_split(qwen3_tts_config(row["model"]), row["gpu_memory_utilization"], row["total_vram"])
        for k in ("talker_util", "pred_util", "process_gpu_memory_fraction"):
            assert got[k] == row[k], (row, k, got[k])            # same operand order: bit-identical floats
        assert mem.worker_utilization(got, talker=False) == 0.85          # predictor "claims the rest", capped
        assert mem.worker_utilization(got, talker=True) == min(row["talker_util"], 0.85)


def test_num_kv_blocks_matches_reference(table):
    from nano_qwen3tts_vllm_b200.engine import memory as mem
    seen_min = False
    for row in table["blocks"]:
        got, raw = mem.num_kv_blocks(row["total"], row["free"], row["allocated"], row["peak"], row["current"],
                                     row["gpu_memory_utilization"], row["process_gpu_memory_fraction"], row["block_bytes"])
        assert got == row["num_kvcache_blocks"], row
        seen_min |= raw <= 0 and got == 1
    assert seen_min                                               # the "would be 0 -> 1 page" branch is in the table


def test_block_bytes_of_the_named_models():
    from nano_qwen3tts_vllm_b200.config import qwen3_tts_config
    from nano_qwen3tts_vllm_b200.engine import memory as mem
    mc = qwen3_tts_config("1.7B")
    assert mem.kv_block_bytes(mc.talker) == 28 << 20              # 28 MiB / 5 MiB
    assert mem.kv_block_bytes(mc.predictor) == 5 << 20
    split = mem.compute_memory_split(mc, 0.9, 191_514_542_080)
    assert split["blocks_each"] > 4000 and 0.05 <= split["talker_util"] <= 0.95
