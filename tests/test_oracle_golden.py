"""CPU: the oracle restatement reproduces the committed outputs of the LIVE reference
(tests/golden/*.pt, written by oracle/pin_against_reference.py) bit for bit."""
import torch

from oracle import ops as O

BF16 = torch.bfloat16


def test_rmsnorm(golden_ops):
    for name in ("rms_hidden", "rms_head", "rms_1024"):
        v = golden_ops[name]
        assert torch.equal(O.rmsnorm(v["x"], v["w"], v["eps"]), v["y"])
        y, r = O.add_rmsnorm(v["x"], v["res"], v["w"], v["eps"])
        assert torch.equal(y, v["y_add"]) and torch.equal(r, v["res_add"])


def test_rope(golden_ops):
    v = golden_ops["rope_1d"]
    q, k = O.rope(v["positions"], v["q"], v["k"], v["inv_freq"])
    assert torch.equal(q, v["q_out"]) and torch.equal(k, v["k_out"])
    for name in ("mrope_interleaved", "mrope_sectioned"):
        v = golden_ops[name]
        mod = O.mrope_dim_modality(128, v["mrope_section"], v["interleaved"])
        assert mod == v["modality"].tolist()
        q, k = O.rope(v["positions3"], v["q"], v["k"], v["inv_freq"], 1.0, mod)
        assert torch.equal(q, v["q_out"]) and torch.equal(k, v["k_out"])


def test_silu(golden_ops):
    v = golden_ops["silu"]
    assert torch.equal(O.silu(v["gate"]), v["silu"])
    assert torch.equal(O.silu_and_mul(v["gate"], v["up"]), v["silu_mul_eager"])
    assert torch.equal(O.silu_and_mul(v["gate"], v["up"], fused_rounding=True), v["silu_mul_fused"])


def test_sampler(golden_ops):
    for name in ("sampler_talker", "sampler_predictor", "sampler_greedy"):
        v = golden_ops[name]
        p = O.topk_probs(v["logits"], v["temperatures"], v["top_k"])
        assert torch.equal(p, v["probs"])
        assert torch.equal(O.sample_from_probs(p, v["noise"]), v["index"])


def test_store_kvcache_semantics():
    g = torch.Generator().manual_seed(0)
    k = torch.randn(5, 8, 128, generator=g).to(BF16)
    v = torch.randn(5, 8, 128, generator=g).to(BF16)
    kc = torch.zeros(2, 256, 8, 128, dtype=BF16)
    vc = torch.zeros_like(kc)
    slots = torch.tensor([3, -1, 511, 256, 0], dtype=torch.int32)
    O.store_kvcache(k, v, kc, vc, slots)
    flat = kc.view(-1, 1024)
    assert torch.equal(flat[3], k[0].reshape(-1)) and torch.equal(flat[511], k[2].reshape(-1))
    assert torch.equal(flat[256], k[3].reshape(-1)) and torch.equal(flat[0], k[4].reshape(-1))
    assert float(flat[1].abs().sum()) == 0.0 and int((kc != 0).any(-1).any(-1).sum()) == 4


def test_attention_restatement_consistency():
    """decode over a paged cache == last row of varlen prefill over the same tokens."""
    g = torch.Generator().manual_seed(1)
    t = 300
    q = torch.randn(t, 16, 128, generator=g).to(BF16)
    k = torch.randn(t, 8, 128, generator=g).to(BF16)
    v = torch.randn(t, 8, 128, generator=g).to(BF16)
    cu = torch.tensor([0, t], dtype=torch.int32)
    full = O.varlen_attn_prefill(q, k, v, cu, cu, 128 ** -0.5)
    kc = torch.zeros(3, 256, 8, 128, dtype=BF16)
    vc = torch.zeros_like(kc)
    slots = torch.cat([torch.arange(512, 768), torch.arange(0, 44)]).to(torch.int32)
    O.store_kvcache(k, v, kc, vc, slots)
    bt = torch.tensor([[2, 0, -1]], dtype=torch.int32)
    dec = O.paged_attn_decode(q[-1:], kc, vc, torch.tensor([t], dtype=torch.int32), bt, 128 ** -0.5)
    assert torch.equal(dec[0], full[-1])


def test_prompt_oracle_reproduces_the_reference_golden():
    """oracle/prompt.py against tests/golden/prompt.pt (outputs of the live utils/generation.py, written by
    oracle/pin_against_reference.py): every mode, every byte."""
    import os
    import types
    import torch
    from oracle import prompt as op
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    gold = torch.load(os.path.join(root, "tests", "golden", "prompt.pt"))
    w = gold["weights"]
    tc = types.SimpleNamespace(
        num_code_groups=16, codec_nothink_id=5, codec_think_id=6, codec_think_bos_id=7, codec_think_eos_id=8,
        codec_pad_id=9, codec_bos_id=10, codec_language_id={"english": 20, "chinese": 21, "sichuan": 22},
        spk_id={"anna": 30, "li": 31}, spk_is_dialect={"li": "sichuan"})
    cfg = types.SimpleNamespace(talker_config=tc, tts_bos_token_id=3, tts_eos_token_id=4, tts_pad_token_id=2)
    lin = torch.nn.functional.linear
    proj = lambda x: lin(torch.nn.functional.silu(lin(x, w["fc1_w"], w["fc1_b"])), w["fc2_w"], w["fc2_b"])
    books = [(lambda ids, i=i: w["books"][i][ids]) for i in range(15)]
    for name, case in gold["cases"].items():
        kw = case["inputs"]
        got = op.prepare_inputs(cfg, kw["input_ids"], instruct_ids=kw.get("instruct_ids"), ref_ids=kw.get("ref_ids"),
                                voice_clone_prompt=kw.get("voice_clone_prompt"), languages=kw.get("languages"),
                                speakers=kw.get("speakers"), non_streaming_mode=kw.get("non_streaming_mode", False),
                                text_embedding=lambda ids: w["text_embedding"][ids],
                                input_embedding=lambda ids: w["codec_embedding"][ids], text_projection=proj,
                                predictor_embeddings=books)
        for g, key in zip(got, ("prompts", "trailing", "tts_pad", "mask")):
            assert torch.equal(g, case[key]), f"{name}: {key}"
