"""GPU parity of the model runners / step API against the committed outputs of the LIVE reference
(tests/golden/models.pt, written by oracle/pin_against_reference.py) and against the CPU oracle.

Tolerance (BASELINE.json north_star): logits / hidden within 2e-2 relative (max|a-b| / max|b|);
KV slot contents compared at the same tolerance here because K depends on GEMM accumulation
order upstream of the write (the write itself is bit-exact, tests/test_gpu_ops.py)."""
import pytest
import torch

from oracle import model as OM

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
TOL = 2e-2


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def small_model_config():
    from nano_qwen3tts_vllm_b200.config import ModelConfig, StackConfig
    mrope = {"mrope_section": [24, 20, 20], "interleaved": True}
    t = StackConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, vocab_size=3072, rope_scaling=mrope,
                    text_vocab_size=64, text_hidden_size=96)
    p = StackConfig(hidden_size=128, intermediate_size=384, num_hidden_layers=2, vocab_size=2048)
    return ModelConfig(t, p)


def small_weights(seed=1234):
    td = OM.StackDims(hidden_size=256, intermediate_size=512, num_hidden_layers=2, vocab_size=3072,
                      rope_theta=1e6, mrope_section=[24, 20, 20], mrope_interleaved=True, page_size=256)
    pd = OM.StackDims(hidden_size=128, intermediate_size=384, num_hidden_layers=2, vocab_size=2048,
                      rope_theta=1e6, page_size=256)
    tw, pw = OM.synthetic_weights(td, pd, seed=seed, text_vocab=64, text_hidden=96)
    return td, pd, tw, pw


@pytest.fixture(scope="module")
def engines():
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    td, pd, tw, pw = small_weights()
    mc = small_model_config()
    kw = dict(num_kvcache_blocks=8, max_num_seqs=16, max_model_len=1024, graph_batch_sizes=[1, 2, 4, 8])
    talker = TalkerLLM(mc, state_dict=tw, **kw)
    predictor = PredictorLLM(mc, state_dict=pw, **kw)
    yield talker, predictor, (td, pd, tw, pw)
    talker.exit()
    predictor.exit()


def test_talker_runner_vs_live_reference(engines, golden_models):
    """Prefill of two ragged prompts, then three decode steps, through ModelRunner.run_model."""
    from nano_qwen3tts_vllm_b200.engine.sequence import Sequence
    from nano_qwen3tts_vllm_b200.utils.context import reset_context
    talker, _, _ = engines
    r = talker.model_runner
    gp = golden_models["talker_prefill"]
    seqs = []
    for prompt, table in zip(gp["prompts"], gp["tables"]):
        s = Sequence([], input_embeds=prompt.unsqueeze(0).cuda())
        s.block_table = list(table)
        seqs.append(s)
    logits, hidden = r.run_model(seqs, True)
    reset_context()
    assert rel(logits, gp["logits"]) <= TOL and rel(hidden, gp["hidden"]) <= TOL
    for step in golden_models["talker_decode"]:
        for i, s in enumerate(seqs):
            s.num_tokens = step["seq_lens"][i]
            s.decode_input_embeds = step["embeds"][i:i + 1].cuda()
        logits, hidden = r.run_model(seqs, False)          # CUDA-graph path (bs 2)
        reset_context()
        assert rel(logits, step["logits"]) <= TOL, rel(logits, step["logits"])
        assert rel(hidden, step["hidden"]) <= TOL
    pages = r.kv_cache[:, :, [5, 2], :16].cpu()
    assert rel(pages, golden_models["talker_kv_pages"]) <= TOL
    # untouched slots of the pool stay zero
    assert float(r.kv_cache[:, :, [0, 1, 3, 4, 6, 7]].abs().max()) == 0.0


def test_predictor_runner_vs_live_reference(engines, golden_models):
    from nano_qwen3tts_vllm_b200.engine.sequence import PredictorSequence
    from nano_qwen3tts_vllm_b200.utils.context import reset_context
    _, predictor, _ = engines
    r = predictor.model_runner
    gp = golden_models["predictor_prefill"]
    seqs = []
    for x, table in zip(gp["inputs"], gp["tables"]):
        s = PredictorSequence([], input_embeds=x.unsqueeze(0).cuda())
        s.block_table = list(table)
        seqs.append(s)
    logits = r.run_model(seqs, True)
    reset_context()
    assert rel(logits, gp["logits"]) <= TOL
    for step in golden_models["predictor_decode"]:
        for i, s in enumerate(seqs):
            s.num_tokens = step["seq_lens"][i]
            s.last_token = step["tokens"][i]
            s.generation_steps = step["gen_steps"][i]
        logits = r.run_model(seqs, False)
        reset_context()
        assert rel(logits, step["logits"]) <= TOL, (step["gen_steps"], rel(logits, step["logits"]))


def test_predictor_mixed_generation_steps_eager_equals_graph(engines):
    """Rows with different generation steps in one batch (staggered arrivals): per-row embedding
    table and per-group head; eager and graph paths agree bit for bit."""
    from nano_qwen3tts_vllm_b200.engine.sequence import PredictorSequence
    from nano_qwen3tts_vllm_b200.utils.context import reset_context
    _, predictor, (td, pd, tw, pw) = engines
    r = predictor.model_runner
    g = torch.Generator().manual_seed(5)
    seqs = []
    for i, (gs, ln) in enumerate([(1, 3), (7, 9), (14, 16)]):
        s = PredictorSequence([], input_embeds=torch.zeros(1, 2, td.hidden_size, dtype=BF16).cuda())
        s.block_table = [i + 3]
        s.num_tokens, s.generation_steps = ln, gs
        s.last_token = int(torch.randint(0, 2048, (1,), generator=g))
        seqs.append(s)
    r.kv_cache.normal_(generator=None)
    kv = r.kv_cache.clone()
    a = r.run_model(seqs, False).clone()
    reset_context()
    r.kv_cache.copy_(kv)
    r.enforce_eager = True
    try:
        b = r.run_model(seqs, False).clone()
    finally:
        r.enforce_eager = False
        reset_context()
    assert torch.equal(a, b)
    # oracle on the same cache contents
    op = OM.Predictor(pd, pw, 8, td.hidden_size)
    op.stack.kv_cache.copy_(kv.cpu())
    want = op.decode([s.last_token for s in seqs], [s.generation_steps for s in seqs], [len(s) for s in seqs],
                     [s.block_table for s in seqs])
    assert rel(a, want) <= TOL
    r.kv_cache.zero_()


def test_step_api_frame_protocol(engines):
    """TalkerLLM / PredictorLLM driven exactly as the reference's frame loop drives them
    (interface.py:1207-1290): add_request -> step_with_outputs -> predictor add_request -> 15 steps."""
    from nano_qwen3tts_vllm_b200.sampling_params import SamplingParams
    talker, predictor, (td, pd, tw, pw) = engines
    g = torch.Generator().manual_seed(9)
    prompts = [(torch.randn(1, n, td.hidden_size, generator=g) * 0.05).to(BF16) for n in (7, 12)]
    for i, p in enumerate(prompts):
        talker.add_request([p], SamplingParams(temperature=1.0, max_tokens=64, ignore_eos=True), request_id=f"r{i}")
    _, ntok, outs = talker.step_with_outputs()
    assert ntok == 21 and len(outs) == 2      # counted after the append, like the reference
    emb0 = tw["model.codec_embedding.weight"].cuda()
    for frame in range(2):
        for rid, _, toks, hidden, fin in outs:
            assert hidden.shape == (td.hidden_size,) and not fin and len(toks) == frame + 1
            x = torch.stack([hidden, emb0[toks[-1]]], 0).unsqueeze(0)            # [1, 2, H]
            predictor.add_request([x.cpu()], SamplingParams(temperature=0.9, max_tokens=17), request_id=rid)
        done = {}
        for _ in range(15):
            fin_out, _ = predictor.step()
            for rid, _, codes in fin_out:
                done[rid] = codes
        assert sorted(done) == ["r0", "r1"] and all(len(c) == 15 for c in done.values())
        assert predictor.is_finished()
        for rid in done:
            talker.add_request([torch.zeros(1, 1, td.hidden_size, dtype=BF16)], SamplingParams(), request_id=rid)
        _, ntok, outs = talker.step_with_outputs()
        assert ntok == -2
    for rid in ("r0", "r1"):
        talker.clear_request(rid)
    assert talker.is_finished()
    assert len(talker.scheduler.block_manager.free_block_ids) == 8


def test_predictor_deferred_tokens_equal_per_step_readback(engines):
    """``PredictorLLMEngine.defer_tokens`` (codes stay on the device until a sequence finishes: one read per burst)
    returns exactly the codes of the reference's per-step ``.tolist()`` protocol, including staggered arrivals
    (a request added in the middle of another one's burst: mixed generation steps, fallback to host tokens)."""
    from nano_qwen3tts_vllm_b200.sampling_params import SamplingParams
    _, predictor, (td, pd, tw, pw) = engines
    g = torch.Generator().manual_seed(77)
    inputs = [(torch.randn(1, 2, td.hidden_size, generator=g) * 0.05).to(BF16).pin_memory() for _ in range(5)]
    sp = SamplingParams(temperature=0.9, max_tokens=17)      # 2 prompt rows count as completion tokens (interface.py:1196)

    def burst(defer: bool):
        predictor.defer_tokens = defer
        predictor.model_runner.kv_cache.zero_()
        torch.manual_seed(1234)                    # same sampler noise in both runs
        torch.cuda.manual_seed(1234)
        done, steps = {}, 0
        for i in range(3):
            predictor.add_request([inputs[i]], sp, request_id=f"r{i}")
        while len(done) < 5:
            if steps == 4:                         # two late arrivals while r0..r2 are mid-burst
                for i in (3, 4):
                    predictor.add_request([inputs[i]], sp, request_id=f"r{i}")
            fin, _ = predictor.step()
            steps += 1
            for rid, _, codes in fin:
                done[rid] = list(codes)
            assert steps < 200
        return done

    try:
        want = burst(False)
        got = burst(True)
    finally:
        predictor.defer_tokens = True
    assert set(got) == set(want) == {f"r{i}" for i in range(5)}
    for rid in want:
        assert len(want[rid]) == 15 and all(0 <= c < pd.vocab_size for c in got[rid]), rid
        assert got[rid] == want[rid], rid


def test_talker_prefill_graph_equals_eager_prefill(engines):
    """Single-sequence prompts of <= 256 tokens run through a CUDA graph per length (captured on first use,
    reference talker_mode_runner.py:194-234 captures all 256 at start-up): same sampled token, same hidden state and
    the same KV pages as the eager prefill, on capture AND on replay."""
    from nano_qwen3tts_vllm_b200.engine.sequence import Sequence
    from nano_qwen3tts_vllm_b200.sampling_params import SamplingParams
    talker, _, (td, pd, tw, pw) = engines
    r = talker.model_runner
    r.top_k = 1
    g = torch.Generator().manual_seed(21)
    try:
        for n in (5, 64, 5, 200, 64):
            prompt = (torch.randn(1, n, td.hidden_size, generator=g) * 0.05).to(BF16)
            out = {}
            for mode in ("eager", "graph"):
                r.enforce_eager = mode == "eager"
                r.kv_cache.zero_()
                seq = Sequence([], SamplingParams(temperature=1.0, max_tokens=4, ignore_eos=True), input_embeds=prompt)
                seq.block_table = [3]
                assert r._prefill_graph_ok([seq]) == (mode == "graph")
                toks, hidden = r.run([seq], True)
                torch.cuda.synchronize()
                out[mode] = (toks, hidden.clone(), r.kv_cache.clone())
            assert out["eager"][0] == out["graph"][0]
            assert torch.equal(out["eager"][1], out["graph"][1]), f"hidden differs at T={n}"
            assert torch.equal(out["eager"][2], out["graph"][2]), f"KV pages differ at T={n}"
        assert sorted(r._pf_static["graphs"]) == [5, 64, 200]
    finally:
        r.enforce_eager = False
        r.top_k = 50
