"""The decoder layer has two ways to apply its RMSNorms: inside the consuming GEMM (9 ... 64 rows by default) or as a
separate rmsnorm launch in front of a plain GEMM (models/qwen3_tts_share.py FUSED_NORM_MIN_ROWS / FUSED_NORM_MAX_ROWS).
The toy-size golden vectors of the live reference (layers/layernorm.py:18-41, models/qwen3_tts_share.py:163-176) must hold
on BOTH, whatever the default range says about the batch sizes the other tests happen to use."""
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("min_rows, max_rows", [(1, 64), (10 ** 6, 64)], ids=["norms-in-gemm", "norms-separate"])
def test_model_goldens_on_both_norm_paths(min_rows, max_rows, golden_models, monkeypatch):
    import nano_qwen3tts_vllm_b200.models.qwen3_tts_share as share
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    from tests import test_gpu_models as M
    monkeypatch.setattr(share, "FUSED_NORM_MIN_ROWS", min_rows)
    monkeypatch.setattr(share, "FUSED_NORM_MAX_ROWS", max_rows)
    td, pd, tw, pw = M.small_weights()
    mc = M.small_model_config()
    kw = dict(num_kvcache_blocks=8, max_num_seqs=16, max_model_len=1024, graph_batch_sizes=[1, 2, 4, 8])
    talker = TalkerLLM(mc, state_dict=tw, **kw)           # graphs are captured under the patched range
    predictor = PredictorLLM(mc, state_dict=pw, **kw)
    try:
        engines = (talker, predictor, (td, pd, tw, pw))
        M.test_talker_runner_vs_live_reference(engines, golden_models)
        M.test_predictor_runner_vs_live_reference(engines, golden_models)
    finally:
        talker.exit()
        predictor.exit()


def test_wide_batch_frames_teacher_forced_vs_oracle():
    """80 streams at toy size: the talker's decode step (80 rows: chunks of 64 + 16), the predictor's prefill (160 rows,
    3 chunks) and its decode passes all take the chunked plain-x GEMMs with separate, programmatically launched
    RMSNorms.  Every decision of 8 frames is compared on the oracle's own history (interface.py:1207-1290)."""
    import torch
    from oracle import model as OM
    from oracle import ops as O
    from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    from tests import test_gpu_frame_loop as F
    from tests import test_gpu_models as M
    b, nframes = 80, 8
    td, pd, tw, pw = M.small_weights()
    g = torch.Generator().manual_seed(77)
    prompts = [(torch.randn(5 + i % 9, td.hidden_size, generator=g) * 0.05).to(torch.bfloat16) for i in range(b)]
    pad = (torch.randn(1, td.hidden_size, generator=torch.Generator().manual_seed(999)) * 0.05).to(torch.bfloat16)
    margins = []

    def greedy(logits, temps):
        top2 = logits.float().topk(2, dim=-1).values
        margins.append((top2[:, 0] - top2[:, 1], top2[:, 0]))
        return O.topk_sample(logits, temps, torch.ones(logits.shape), top_k=1)

    want = torch.tensor(OM.run_frames(OM.Talker(td, tw, b * 2), OM.Predictor(pd, pw, b, td.hidden_size), prompts, nframes,
                                      pad, greedy).frames)
    kw = dict(num_kvcache_blocks=2 * b, max_num_seqs=b, max_model_len=256, enforce_eager=True)
    talker = TalkerLLM(M.small_model_config(), state_dict=tw, **kw)
    predictor = PredictorLLM(M.small_model_config(), state_dict=pw, **kw)
    try:
        loop = DeviceFrameLoop(talker.model_runner, predictor.model_runner, b, nframes, 16, top_k=1)
        loop.unit_noise = True
        loop.force(want)
        loop.start([x.cuda() for x in prompts], pad.cuda())
        loop.capture()
        loop.run(nframes)
        assert torch.equal(loop.result(), want)
        F._check_teacher_forced(loop.sampled[:nframes].permute(0, 2, 1).cpu(), want, margins)
    finally:
        talker.exit()
        predictor.exit()
