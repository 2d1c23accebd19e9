"""GPU parity of the device-resident frame loop: greedy (top_k = 1) codec
frames of DeviceFrameLoop / StepApiFrameLoop against the CPU oracle's run_frames on the
same random-init weights and prompts (north_star: greedy codec-token sequences identical).

A greedy sequence can only be required to be identical while the oracle's own arg-max is decided by more
than the arithmetic noise between two correct bf16 pipelines: where the oracle's top-2 logit margin is below
one bf16 ulp of the top logit the choice is a coin flip of GEMM accumulation order (fp32 sums are formed in a
different order on the tensor cores than in the CPU's blocked GEMM).  The test therefore demands identity up to the
first frame whose oracle margin is under that ulp, and demands that any divergence sits exactly on such a
near-tie."""
import pytest
import torch

from oracle import model as OM
from oracle import ops as O
from tests.test_gpu_models import small_model_config, small_weights

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
NFRAMES = 24
B = 4


def _prompts(td, b):
    g = torch.Generator().manual_seed(1000)
    prompts = [(torch.randn(n, td.hidden_size, generator=g) * 0.05).to(BF16) for n in ([9, 14, 5, 11] * 4)[:b]]
    pad = (torch.randn(1, td.hidden_size, generator=torch.Generator().manual_seed(999)) * 0.05).to(BF16)
    return prompts, pad


@pytest.fixture(scope="module")
def setup():
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    td, pd, tw, pw = small_weights()
    prompts, pad = _prompts(td, B)
    margins = []

    def greedy(logits, temps):
        top2 = logits.float().topk(2, dim=-1).values
        margins.append((top2[:, 0] - top2[:, 1], top2[:, 0]))
        return O.topk_sample(logits, temps, torch.ones(logits.shape), top_k=1)

    ot = OM.Talker(td, tw, B * 2)
    op = OM.Predictor(pd, pw, B, td.hidden_size)
    want = torch.tensor(OM.run_frames(ot, op, prompts, NFRAMES, pad, greedy).frames)      # [frames, B, 16]
    mc = small_model_config()
    kw = dict(num_kvcache_blocks=4 * B, max_num_seqs=max(B, 2), max_model_len=512, enforce_eager=True)
    talker = TalkerLLM(mc, state_dict=tw, **kw)
    predictor = PredictorLLM(mc, state_dict=pw, **kw)
    yield dict(td=td, prompts=prompts, pad=pad, want=want, margins=margins, talker=talker, predictor=predictor)
    talker.exit()
    predictor.exit()


def _check_against_oracle(got, want, margins):
    """Identity per stream up to its first near-tie; every divergence must sit on a near-tie."""
    assert got.shape == want.shape
    nframes, b, ncb = want.shape
    exact = 0
    for s in range(b):
        ne = (got[:, s] != want[:, s]).nonzero()
        if ne.numel() == 0:
            exact += 1
            continue
        f, c = ne[0].tolist()
        m, top = margins[ncb * f + c]
        ulp = 2.0 ** (torch.tensor(float(top[s])).abs().log2().floor().item() - 7)
        assert float(m[s]) <= 2 * ulp, (f"stream {s} diverges at frame {f} codebook {c} where the oracle's top-2 margin "
                                        f"{float(m[s]):.4g} is well above a bf16 ulp ({ulp:.4g})")
    return exact


def test_device_frame_loop_greedy_vs_oracle(setup):
    from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
    t, p = setup["talker"].model_runner, setup["predictor"].model_runner
    t.kv_cache.zero_(); p.kv_cache.zero_()
    loop = DeviceFrameLoop(t, p, B, NFRAMES, 16, top_k=1, scatter_pages=True)
    loop.unit_noise = True
    loop.start([x.cuda() for x in setup["prompts"]], setup["pad"].cuda())
    loop.capture()
    loop.run(NFRAMES)
    got = loop.result()
    _check_against_oracle(got, setup["want"], setup["margins"])


def test_graph_replay_equals_eager_frames(setup):
    """The captured frame graph and the same frame launched eagerly produce the same codes."""
    from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
    t, p = setup["talker"].model_runner, setup["predictor"].model_runner
    outs = []
    for capture in (False, True):
        t.kv_cache.zero_(); p.kv_cache.zero_()
        loop = DeviceFrameLoop(t, p, B, 6, 16, top_k=1, scatter_pages=False)
        loop.unit_noise = True       # exact bf16 logit ties are broken by the sampler's noise otherwise
        loop.start([x.cuda() for x in setup["prompts"]], setup["pad"].cuda())
        if capture:
            loop.capture()
        loop.run(6)
        outs.append(loop.result())
    assert torch.equal(outs[0], outs[1])


def test_step_api_frame_loop_greedy_vs_oracle(setup):
    """The reference's per-frame protocol with host buffers (StepApiFrameLoop) gives the oracle's frames."""
    from nano_qwen3tts_vllm_b200.engine.frame_loop import StepApiFrameLoop
    talker, predictor = setup["talker"], setup["predictor"]
    talker.model_runner.kv_cache.zero_(); predictor.model_runner.kv_cache.zero_()
    old = talker.model_runner.top_k, predictor.model_runner.top_k
    talker.model_runner.top_k = predictor.model_runner.top_k = 1
    try:
        api = StepApiFrameLoop(talker, predictor, setup["pad"])
        api.add_streams([x.unsqueeze(0).pin_memory() for x in setup["prompts"]])
        for _ in range(NFRAMES):
            api.step_frame()
        got = torch.tensor([api.frames[f"s{i}"] for i in range(B)]).permute(1, 0, 2)
        api.close()
    finally:
        talker.model_runner.top_k, predictor.model_runner.top_k = old
    _check_against_oracle(got, setup["want"], setup["margins"])
