"""GPU parity of the prompt-prep embedding assembly (nano_qwen3tts_vllm_b200/utils/generation.py, b2tts_prompt_assemble)
against the reference's utils/generation.py:18-379, 412-515 - through the golden vectors written by
oracle/pin_against_reference.py from the LIVE reference functions (tests/golden/prompt.pt), and against the CPU
oracle (oracle/prompt.py) fed with the very same projected rows, where every byte must agree."""
import os
import types

import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    return torch.load(os.path.join(ROOT, "tests", "golden", "prompt.pt"))


def _config():
    tc = types.SimpleNamespace(
        num_code_groups=16, codec_nothink_id=5, codec_think_id=6, codec_think_bos_id=7, codec_think_eos_id=8,
        codec_pad_id=9, codec_bos_id=10, codec_language_id={"english": 20, "chinese": 21, "sichuan": 22},
        spk_id={"anna": 30, "li": 31}, spk_is_dialect={"li": "sichuan"})
    return types.SimpleNamespace(talker_config=tc, tts_bos_token_id=3, tts_eos_token_id=4, tts_pad_token_id=2)


def _modules(w):
    from nano_qwen3tts_vllm_b200.models.qwen3_tts_talker import Qwen3TTSTalkerResizeMLP
    text = torch.nn.Embedding(*w["text_embedding"].shape).to(BF16).cuda()
    codec = torch.nn.Embedding(*w["codec_embedding"].shape).to(BF16).cuda()
    torch.set_default_dtype(BF16)
    try:
        proj = Qwen3TTSTalkerResizeMLP(w["fc1_w"].shape[1], w["fc1_w"].shape[0], w["fc2_w"].shape[0], bias=True).cuda()
    finally:
        torch.set_default_dtype(torch.float32)
    with torch.no_grad():
        text.weight.copy_(w["text_embedding"]); codec.weight.copy_(w["codec_embedding"])
        proj.linear_fc1.weight.copy_(w["fc1_w"]); proj.linear_fc1.bias.copy_(w["fc1_b"])
        proj.linear_fc2.weight.copy_(w["fc2_w"]); proj.linear_fc2.bias.copy_(w["fc2_b"])
    return text, codec, proj, w["books"].cuda().contiguous()


def _to(x, dev):
    if torch.is_tensor(x):
        return x.to(dev)
    if isinstance(x, list):
        return [_to(v, dev) for v in x]
    if isinstance(x, dict):
        return {k: _to(v, dev) for k, v in x.items()}
    return x


def _call(fn, cfg, kw, dev, **mods):
    return fn(cfg, _to(kw["input_ids"], dev), instruct_ids=_to(kw.get("instruct_ids"), dev), ref_ids=_to(kw.get("ref_ids"), dev),
              voice_clone_prompt=_to(kw.get("voice_clone_prompt"), dev), languages=kw.get("languages"),
              speakers=kw.get("speakers"), non_streaming_mode=kw.get("non_streaming_mode", False), **mods)


def _close(got, want, what, frac=0.02):
    got, want = got.float().cpu(), want.float()
    assert got.shape == want.shape, (what, got.shape, want.shape)
    err = (got - want).abs().max().item()
    assert err <= frac * want.abs().max().item() + 1e-6, f"{what}: max abs err {err}"


def test_prompt_assembly_matches_reference_golden(golden):
    """Every mode of the reference against its own outputs (bf16, 2e-2 relative: the projection MLP runs on cuBLAS
    / the native GEMM here and on the CPU there); masks and shapes exactly."""
    from nano_qwen3tts_vllm_b200.utils.generation import generate_speaker_prompt, prepare_inputs
    cfg = _config()
    text, codec, proj, books = _modules(golden["weights"])
    for name, case in golden["cases"].items():
        got = _call(prepare_inputs, cfg, case["inputs"], "cuda", text_embedding=text, input_embedding=codec,
                    text_projection=proj, device="cuda", code_predictor_embeddings=books,
                    generate_speaker_prompt_fn=lambda v: generate_speaker_prompt(v, "cuda"))
        _close(got[0], case["prompts"], f"{name}: prompts")
        _close(got[1], case["trailing"], f"{name}: trailing")
        _close(got[2], case["tts_pad"], f"{name}: tts_pad")
        assert torch.equal(got[3].cpu(), case["mask"]), f"{name}: attention mask"
        # padding rows are exact zeros, as pad_sequence leaves them (generation.py:329-341)
        assert torch.equal(got[0][got[3] == 0], torch.zeros_like(got[0][got[3] == 0]))


def test_prompt_assembly_is_bit_exact_given_the_same_projected_rows(golden):
    """The oracle (pinned bit-exact against the reference) and the product fed with ONE table of projected rows:
    gathers, bf16 adds, ICL frame sums, left padding, pad-row padding - every byte equal."""
    from oracle import prompt as op
    from nano_qwen3tts_vllm_b200.utils.generation import generate_icl_prompt, generate_speaker_prompt, prepare_inputs
    cfg = _config()
    text, codec, proj, books = _modules(golden["weights"])
    with torch.no_grad():
        table = proj(text.weight.data)                                # [vocab, H]: one projected row per text token
    lookup_gpu = lambda rows: rows                                     # "projection" of a row that is already projected
    emb_gpu = types.SimpleNamespace(weight=table.contiguous())          # text_embedding := the projected table
    table_cpu = table.cpu()
    emb_cpu = lambda ids: table_cpu[ids]
    codec_cpu = lambda ids: golden["weights"]["codec_embedding"][ids]
    books_cpu = [(lambda ids, i=i: golden["weights"]["books"][i][ids]) for i in range(15)]
    for name, case in golden["cases"].items():
        want = _call(op.prepare_inputs, cfg, case["inputs"], "cpu", text_embedding=emb_cpu, input_embedding=codec_cpu,
                     text_projection=lambda x: x, predictor_embeddings=books_cpu)
        got = _call(prepare_inputs, cfg, case["inputs"], "cuda", text_embedding=emb_gpu, input_embedding=codec,
                    text_projection=lookup_gpu, device="cuda", code_predictor_embeddings=books,
                    generate_speaker_prompt_fn=lambda v: generate_speaker_prompt(v, "cuda"))
        for g, w, what in zip(got, want, ("prompts", "trailing", "tts_pad", "mask")):
            assert g.shape == w.shape, (name, what, g.shape, w.shape)
            if "icl" in name and what in ("prompts", "trailing"):
                # 16-codebook sums: fp32 accumulation in both, but not the same order -> at most one bf16 ulp apart
                d = (g.float().cpu() - w.float()).abs()
                assert (d <= w.float().abs() * 2 ** -7 + 1e-8).all(), f"{name}: {what} beyond one ulp"
                assert (d == 0).float().mean().item() > 0.98, f"{name}: {what}"
            else:
                assert torch.equal(g.cpu(), w), f"{name}: {what} differs"
    # the standalone ICL entry point equals what prepare_inputs plans for the same request
    case = golden["cases"]["icl_text_longer"]["inputs"]
    ids, ref = case["input_ids"][0].cuda(), case["ref_ids"][0].cuda()
    full = _call(prepare_inputs, cfg, case, "cuda", text_embedding=emb_gpu, input_embedding=codec, text_projection=lookup_gpu,
                 device="cuda", code_predictor_embeddings=books,
                 generate_speaker_prompt_fn=lambda v: generate_speaker_prompt(v, "cuda"))
    rows, trail = generate_icl_prompt(ids[:, 3:-5], ref[:, 3:-2], case["voice_clone_prompt"]["ref_code"][0].cuda(),
                                      full[2], table[cfg.tts_eos_token_id].view(1, 1, -1), False, cfg, emb_gpu, codec, lookup_gpu,
                                      books, "cuda")
    assert torch.equal(full[0][0, -rows.shape[1]:], rows[0]) and torch.equal(full[1][0, :trail.shape[1]], trail[0])
    # and the reference's hook protocol (generate_icl_prompt_fn, no codebook tables handed to prepare_inputs)
    hook = lambda text_id, ref_id, ref_code, tts_pad_embed, tts_eos_embed, non_streaming_mode: generate_icl_prompt(
        text_id, ref_id, ref_code, tts_pad_embed, tts_eos_embed, non_streaming_mode, cfg, emb_gpu, codec, lookup_gpu, books, "cuda")
    hooked = _call(prepare_inputs, cfg, case, "cuda", text_embedding=emb_gpu, input_embedding=codec, text_projection=lookup_gpu,
                   device="cuda", generate_icl_prompt_fn=hook,
                   generate_speaker_prompt_fn=lambda v: generate_speaker_prompt(v, "cuda"))
    for g, w in zip(hooked, full):
        assert torch.equal(g, w)


def test_prompt_assembly_at_model_width_and_4k_icl():
    """1.7B widths (text 2048 -> hidden 2048, native GEMM projection) and a 4096-frame voice-clone reference
    (BASELINE configs[3]): shapes, mask, and the projection within 2e-2 of an fp32 evaluation."""
    from nano_qwen3tts_vllm_b200.models.qwen3_tts_talker import Qwen3TTSTalkerResizeMLP
    from nano_qwen3tts_vllm_b200.utils.generation import prepare_inputs
    cfg = _config()
    g = torch.Generator(device="cuda").manual_seed(5)
    rnd = lambda *s, std=0.02: (torch.randn(*s, device="cuda", generator=g) * std).to(BF16)
    text = types.SimpleNamespace(weight=rnd(1000, 2048, std=0.05))
    codec = types.SimpleNamespace(weight=rnd(3072, 2048, std=0.05))
    books = rnd(15, 2048, 2048, std=0.05)
    torch.set_default_dtype(BF16)
    try:
        proj = Qwen3TTSTalkerResizeMLP(2048, 2048, 2048, bias=True).cuda()
    finally:
        torch.set_default_dtype(torch.float32)
    with torch.no_grad():
        for p in proj.parameters():
            p.copy_(rnd(*p.shape))
    ids = torch.randint(11, 990, (1, 60), device="cuda", generator=g)
    ref = torch.randint(11, 990, (1, 40), device="cuda", generator=g)
    code = torch.randint(0, 2048, (4096, 16), device="cuda", generator=g)
    clone = {"ref_code": [code], "ref_spk_embedding": [rnd(2048)], "x_vector_only_mode": [False], "icl_mode": [True]}
    prompts, trailing, pad, mask = prepare_inputs(cfg, [ids], ref_ids=[ref], voice_clone_prompt=clone, languages=["english"],
                                                  text_embedding=text, input_embedding=codec, text_projection=proj,
                                                  device="cuda", voice_clone_spk_embeds=clone["ref_spk_embedding"],
                                                  code_predictor_embeddings=books)
    n_tags = 4 + 1 + 2
    assert prompts.shape == (1, 3 + n_tags - 1 + 4097, 2048) and mask.all() and trailing.shape == (1, 1, 2048)
    x = text.weight[[cfg.tts_pad_token_id]].float()
    want = torch.nn.functional.linear(torch.nn.functional.silu(torch.nn.functional.linear(
        x, proj.linear_fc1.weight.float(), proj.linear_fc1.bias.float())), proj.linear_fc2.weight.float(), proj.linear_fc2.bias.float())
    _close(pad.view(1, -1), want.cpu(), "tts_pad at model width")
    # a middle ICL row: tts pad (text exhausted) + the 16-codebook sum of frame 1000
    t = 3 + n_tags - 1 + 1 + 1000
    frame = codec.weight[code[1000, 0]].float() + sum(books[i - 1][code[1000, i]].float() for i in range(1, 16))
    _close(prompts[0, t].view(1, -1), (pad.view(-1).float() + frame.to(BF16).float()).view(1, -1).cpu(), "ICL row 1000")
