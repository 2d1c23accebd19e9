"""f4 (what can be built of it): the graph-per-length wrapper and the streaming window logic of the reference's
codec decode path (utils/speech_tokenizer_cudagraph.py:18-168, examples/server.py:196-222, 338-349) around a STAND-IN
decoder - the real vocoder (qwen-tts 0.0.5) is not available."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


class ToyDecoder(torch.nn.Module):
    """codes [B, 16, T] -> wav [B, 1, T * 8]: 16-codebook embedding sum, two transposed convolutions, tanh."""
    total_upsample = 8

    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.emb = torch.nn.Embedding(16 * 100, 32)
        self.up1 = torch.nn.ConvTranspose1d(32, 16, 4, stride=2, padding=1)
        self.up2 = torch.nn.ConvTranspose1d(16, 1, 8, stride=4, padding=2)

    def forward(self, codes):
        off = torch.arange(16, device=codes.device).view(1, 16, 1) * 100
        x = self.emb(codes + off).sum(1).transpose(1, 2)                 # [B, 32, T]
        return torch.tanh(self.up2(torch.relu(self.up1(x))))

    def chunked_decode(self, codes, chunk_size, left_context_size):
        outs, t = [], codes.shape[2]
        for s in range(0, t, chunk_size):
            lo = max(0, s - left_context_size)
            w = self.forward(codes[:, :, lo:min(t, s + chunk_size)])
            outs.append(w[..., (s - lo) * self.total_upsample:])
        return torch.cat(outs, -1)


def _tokenizer():
    dec = ToyDecoder().cuda().eval()
    model = types.SimpleNamespace(decoder=dec, get_output_sample_rate=lambda: 24000)

    def decode(inputs):
        with torch.inference_mode():
            return [dec(torch.as_tensor(r["audio_codes"]).cuda().long().T.unsqueeze(0)).reshape(-1).float().cpu().numpy()
                    for r in inputs], 24000
    return types.SimpleNamespace(model=model, decode=decode, config=types.SimpleNamespace(sample_rate=24000))


def test_decoder_graphs_replay_equals_eager_and_fall_back_beyond_50_frames():
    from nano_qwen3tts_vllm_b200.utils.speech_tokenizer_cudagraph import SpeechTokenizerCUDAGraph
    tok = _tokenizer()
    ref = ToyDecoder().cuda().eval()                                     # same seed -> same weights, never patched
    st = SpeechTokenizerCUDAGraph(tokenizer=tok, device="cuda:0", num_graph_lengths=50)
    dec = tok.model.decoder
    assert sorted(dec._b2_graphs.graphs) == list(range(1, 51)) and st.sample_rate == 24000 and st.samples_per_frame == 8
    g = torch.Generator(device="cuda").manual_seed(1)
    with torch.inference_mode():
        for t in (1, 2, 17, 50, 51, 120):
            codes = torch.randint(0, 100, (1, 16, t), device="cuda", generator=g)
            want = ref(codes)
            got = dec(codes)                                             # patched forward: graph replay for t <= 50
            assert got.shape == (1, 1, 8 * t) and torch.equal(got, want), t
            assert torch.equal(dec(codes), want)                         # replay again: static buffers re-filled
    wavs, sr = st.decode_codec_ids(torch.randint(0, 100, (2, 16, 9), device="cuda", generator=g))
    assert sr == 24000 and len(wavs) == 2 and wavs[0].shape == (72,)


def test_decode_window_trims_the_left_context_and_emits_pcm16():
    from nano_qwen3tts_vllm_b200.utils.speech_tokenizer_cudagraph import SpeechTokenizerCUDAGraph, float_to_pcm16
    tok = _tokenizer()
    st = SpeechTokenizerCUDAGraph(tokenizer=tok, device="cuda:0", num_graph_lengths=50)
    g = torch.Generator().manual_seed(2)
    frames = torch.randint(0, 100, (40, 16), generator=g)
    batch = [{"audio_codes": frames[:30].tolist(), "left_context_frames": 25},
             {"audio_codes": frames[5:40].tolist(), "left_context_frames": None},
             {"audio_codes": frames[:3].tolist(), "left_context_frames": 7}]           # context longer than the window
    wavs, sr = st.decode_window_batched(batch)
    full, _ = tok.decode([{"audio_codes": b["audio_codes"]} for b in batch])          # server.py:201-222 fallback
    assert sr == 24000
    for w, f, b in zip(wavs, full, batch):
        skip = (b["left_context_frames"] or 0) * 8
        want = f[skip:] if skip < len(f) else np.array([], dtype=np.float32)
        assert w.dtype == np.float32 and np.array_equal(w, want)
    pcm, _ = st.decode_window_batched(batch[:2], pcm16=True, volume=0.7)
    for p, f, b in zip(pcm, full, batch):
        x = np.clip(f[(b["left_context_frames"] or 0) * 8:], -1.0, 1.0)
        want = (x * 32767).astype(np.int16)                                           # server.py:338-341
        want = np.clip((want.astype(np.float32) * 0.7).round(), -32768, 32767).astype(np.int16)    # :344-349
        assert p.dtype == np.int16 and np.array_equal(p, want)
    # the conversion alone, full range, ties and out-of-range samples
    x = torch.tensor([0.0, 1.0, -1.0, 1.5, -2.0, 0.5, -0.5, 1e-5, 0.999999, 1.5258789e-05 * 3], device="cuda")
    want = (np.clip(x.cpu().numpy(), -1, 1) * 32767).astype(np.int16)
    assert np.array_equal(float_to_pcm16(x).cpu().numpy(), want)
    big = (torch.rand(100003, device="cuda") * 2.4 - 1.2)
    want = (np.clip(big.cpu().numpy(), -1, 1) * 32767).astype(np.int16)
    assert np.array_equal(float_to_pcm16(big).cpu().numpy(), want)


def test_missing_vocoder_raises_like_the_reference():
    from nano_qwen3tts_vllm_b200.utils.speech_tokenizer_cudagraph import SpeechTokenizerCUDAGraph
    with pytest.raises(ImportError, match="qwen_tts not found"):
        SpeechTokenizerCUDAGraph("/nonexistent")
