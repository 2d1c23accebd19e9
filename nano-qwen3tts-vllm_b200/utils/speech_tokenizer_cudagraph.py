"""Codec decode windows: the graph-per-length wrapper the reference puts around the
speech tokenizer's decoder (utils/speech_tokenizer_cudagraph.py:18-56, 59-168) and the streaming window logic of its
server (examples/server.py:196-222, 338-349) - WITHOUT the vocoder itself.

The vocoder (``qwen_tts.inference.qwen3_tts_tokenizer.Qwen3TTSTokenizer``, qwen-tts 0.0.5) is a third-party model that
is neither in the reference tree nor installed here, so its conv / transformer stack cannot be restated; what the
reference adds on top of it can, and is, for ANY decoder module with the same call surface
(``decoder(codes [1, 16, T] int64) -> wav [1, 1, T * total_upsample]``):

* ``DecoderGraphs`` - one CUDA graph per window length T = 1 .. 50 (captured largest first into one pool, as the
  reference does), replay on a matching length, eager otherwise;
* ``SpeechTokenizerCUDAGraph`` - the reference class: ``decode``, ``chunked_decode``, ``decode_codec_ids``, plus
  ``decode_window_batched`` (the optional fast path its server looks for): decode (left context + new frames), trim the
  context's samples, and leave the GPU as PCM16 (``b2tts_float_to_pcm16``) instead of fp32.
"""
from __future__ import annotations

from typing import List, Tuple

import torch

from .. import _native as nat


class DecoderGraphs:
    """CUDA graphs of ``decoder(codes)`` for codes ``[1, groups, T]``, T in ``lengths``."""

    def __init__(self, decoder, device, lengths: List[int], num_code_groups: int = 16, code_high: int = 100):
        self.decoder, self.device = decoder, torch.device(device)
        self.graphs, self.inputs, self.outputs = {}, {}, {}
        self.eager = decoder.forward
        pool = None
        with torch.inference_mode():
            self.eager(torch.randint(0, code_high, (1, num_code_groups, max(lengths + [1])), device=self.device))   # warm-up
            torch.cuda.synchronize(self.device)
            for t in sorted(lengths, reverse=True):                    # largest first: one pool serves all
                buf = torch.randint(0, code_high, (1, num_code_groups, t), device=self.device, dtype=torch.long)
                self.eager(buf)
                torch.cuda.synchronize(self.device)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, pool):
                    out = self.eager(buf)
                pool = pool or g.pool()
                self.graphs[t], self.inputs[t], self.outputs[t] = g, buf, out
            torch.cuda.synchronize(self.device)

    def __call__(self, codes: torch.Tensor) -> torch.Tensor:
        t = codes.shape[2]
        g = self.graphs.get(t) if codes.shape[0] == 1 else None
        if g is None:
            return self.eager(codes)
        self.inputs[t].copy_(codes)
        g.replay()
        return self.outputs[t]

    def patch(self):
        """Route ``decoder.forward`` through the graphs in place (what the reference's capture function does)."""
        self.decoder.original_forward = self.eager
        self.decoder.forward = self.__call__
        return self


def _capture_decoder_cudagraphs(decoder, device: str, graph_lengths: List[int]):
    """Reference name and behaviour (speech_tokenizer_cudagraph.py:18-56): patches ``decoder.forward`` in place."""
    decoder.eval()
    decoder._b2_graphs = DecoderGraphs(decoder, device, list(graph_lengths)).patch()


class SpeechTokenizerCUDAGraph:
    """The reference's class (speech_tokenizer_cudagraph.py:59-168).  ``tokenizer`` is a loaded speech tokenizer
    (``.model.decoder``, ``.decode``, ``.model.get_output_sample_rate``); without one the constructor tries to load
    qwen_tts' from ``model_path`` exactly as the reference does and raises its ImportError when qwen_tts is absent."""

    def __init__(self, model_path: str | None = None, device: str | None = None, dtype: torch.dtype = torch.bfloat16,
                 num_graph_lengths: int = 50, tokenizer=None):
        device = device or "cuda:0"
        if tokenizer is None:
            try:
                from qwen_tts.inference.qwen3_tts_tokenizer import Qwen3TTSTokenizer
            except ImportError as e:
                raise ImportError("qwen_tts not found. Install Qwen3-TTS and set QWEN_TTS_PATH or add to path.") from e
            tokenizer = Qwen3TTSTokenizer.from_pretrained(model_path, device_map=device)
            tokenizer.model = tokenizer.model.to(dtype)
        self.tokenizer, self.device, self.dtype = tokenizer, device, dtype
        cfg = getattr(tokenizer, "config", None)
        if hasattr(cfg, "sample_rate"):
            self.sample_rate = cfg.sample_rate
        elif hasattr(getattr(tokenizer, "feature_extractor", None), "sampling_rate"):
            self.sample_rate = tokenizer.feature_extractor.sampling_rate
        else:
            self.sample_rate = 12500
        if str(device).startswith("cuda") and num_graph_lengths > 0:
            _capture_decoder_cudagraphs(tokenizer.model.decoder, device, list(range(1, num_graph_lengths + 1)))

    @property
    def samples_per_frame(self) -> int:
        return int(self.tokenizer.model.decoder.total_upsample)                 # server.py:196-198

    @torch.inference_mode()
    def decode(self, inputs: List[dict]) -> Tuple[List, int]:
        return self.tokenizer.decode(inputs)

    def _codes(self, audio_codes) -> torch.Tensor:
        codes = torch.tensor(audio_codes, dtype=torch.long, device=self.device) if isinstance(audio_codes, list) \
            else audio_codes.to(self.device)
        return codes.transpose(0, 1).unsqueeze(0) if codes.dim() == 2 else codes      # [time, 16] -> [1, 16, time]

    @torch.inference_mode()
    def chunked_decode(self, inputs: List[dict], chunk_size: int = 300, left_context_size: int = 25) -> Tuple[List, int]:
        wav = self.tokenizer.model.decoder.chunked_decode(self._codes(inputs[0]["audio_codes"]), chunk_size, left_context_size)
        return [wav.squeeze().to(torch.float32).cpu().numpy()], int(self.tokenizer.model.get_output_sample_rate())

    @torch.inference_mode()
    def decode_codec_ids(self, codec_ids: torch.Tensor) -> Tuple[List, int]:
        inputs = []
        for i in range(codec_ids.shape[0]):
            c = codec_ids[i]
            inputs.append({"audio_codes": c.transpose(0, 1) if c.dim() == 2 else c})
        return self.decode(inputs)

    @torch.inference_mode()
    def decode_window_batched(self, batch_inputs: List[dict], pcm16: bool = False, volume: float = 1.0) -> Tuple[List, int]:
        """Streaming windows (server.py:201-222, 372-478): each item is ``{"audio_codes": [context + new frames][16],
        "left_context_frames": n}``; the window is decoded through the graph of its length and the context's
        ``n * total_upsample`` samples are dropped.  ``pcm16``: the kept samples leave the GPU as int16
        (server.py:338-349 on the device) instead of float32."""
        spf, dec = self.samples_per_frame, self.tokenizer.model.decoder
        out = []
        for r in batch_inputs:
            wav = dec(self._codes(r["audio_codes"])).reshape(-1).to(torch.float32)
            skip = (r.get("left_context_frames") or 0) * spf
            wav = wav[skip:] if skip < wav.numel() else wav[:0]
            if pcm16:
                out.append(nat.float_to_pcm16(wav.contiguous(), volume).cpu().numpy())
            else:
                out.append(wav.cpu().numpy())
        return out, int(self.tokenizer.model.get_output_sample_rate())


def float_to_pcm16(wav: torch.Tensor, volume: float = 1.0) -> torch.Tensor:
    """``_float_to_pcm16`` (+ ``_apply_volume_pcm16``) of examples/server.py:338-349 on the device."""
    return nat.float_to_pcm16(wav.to(torch.float32), volume)
