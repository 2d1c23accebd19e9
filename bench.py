#!/usr/bin/env python
"""bench.py - codec frames/s of the Qwen3-TTS decode hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--streams B] [--impl ours|reference]

A STEP is one codec frame (16 codes) for every stream: 1 talker decode pass (28 layers) + 1 predictor
prefill (2 tokens/stream) + 14 predictor decode passes (5 layers) + samplers + the 16-codebook embedding
sum.  Workload: Qwen3-TTS-12Hz-1.7B shapes, random-init bf16 weights, B concurrent
synthetic streams with a 64-token prompt, top-k 50 sampling at T=1.0/0.9, EOS ignored.

  value     frames/s with everything resident in HBM: one CUDA-graph launch per frame (DeviceFrameLoop)
  e2e       frames/s through the reference-facing step API (TalkerLLM / PredictorLLM add_request/step)
            with HOST buffers crossing both ways every frame (StepApiFrameLoop)
  roofline  the dominant kernel (weight-streaming tcgen05 GEMM) + roofline_kernels for every other kernel of the path
  reference_kernels   cuBLAS / flash-attn / Triton on the same shapes, timed the same way (bench_kernels.py)
  other_workloads     1 / 8 / 128 / 256 / 512 streams;  icl_workload = BASELINE configs[3] (4k-token prompts)
  prompt_prep         f3: prompt assembly for a 64-request burst and for one 4096-frame ICL request
  cpu_baseline   the CPU oracle (port of the reference's torch path) on the host cores, bounded sample

N > 1: independent per-GPU replicas (no collective on the data path); barrier + max-over-ranks timing.
`--impl reference` times the oracle port on the host cores (the reference's own modules need
/root/reference, which does not exist on the GPU box).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME_HZ = 12.5            # codec frames per second of audio (12 Hz tokenizer, interface.py:426)
PROMPT_LEN = 64


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--streams", type=int, default=64, help="concurrent streams per GPU")
    ap.add_argument("--model", default="1.7B")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=24)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-streams", type=int, default=0, help="streams of the CPU legs (0 = the same as --streams)")
    ap.add_argument("--also-streams", default="1,8,128,256,512",
                    help="extra stream counts reported under other_workloads (BASELINE configs[1], [2], [4])")
    ap.add_argument("--no-icl", action="store_true", help="skip the BASELINE configs[3] leg (4k-token prompts, 8 streams)")
    ap.add_argument("--no-reference-kernels", action="store_true")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak():
    """Dense bf16 TFLOP/s: the sustained figure (kernels timed back to back inside a graph)."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        for key in ("bf16_tflops_sustained", "bf16_sustained_tflops", "bf16_tflops"):
            if key in p:
                return float(p[key]), f"measured (MEASURED_PEAKS.json {key})"
        raise KeyError
    except Exception:
        return 1400.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ GPU arm
def build_engines(model_size: str, streams: int, max_frames: int, device: str, prompt_len: int = PROMPT_LEN):
    from nano_qwen3tts_vllm_b200.config import qwen3_tts_config
    from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
    mc = qwen3_tts_config(model_size)
    page = 256
    width = (prompt_len + max_frames + 2 + page - 1) // page
    max_len = width * page
    common = dict(max_num_seqs=max(streams, 2), max_model_len=max_len, max_num_batched_tokens=max(16384, max_len),
                  device=device, graph_batch_sizes=[streams])
    talker = TalkerLLM(mc, state_dict="random:1234", num_kvcache_blocks=2 * streams * width, **common)
    predictor = PredictorLLM(mc, state_dict="random:1234", num_kvcache_blocks=2 * streams, **common)
    return mc, talker, predictor


def synthetic_prompts(streams: int, hidden: int, pinned: bool, prompt_len: int = PROMPT_LEN):
    import torch
    out = []
    for s in range(streams):
        g = torch.Generator().manual_seed(1000 + s)
        p = (torch.randn(1, prompt_len, hidden, generator=g) * 0.05).to(torch.bfloat16)
        out.append(p.pin_memory() if pinned else p)
    pad = (torch.randn(1, hidden, generator=torch.Generator().manual_seed(999)) * 0.05).to(torch.bfloat16)
    return out, pad


def frame_bytes(mc, lens_sum: float) -> float:
    """Algorithmic HBM bytes of one frame for all streams: every weight matrix streamed
    once per pass + the talker's KV read."""
    t, p = mc.talker, mc.predictor

    def layer_params(c):
        return c.hidden_size * (c.q_size + 2 * c.kv_size) + c.q_size * c.hidden_size + 3 * c.hidden_size * c.intermediate_size
    talker_w = (t.num_hidden_layers * layer_params(t) + t.vocab_size * t.hidden_size) * 2
    pred_pass = (p.num_hidden_layers * layer_params(p) + p.vocab_size * p.hidden_size
                 + (t.hidden_size * p.hidden_size if t.hidden_size != p.hidden_size else 0)) * 2
    kv = t.num_hidden_layers * 2 * t.kv_size * 2 * lens_sum
    return talker_w + (p.num_code_groups - 1) * pred_pass + kv


def ncu_traffic() -> dict:
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures
    (profiles/ncu_traffic.json, written from the .ncu-rep files by hand; see profiles/*_ncu_*.md)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def _graph_ms(fn, launches: int, iters: int = 3):
    """Median duration of one launch: `fn` issues `launches` launches; they are captured into one CUDA graph (no host
    gaps) and the replay is bracketed by CUDA events on the launching stream."""
    import torch
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    times = []
    for it in range(iters + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        if it:
            times.append(e0.elapsed_time(e1) / launches)
    return sorted(times)[len(times) // 2]


def time_attention_long(device, streams: int = 64, ctx_len: int = 4096, layers: int = 3):
    """paged_attn_kernel at the north-star shape (BASELINE configs[3]: ~4k-token KV): `streams` sequences of `ctx_len`
    cached tokens, scattered pages, `layers` launches over different caches (each 2 x streams x ctx_len x 2 KiB)."""
    import torch
    import nano_qwen3tts_vllm_b200._native as nat
    page, kvh, h, hd = 256, 8, 16, 128
    pages_per = (ctx_len + page - 1) // page
    npages = streams * pages_per
    caches = [(torch.randn(npages, page, kvh, hd, device=device, dtype=torch.bfloat16),
               torch.randn(npages, page, kvh, hd, device=device, dtype=torch.bfloat16)) for _ in range(layers)]
    bt = torch.randperm(npages, generator=torch.Generator().manual_seed(13)).to(torch.int32).view(streams, pages_per).to(device)
    lens = torch.full((streams,), ctx_len, dtype=torch.int32, device=device)
    q = torch.randn(streams, h, hd, device=device, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    max_ctx = pages_per * page
    ws = torch.empty(max(16, nat.paged_attn_decode_workspace_bytes(streams, h, hd, max_ctx)), dtype=torch.uint8, device=device)

    def launches():
        for kc, vc in caches:
            nat.paged_attn_decode(q, kc, vc, bt, lens, hd ** -0.5, max_ctx, out=out, workspace=ws)
    ms = _graph_ms(launches, layers)
    return ms, float(streams * (4096.0 * ctx_len + 8192.0))


def run_device_loop(talker, predictor, mc, streams, steps, warmup, sampler: ClockSampler | None, dist_barrier=None,
                    prompt_len: int = PROMPT_LEN, prefill_ms: list | None = None):
    import torch
    import nano_qwen3tts_vllm_b200._native as nat
    from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
    dev = talker.model_runner.device
    loop = DeviceFrameLoop(talker.model_runner, predictor.model_runner, streams, steps + warmup + 4, prompt_len)
    prompts, pad = synthetic_prompts(streams, mc.talker.hidden_size, pinned=False, prompt_len=prompt_len)
    prompts_dev, pad_dev = [p[0].to(dev) for p in prompts], pad.to(dev)
    loop.start(prompts_dev, pad_dev)
    if prefill_ms is not None:                 # the packed varlen prefill of all streams, timed warm (second run)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loop.start(prompts_dev, pad_dev)
        e1.record()
        torch.cuda.synchronize()
        prefill_ms.append(e0.elapsed_time(e1))
    n0 = nat.launch_count()
    loop.capture()
    # capture() runs one warm-up frame eagerly and records one: half of the delta is one frame
    lib_launches_per_frame = (nat.launch_count() - n0) // 2
    loop.run(max(warmup, 3))
    torch.cuda.synchronize()
    if dist_barrier:
        dist_barrier()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop.run(steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    if dist_barrier:
        dist_barrier()
    return loop, ms, lib_launches_per_frame, clocks


def run_e2e(talker, predictor, mc, streams, steps):
    """Frames/s through the step API with host buffers (see StepApiFrameLoop)."""
    import torch
    from nano_qwen3tts_vllm_b200.engine.frame_loop import StepApiFrameLoop
    talker.model_runner.kv_cache.zero_()
    prompts, pad = synthetic_prompts(streams, mc.talker.hidden_size, pinned=True)
    api = StepApiFrameLoop(talker, predictor, pad)
    api.add_streams(prompts)
    for _ in range(3):                         # prefill + 2 warm frames
        api.step_frame()
    torch.cuda.synchronize()
    h0, d0 = api.h2d_bytes, api.d2h_bytes
    # three back-to-back segments of `steps` frames, the median reported: this leg is host-paced (4 host <-> device
    # round trips per frame) and a single 0.2 s window is at the mercy of whatever else the box's cores are doing
    rates = []
    for _ in range(3):
        t0 = time.perf_counter()
        done = 0
        for _ in range(steps):
            done += api.step_frame()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        rates.append(done / dt)
    api.close()
    return sorted(rates)[1], (api.h2d_bytes - h0) / (3 * steps), (api.d2h_bytes - d0) / (3 * steps), dt


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_oracle_frames(model_size: str, streams: int, frames: int, warm_frames: int, budget_s: float,
                      prompt_len: int = PROMPT_LEN):
    """The CPU oracle (oracle/model.py: the reference's torch path restated op by op) timed on the
    host cores: `streams` streams with a PROMPT_LEN-token prompt, `frames` whole frames after
    `warm_frames`.  Returns (frames/s, cores, description, seconds)."""
    import torch
    from nano_qwen3tts_vllm_b200.config import qwen3_tts_config
    from oracle import model as OM
    from oracle import ops as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    mc = qwen3_tts_config(model_size)

    def dims(c, mrope):
        return OM.StackDims(hidden_size=c.hidden_size, intermediate_size=c.intermediate_size,
                            num_hidden_layers=c.num_hidden_layers, vocab_size=c.vocab_size, rope_theta=c.rope_theta,
                            mrope_section=[24, 20, 20] if mrope else None, mrope_interleaved=mrope)
    td, pd = dims(mc.talker, True), dims(mc.predictor, False)

    def fast_weights(d, extra):
        g = torch.Generator().manual_seed(7)
        w = {}
        qs, kvs = d.num_attention_heads * d.head_dim, d.num_key_value_heads * d.head_dim

        def mat(*shape):
            return torch.empty(*shape, dtype=torch.bfloat16).normal_(0, 0.02, generator=g)
        for i in range(d.num_hidden_layers):
            p = f"model.layers.{i}."
            w[p + "self_attn.q_proj.weight"] = mat(qs, d.hidden_size)
            w[p + "self_attn.k_proj.weight"] = mat(kvs, d.hidden_size)
            w[p + "self_attn.v_proj.weight"] = mat(kvs, d.hidden_size)
            w[p + "self_attn.o_proj.weight"] = mat(d.hidden_size, qs)
            w[p + "self_attn.q_norm.weight"] = torch.ones(d.head_dim, dtype=torch.bfloat16)
            w[p + "self_attn.k_norm.weight"] = torch.ones(d.head_dim, dtype=torch.bfloat16)
            w[p + "mlp.gate_proj.weight"] = mat(d.intermediate_size, d.hidden_size)
            w[p + "mlp.up_proj.weight"] = mat(d.intermediate_size, d.hidden_size)
            w[p + "mlp.down_proj.weight"] = mat(d.hidden_size, d.intermediate_size)
            w[p + "input_layernorm.weight"] = torch.ones(d.hidden_size, dtype=torch.bfloat16)
            w[p + "post_attention_layernorm.weight"] = torch.ones(d.hidden_size, dtype=torch.bfloat16)
        w["model.norm.weight"] = torch.ones(d.hidden_size, dtype=torch.bfloat16)
        w.update({k: mat(*shape) for k, shape in extra.items()})
        return w
    tw = fast_weights(td, {"model.codec_embedding.weight": (td.vocab_size, td.hidden_size),
                           "codec_head.weight": (td.vocab_size, td.hidden_size)})
    pextra = {}
    for j in range(15):
        pextra[f"model.codec_embedding.{j}.weight"] = (pd.vocab_size, td.hidden_size)
        pextra[f"lm_head.{j}.weight"] = (pd.vocab_size, pd.hidden_size)
    if td.hidden_size != pd.hidden_size:
        pextra["small_to_mtp_projection.weight"] = (pd.hidden_size, td.hidden_size)
    pw = fast_weights(pd, pextra)
    if td.hidden_size != pd.hidden_size:
        pw["small_to_mtp_projection.bias"] = torch.zeros(pd.hidden_size, dtype=torch.bfloat16)
    pages = (prompt_len + frames + warm_frames + 2 + 255) // 256
    talker = OM.Talker(td, tw, streams * pages)
    predictor = OM.Predictor(pd, pw, streams, td.hidden_size)
    prompts, pad = synthetic_prompts(streams, td.hidden_size, pinned=False, prompt_len=prompt_len)
    gen = torch.Generator().manual_seed(42)
    stamps = []

    def sampler(logits, temps):                       # sampler.py:15-24 with an explicit Exp(1) draw
        noise = torch.empty(logits.shape, dtype=torch.float32).exponential_(1, generator=gen)
        if logits.shape[-1] == td.vocab_size:          # one talker call per frame: a frame boundary
            stamps.append(time.perf_counter())
            if len(stamps) > warm_frames + 1 and stamps[-1] - stamps[warm_frames] > budget_s:
                raise TimeoutError
        return O.topk_sample(logits, temps, noise, top_k=50)
    with torch.inference_mode():
        try:
            OM.run_frames(talker, predictor, [p[0] for p in prompts], frames + warm_frames, pad, sampler)
        except TimeoutError:
            pass
    # stamps[0] = prefill done; stamps[i] = frame i done
    done = len(stamps) - 1 - warm_frames
    if done < 1:
        done, dt = len(stamps) - 1, stamps[-1] - stamps[0]
    else:
        dt = stamps[-1] - stamps[warm_frames]
    fps = streams * done / dt
    what = (f"oracle/model.py run_frames, {model_size} shapes, {streams} streams x {done} frames after {warm_frames} warm, "
            f"prompt {prompt_len}, torch {torch.__version__} CPU bf16, {cores} threads")
    return fps, cores, what, dt


def reference_arm(a):
    """`--impl reference`: the reference's CPU path (oracle port) on the host cores, one step = one frame of
    the SAME workload as the GPU arm (the same stream count: `config` describes the work that ran), bounded in
    time: the run stops after ~150 s of frames and reports the frames it completed."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, a.steps)
    streams = a.cpu_streams or a.streams
    fps, cores, what, dt = cpu_oracle_frames(a.model, streams, steps, max(1, min(a.warmup, 2)), budget_s=150.0)
    line = {"impl": "reference", "metric": "codec frames/s", "value": round(fps, 3), "unit": "frames/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(1000 * streams / fps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(a, streams),
            "cpu_baseline": {"value": round(fps, 3), "unit": "frames/s", "cores": cores, "kind": "port", "sample": what},
            "e2e": {"value": round(fps, 3), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "rtf_per_stream": round(FRAME_HZ * streams / fps, 4)}
    print(json.dumps(line), flush=True)


def time_prompt_prep(device):
    """Prompt assembly for (a) 64 simultaneous arrivals of ~20-word prompts (BASELINE configs[4]'s admission burst) and
    (b) one voice-clone ICL request with a 4096-frame reference (configs[3]); random-init embeddings at the 1.7B widths
    (text 151936 x 2048 -> hidden 2048).  CUDA events around the public call, host planning included."""
    import types
    import torch
    from nano_qwen3tts_vllm_b200.models.qwen3_tts_talker import Qwen3TTSTalkerResizeMLP
    from nano_qwen3tts_vllm_b200.utils.generation import prepare_inputs
    g = torch.Generator(device=device).manual_seed(77)
    rnd = lambda *s: (torch.randn(*s, device=device, generator=g) * 0.02).to(torch.bfloat16)
    tc = types.SimpleNamespace(num_code_groups=16, codec_nothink_id=2155, codec_think_id=2154, codec_think_bos_id=2156,
                               codec_think_eos_id=2157, codec_pad_id=2148, codec_bos_id=2149,
                               codec_language_id={"english": 2050}, spk_id={}, spk_is_dialect={})
    cfg = types.SimpleNamespace(talker_config=tc, tts_bos_token_id=151672, tts_eos_token_id=151673, tts_pad_token_id=151671)
    text = types.SimpleNamespace(weight=rnd(151936, 2048))
    codec = types.SimpleNamespace(weight=rnd(3072, 2048))
    books = rnd(15, 2048, 2048)
    torch.set_default_dtype(torch.bfloat16)
    try:
        proj = Qwen3TTSTalkerResizeMLP(2048, 2048, 2048, bias=True).to(device)
    finally:
        torch.set_default_dtype(torch.float32)
    with torch.no_grad():
        for p in proj.parameters():
            p.copy_(rnd(*p.shape))
    ids = lambda n: torch.randint(1000, 150000, (1, n), device=device, generator=g)
    kw = dict(text_embedding=text, input_embedding=codec, text_projection=proj, device=device, code_predictor_embeddings=books)

    def timed(fn, iters=10):
        fn(); fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters, out

    burst = [ids(40) for _ in range(64)]
    ms_a, out_a = timed(lambda: prepare_inputs(cfg, burst, languages=["english"] * 64, **kw))
    code = torch.randint(0, 2048, (4096, 16), device=device, generator=g)
    clone = {"ref_code": [code], "ref_spk_embedding": [rnd(2048)], "x_vector_only_mode": [False], "icl_mode": [True]}
    one, ref = [ids(80)], [ids(60)]
    ms_b, out_b = timed(lambda: prepare_inputs(cfg, one, ref_ids=ref, voice_clone_prompt=clone, languages=["english"],
                                               voice_clone_spk_embeds=clone["ref_spk_embedding"], **kw))
    return {"what": "prepare_inputs (utils/generation.py:18-379, 412-515): one text_embedding gather + one text_projection "
                    "over all tokens, one codec gather / 16-codebook gather-sum, ONE b2tts_prompt_assemble launch; host "
                    "planning inside the timed region",
            "burst_64_requests_ms": round(ms_a, 3), "burst_prompt_rows": int(out_a[0].shape[1]),
            "icl_4096_frames_ms": round(ms_b, 3), "icl_prompt_rows": int(out_b[0].shape[1])}


def workload_config(a, streams):
    return {"workload": f"Qwen3-TTS-12Hz-{a.model} talker+code-predictor, {streams} concurrent streams per GPU "
                        f"(BASELINE configs[4]; configs[1]/[2] = 1/8 streams under other_workloads), prompt {PROMPT_LEN} tokens, "
                        f"random-init weights, top-k 50 sampling T=1.0/0.9, EOS ignored, paged KV block 256",
            "streams_per_gpu": streams, "prompt_len": PROMPT_LEN, "frame": "1 talker decode + predictor prefill + 14 predictor decodes",
            "l2": "no flush needed: each step streams >5 GB of weights + KV through the 126 MB L2",
            "parallelism": f"{a.gpus} independent replicas"}


# ------------------------------------------------------------------------------------------ main
def main():
    a = parse()
    if a.impl == "reference":
        reference_arm(a)
        return
    import torch
    import nano_qwen3tts_vllm_b200._native as nat
    nat.lib()                                            # no library, no bench: there is no fallback
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = f"cuda:{local}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(device))
    barrier = (lambda: (dist.barrier(), torch.cuda.synchronize())) if dist else None

    total_frames = a.steps + a.warmup + 8
    mc, talker, predictor = build_engines(a.model, a.streams, total_frames + 3 * a.e2e_steps + 8, device)
    sampler = ClockSampler(local) if rank == 0 else None
    loop, ms, lib_per_frame, clocks = run_device_loop(talker, predictor, mc, a.streams, a.steps, a.warmup, sampler, barrier)
    from nano_qwen3tts_vllm_b200 import replicas
    frames, ms = replicas.aggregate(a.streams * a.steps, ms, dist, device)     # sum of units, max of time over ranks
    value = frames / (ms / 1e3)
    ms_per_step = ms / a.steps

    line = None
    if rank == 0:
        import bench_kernels as BK
        peak, peak_src = peaks()
        tpeak, tpeak_src = tensor_peak()
        ctx_now = loop.seq_lens.to(torch.int32)                   # context lengths at the end of the timed run
        mean_ctx = float(ctx_now.float().mean())
        gemm_ms, gemm_alg, gemm_n = BK.gemm_stack(talker.model_runner, a.streams)
        pg_ms, pg_alg, pg_n = BK.gemm_stack(predictor.model_runner, a.streams)
        from nano_qwen3tts_vllm_b200.layers.linear import GEMM_NATIVE_MAX_M
        m2 = min(2 * a.streams, GEMM_NATIVE_MAX_M)                # longer prefills take cuBLAS (layers/linear.py)
        pg2_ms, pg2_alg, _ = BK.gemm_stack(predictor.model_runner, m2)
        att_ms, att_alg = BK.decode_attention(a.streams, ctx_now, device, layers=8)
        a512_ms, a512_alg = BK.decode_attention(64, 512, device, layers=6)
        a4k_ms, a4k_alg = BK.decode_attention(64, 4096, device, layers=3)
        long_ms, long_alg = time_attention_long(device)
        pf_ms, pf_flops = BK.prefill_attention([4096], device)
        st_ms, st_alg = BK.kv_store(4096, device)
        sm_ms, sm_alg = BK.sampler(a.streams, mc.predictor.vocab_size, device)
        lens_sum = float((loop.seq_lens - 1).sum())
        step_bytes = frame_bytes(mc, lens_sum)
        traffic = ncu_traffic()

        def hbm(kernel, ms_, alg_, how, **extra):
            g = alg_ / ms_ / 1e6
            d = {"kernel": kernel, "bound": "hbm", "achieved": round(g, 1), "peak": peak, "unit": "GB/s",
                 "frac": round(g / peak, 4), "traffic": None, "algorithmic_bytes_per_launch": alg_,
                 "launch_ms": round(ms_, 5), "how": how}
            d.update(extra)
            return d
        # the dominant kernel of the frame by time share (profiles/*_launches_summary.txt, tools/pass_ablation.py:
        # ~82 % of the frame): the weight-streaming GEMM
        roofline = hbm("gemm_stream_kernel (qkv / o / gate_up / down projections of the talker, tcgen05 + TMEM)", gemm_ms, gemm_alg,
                       f"CUDA events around one graph of the talker's {gemm_n} projection launches at M={a.streams} over the "
                       "model's own weights (2.8 GB >> L2), production epilogues + PDL; bytes = 2*(K*N + M*K + M*N) "
                       "averaged over the 4 projections of a layer",
                       traffic=traffic.get("gemm_stream_kernel"), traffic_note=traffic.get("_gemm_stream_kernel"),
                       peak_source=peak_src, launches_timed=gemm_n)
        roofline_kernels = [
            hbm("gemm_stream_kernel, code-predictor shapes (H 1024, I 3072), decode pass", pg_ms, pg_alg,
                f"graph of the predictor's {pg_n} projection launches at M={a.streams} (5 layers x 4; 31 MB of weights per "
                "layer: latency-bound launches)"),
            hbm("gemm_stream_kernel, code-predictor shapes, 2-token prefill pass", pg2_ms, pg2_alg,
                f"same at M={m2} rows"),
            hbm("fused_decode_attn_kernel (qk-norm + M-RoPE + KV write + paged GQA attention + combine), bench contexts",
                att_ms, att_alg, "graph of 8 launches over different caches at the run's end-of-run context lengths; "
                "bytes = sum_b(4096*L_b + 8192)", mean_context=round(mean_ctx, 1)),
            hbm("fused_decode_attn_kernel, 64 sequences x 512 cached tokens", a512_ms, a512_alg, "graph of 6 launches, cold caches",
                traffic=traffic.get("fused_decode_attn_kernel"), traffic_note=traffic.get("_fused_decode_attn_kernel")),
            hbm("fused_decode_attn_kernel, 64 sequences x 4096 cached tokens (BASELINE configs[3] KV length)", a4k_ms, a4k_alg,
                "graph of 3 launches over different 1 GiB caches, scattered pages"),
            hbm("paged_attn_kernel<2,false> (stand-alone a2 entry point), 64 sequences x 4096 cached tokens", long_ms, long_alg,
                "graph of 3 launches over different 1 GiB caches, scattered pages",
                traffic=traffic.get("paged_attn_kernel"), traffic_note=traffic.get("_paged_attn_kernel")),
            hbm("store_kvcache_kernel (stand-alone a1 entry point), 4096 tokens", st_ms, st_alg,
                "graph of 8 launches; bytes = tokens * 1024 * 2 B * 2 (K, V) * 2 (read + write)"),
            hbm(f"topk_sample_row_kernel, [{a.streams}, {mc.predictor.vocab_size}] bf16 logits, top-k 50", sm_ms, sm_alg,
                "graph of 16 launches; bytes = the logits (latency-bound: one CTA per row)"),
            {"kernel": "prefill_attn_tc_kernel (tcgen05 128x128 tiles, TMEM), one 4096-token sequence (BASELINE configs[3])",
             "bound": "tensor", "achieved": round(pf_flops / pf_ms / 1e9, 1), "peak": tpeak, "unit": "TFLOP/s",
             "frac": round(pf_flops / pf_ms / 1e9 / tpeak, 4), "traffic": None, "flops_per_launch": pf_flops,
             "launch_ms": round(pf_ms, 5), "peak_source": tpeak_src,
             "how": "graph of 4 launches over different QKV buffers; flops = 4*heads*hd*n(n+1)/2 (causal half)"}]
        ref_kernels = None
        if not a.no_reference_kernels:
            ref_kernels = {"about": "the kernels the reference runs for the same ops, timed the same way on this GPU "
                                    "(bench_kernels.py); library code, measured only"}

            def leg(name, fn):
                try:
                    ref_kernels[name] = fn()
                except Exception as e:                                   # a missing library must not void the bench line
                    ref_kernels[name] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}

            def cublas(runner, m, ours_ms):
                ms_, alg_, n_ = BK.gemm_stack(runner, m, impl="cublas")
                return {"library": "cuBLAS via F.linear (layers/linear.py:72-73)", "launch_ms": round(ms_, 5),
                        "GBps": round(alg_ / ms_ / 1e6, 1), "ours_launch_ms": round(ours_ms, 5), "speedup": round(ms_ / ours_ms, 3),
                        "note": "plain products only: the reference's separate RMSNorm / residual / SiLU*mul launches are fused "
                                "into ours and not counted on the library side"}

            def flash_dec(b_, ctx_, layers_, ours_ms):
                ms_, alg_ = BK.decode_attention(b_, ctx_, device, layers=layers_, impl="flash")
                return {"library": "flash_attn_with_kvcache (layers/attention.py:92-95)", "launch_ms": round(ms_, 5),
                        "GBps": round(alg_ / ms_ / 1e6, 1), "ours_launch_ms": round(ours_ms, 5), "speedup": round(ms_ / ours_ms, 3),
                        "note": "attention only: the reference's separate q/k norm, RoPE and Triton KV store launches are "
                                "inside ours and not counted on the library side"}

            def flash_pre():
                ms_, fl_ = BK.prefill_attention([4096], device, impl="flash")
                return {"library": "flash_attn_varlen_func (layers/attention.py:78-91)", "launch_ms": round(ms_, 5),
                        "TFLOPs": round(fl_ / ms_ / 1e9, 1), "ours_launch_ms": round(pf_ms, 5), "speedup": round(ms_ / pf_ms, 3)}

            def triton_store():
                ms_, alg_ = BK.kv_store(4096, device, impl="triton")
                return {"library": "Triton store_kvcache restated (layers/attention.py:10-41)", "launch_ms": round(ms_, 5),
                        "GBps": round(alg_ / ms_ / 1e6, 1), "ours_launch_ms": round(st_ms, 5), "speedup": round(ms_ / st_ms, 3)}
            leg(f"gemm_talker_M{a.streams}", lambda: cublas(talker.model_runner, a.streams, gemm_ms))
            leg(f"gemm_predictor_M{a.streams}", lambda: cublas(predictor.model_runner, a.streams, pg_ms))
            leg(f"gemm_predictor_M{m2}", lambda: cublas(predictor.model_runner, m2, pg2_ms))
            leg("decode_attention_bench_contexts", lambda: flash_dec(a.streams, ctx_now, 8, att_ms))
            leg("decode_attention_64x512", lambda: flash_dec(64, 512, 6, a512_ms))
            leg("decode_attention_64x4096", lambda: flash_dec(64, 4096, 3, a4k_ms))
            leg("prefill_attention_4096", flash_pre)
            leg("kv_store_4096", triton_store)

            def frame_library():
                # the layer-level answer to "the library GEMM is faster per launch": the same frame loop in a child process
                # with every projection on F.linear (cuBLAS) and the norms / residual adds / SiLU*mul as separate launches
                # of this repo's kernels (B2TTS_GEMM_NATIVE_MAX_M=0, tools/library_path_frame.py)
                out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "library_path_frame.py"), "--child",
                                      str(a.streams), "64"], env={**os.environ, "B2TTS_GEMM_NATIVE_MAX_M": "0"},
                                     capture_output=True, text=True, timeout=600)
                d = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
                return {"library": "cuBLAS via F.linear for every projection + separate RMSNorm / add / SiLU*mul launches "
                                   "(measured only: not a product path)",
                        "ms_per_step": round(d["ms_per_frame"], 4), "ours_ms_per_step": round(ms_per_step, 4),
                        "speedup": round(d["ms_per_frame"] / ms_per_step, 3)}
            if world == 1:
                leg(f"frame_{a.streams}_streams_library_gemm", frame_library)
        step_roof = {"bound": "hbm", "achieved": round(step_bytes / (ms_per_step / 1e3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(step_bytes / (ms_per_step / 1e3) / 1e9 / peak, 4), "algorithmic_bytes_per_step": step_bytes,
                     "note": "whole frame: weights streamed once per pass (talker + 15 predictor passes) + talker KV"}
        line = {"metric": "codec frames/s", "value": round(value, 1), "unit": "frames/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": workload_config(a, a.streams), "clocks": clocks,
                "rtf_per_stream": round(ms_per_step / 1e3 * FRAME_HZ, 5),
                "rounding": "eager (the reference source op by op: what its CPU path computes; ROUND_FUSED = its inductor GPU "
                            "path is the other mode, same kernels and launch count)",
                "roofline": roofline, "roofline_kernels": roofline_kernels, "step_roofline": step_roof,
                "gpu_launches": int(lib_per_frame * a.steps),
                "launches_per_step": {"libb200tts": int(lib_per_frame), "note": "includes the tcgen05 GEMMs; exponential_ noise and small torch copies are extra"}}
        if ref_kernels is not None:
            line["reference_kernels"] = ref_kernels
    # ---- end to end through the step API with host buffers
    if not a.no_e2e:
        e2e_fps, h2d, d2h, _ = run_e2e(talker, predictor, mc, a.streams, a.e2e_steps)
        if dist:
            t = torch.tensor([e2e_fps], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            e2e_fps = float(t.item())
        if rank == 0:
            line["e2e"] = {"value": round(e2e_fps, 1), "unit": "frames/s", "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "steps": a.e2e_steps, "segments": "median of 3",
                           "api": "TalkerLLM.add_requests/step_launch/step_collect + PredictorLLM.add_requests/step_launch/step_collect (StepApiFrameLoop, pipelined), pinned host tensors"}
    # ---- other stream counts (BASELINE configs[1], [2], [4]) on rank 0, short runs
    if rank == 0 and world == 1 and a.also_streams:
        others = {}
        for b in [int(x) for x in a.also_streams.split(",") if x]:
            if b == a.streams:
                continue
            talker.exit(); predictor.exit()
            del loop, talker, predictor
            torch.cuda.empty_cache()
            steps_b = a.steps if b <= 64 else max(16, a.steps // 2)
            try:
                mc, talker, predictor = build_engines(a.model, b, steps_b + a.warmup + 8, device)
                loop, ms_b, _, _ = run_device_loop(talker, predictor, mc, b, steps_b, a.warmup, None)
                others[str(b)] = {"frames_per_s": round(b * steps_b / (ms_b / 1e3), 1), "ms_per_step": round(ms_b / steps_b, 4),
                                  "rtf_per_stream": round(ms_b / steps_b / 1e3 * FRAME_HZ, 5), "steps": steps_b}
            except Exception as e:
                others[str(b)] = {"failed": f"{type(e).__name__}: {e}"[:300]}
                mc, talker, predictor = build_engines(a.model, 1, 16, device)
                loop = None
        line["other_workloads"] = others
    # ---- BASELINE configs[3]: voice-clone ICL, ~4k-token prompts (varlen prefill + decode), 8 streams
    if rank == 0 and world == 1 and not a.no_icl:
        talker.exit(); predictor.exit()
        del loop, talker, predictor
        torch.cuda.empty_cache()
        try:
            b, plen, frames = 8, 4096, 256
            mc, talker, predictor = build_engines(a.model, b, frames + 16, device, prompt_len=plen)
            pre = []
            loop, ms_b, _, _ = run_device_loop(talker, predictor, mc, b, frames, 4, None, prompt_len=plen, prefill_ms=pre)
            t = mc.talker
            lin = 2.0 * b * plen * t.num_hidden_layers * (t.hidden_size * (t.q_size + 2 * t.kv_size) + t.q_size * t.hidden_size
                                                          + 3 * t.hidden_size * t.intermediate_size)
            att = b * t.num_hidden_layers * 4.0 * t.num_attention_heads * t.head_dim * plen * (plen + 1) / 2
            line["icl_workload"] = {
                "workload": f"BASELINE configs[3]: {b} streams x {plen}-token prompt (packed varlen prefill: tcgen05 prefill "
                            f"attention + cuBLAS projections above {os.environ.get('B2TTS_GEMM_NATIVE_MAX_M', '512')} rows), "
                            f"then {frames} frames of decode at ~{plen + frames // 2}-token contexts",
                "prefill_ms": round(pre[0], 3), "prefill_tokens_per_s": round(b * plen / (pre[0] / 1e3), 1),
                "prefill_TFLOPs": round((lin + att) / (pre[0] / 1e3) / 1e12, 1),
                "decode_frames_per_s": round(b * frames / (ms_b / 1e3), 1), "decode_ms_per_step": round(ms_b / frames, 4),
                "rtf_per_stream": round(ms_b / frames / 1e3 * FRAME_HZ, 5)}
        except Exception as e:
            line["icl_workload"] = {"failed": f"{type(e).__name__}: {e}"[:300]}
            mc, talker, predictor = None, None, None
    # ---- f3: prompt-prep embedding assembly (utils/generation.py of the reference) at the 1.7B widths
    if rank == 0 and world == 1 and not a.no_icl:
        try:
            line["prompt_prep"] = time_prompt_prep(device)
        except Exception as e:
            line["prompt_prep"] = {"failed": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0 and not a.no_cpu_baseline and world == 1:
        fps, cores, what, dt = cpu_oracle_frames(a.model, a.cpu_streams or a.streams, 2, 1, budget_s=25.0)
        line["cpu_baseline"] = {"value": round(fps, 3), "unit": "frames/s", "cores": cores, "kind": "port", "sample": what}
        # BASELINE configs[0]: 0.6B, 1 stream, ~20-word (40-token) prompt - the reference's own CPU-runnable case
        fps0, cores0, what0, _ = cpu_oracle_frames("0.6B", 1, 8, 2, budget_s=20.0, prompt_len=40)
        line["cpu_baseline_configs0"] = {"value": round(fps0, 3), "unit": "frames/s", "cores": cores0, "kind": "port",
                                         "rtf": round(FRAME_HZ / fps0, 4), "sample": what0}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
