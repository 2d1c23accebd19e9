"""Model runner: batch layout, paged KV pool, CUDA-graph capture/replay of the decode step and the
sampler call (surface S2 of SURVEY.md section 8b; reference engine/model_runner/base.py:30-312).

B200-first differences from the reference, none visible through ``run(seqs, is_prefill)``:
  * the step's layout (positions, slot_mapping, context_lens, block_tables, temperatures, last tokens,
    generation steps) is written with numpy into ONE pinned staging buffer and crosses PCIe as ONE
    async copy, instead of 5-6 tensors built from Python lists and copied one by one
    (base.py:172-241);
  * the graph's static tensors are views of the device mirror of that buffer, so replay needs no
    per-field ``copy_/fill_/zero_`` staging kernels (talker_mode_runner.py:95-101);
  * block tables are sized from the real ``max_model_len`` (the reference pins it to 1024 and
    overflows its 4-column table beyond that);
  * one process per GPU, no NCCL group: replicas only (DESIGN.md "multi-GPU").
"""
from __future__ import annotations

import numpy as np
import torch

from ... import _native as nat
from ...config import Config
from ...layers.sampler import Sampler
from ...utils.context import reset_context, set_context
from ..sequence import Sequence

BF16 = torch.bfloat16


class StepLayout:
    """Pinned host staging + device mirror of the per-step index tensors, one region per field,
    each sized for ``max_bs`` rows (decode) - a whole-buffer copy is ~50 KB."""

    FIELDS = (("positions", np.int64, 1), ("last_tokens", np.int64, 1), ("slot_mapping", np.int32, 1),
              ("context_lens", np.int32, 1), ("gen_steps", np.int32, 1), ("temperatures", np.float32, 1),
              ("block_tables", np.int32, None))

    def __init__(self, max_bs: int, max_blocks: int, device):
        self.max_bs, self.max_blocks = max_bs, max_blocks
        off, self.offsets = 0, {}
        for name, dt, width in self.FIELDS:
            w = max_blocks if width is None else width
            nbytes = max_bs * w * np.dtype(dt).itemsize
            self.offsets[name] = (off, nbytes, dt, w)
            off += (nbytes + 15) // 16 * 16
        self.nbytes = off
        # a RING of pinned staging buffers: a step's upload is asynchronous and the host may already be filling the
        # layout of a later step (the predictor's deferred-token burst never synchronises in between), so a buffer
        # is rewritten only after the copy that read it has completed (per-buffer event)
        self.dev = torch.zeros(off, dtype=torch.uint8, device=device)
        tdt = {np.int64: torch.int64, np.int32: torch.int32, np.float32: torch.float32}
        self.d = {}
        for name, (o, n, dt, w) in self.offsets.items():
            dv = self.dev[o:o + n].view(tdt[dt])
            if w > 1 or name == "block_tables":
                dv = dv.view(max_bs, w)
            self.d[name] = dv
        self.hosts, self.hviews, self.events = [], [], []
        for _ in range(self.RING):
            host = torch.zeros(off, dtype=torch.uint8)
            if torch.cuda.is_available():
                host = host.pin_memory()
            hn = host.numpy()
            views = {}
            for name, (o, n, dt, w) in self.offsets.items():
                hv = hn[o:o + n].view(dt)
                if w > 1 or name == "block_tables":
                    hv = hv.reshape(max_bs, w)
                views[name] = hv
            views["slot_mapping"][:] = -1
            views["block_tables"][:] = -1
            views["temperatures"][:] = 1.0
            views["gen_steps"][:] = 1
            self.hosts.append(host)
            self.hviews.append(views)
            self.events.append(None)
        self.cur = 0
        self.host, self.h = self.hosts[0], self.hviews[0]
        self.upload()

    RING = 24

    def begin(self):
        """Rotate to the next staging buffer (waits only if its last upload is still pending) and return its views."""
        self.cur = (self.cur + 1) % self.RING
        ev = self.events[self.cur]
        if ev is not None:
            ev.synchronize()
        self.host, self.h = self.hosts[self.cur], self.hviews[self.cur]
        return self.h

    def upload(self):
        self.dev.copy_(self.host, non_blocking=True)
        if torch.cuda.is_available():
            ev = self.events[self.cur] or torch.cuda.Event()
            ev.record()
            self.events[self.cur] = ev


class ModelRunner:
    is_talker = False

    def __init__(self, config: Config, rank: int = 0, event=None, state_dict: dict | None = None):
        assert rank == 0, "replicas only: one runner per GPU"
        self.config = config
        self.block_size = config.kvcache_block_size
        self.enforce_eager = config.enforce_eager
        self.device = torch.device(config.device)
        self.rounding = config.rounding
        torch.cuda.set_device(self.device)
        nat.lib()                                               # fail loudly if the library is missing
        self.stack_config = self.pick_stack_config(config.hf_config)
        self.max_blocks = (config.max_model_len + self.block_size - 1) // self.block_size
        self.max_bs = min(config.max_num_seqs, 512)
        prev = torch.get_default_dtype()
        torch.set_default_dtype(BF16)
        try:
            with torch.device(self.device):
                self.model = self.build_model()
        finally:
            torch.set_default_dtype(prev)
        self.load_model(state_dict)
        self.post_init()

    # ------------------------------------------------------------------ subclass hooks
    def pick_stack_config(self, model_config):
        raise NotImplementedError

    def build_model(self):
        raise NotImplementedError

    def graph_batch_sizes(self) -> list[int]:
        raise NotImplementedError

    def hidden_in(self) -> int:
        return self.stack_config.hidden_size

    def load_model(self, state_dict: dict | str | None):
        if isinstance(state_dict, str) and state_dict.startswith("random"):     # "random" | "random:<seed>"
            from ...models.qwen3_tts_share import random_init_
            seed = int(state_dict.split(":")[1]) if ":" in state_dict else 1234
            random_init_(self.model, seed + (0 if self.is_talker else 1))
            return
        if state_dict is None and self.config.model:
            import os
            path = os.path.join(self.config.model, "model.safetensors")
            if os.path.exists(path):
                from safetensors.torch import load_file
                state_dict = load_file(path)
        if state_dict is None:
            raise RuntimeError("no weights: pass state_dict= or a model directory holding model.safetensors")
        self.model.load_state_dict(state_dict)

    # ------------------------------------------------------------------ init
    def post_init(self):
        self.sampler = Sampler()
        self.top_k = 50        # the reference never forwards SamplingParams.top_k; 1 = greedy
        self.layout = StepLayout(self.max_bs, self.max_blocks, self.device)
        self.allocate_kv_cache()
        self.graphs: dict[int, torch.cuda.CUDAGraph] = {}
        if not self.enforce_eager:
            self.capture_cudagraph()

    def exit(self):
        self.graphs = {}
        self.graph_pool = None
        torch.cuda.synchronize()
        from ...layers.linear import release_packed
        release_packed(self.model)

    def call(self, method_name, *args):
        return getattr(self, method_name)(*args)

    def kv_block_bytes(self) -> int:
        c = self.stack_config
        return 2 * c.num_hidden_layers * self.block_size * c.num_key_value_heads * c.head_dim * 2

    def allocate_kv_cache(self):
        """base.py:131-164: size the pool from free memory (or take ``num_kvcache_blocks`` as given),
        allocate ``[2, L, blocks, page, kv_heads, hd]`` and bind per-layer views to every module that
        has ``k_cache`` / ``v_cache``."""
        cfg, c = self.config, self.stack_config
        if cfg.num_kvcache_blocks <= 0:
            # the reference's accounting (base.py:131-151, restated in engine/memory.py and pinned against the
            # reference's own method): pages = (budget - used - peak + current) // page_bytes, at least 1
            from ..memory import num_kv_blocks
            free, total = torch.cuda.mem_get_info(self.device)
            stats = torch.cuda.memory_stats(self.device)
            blocks, raw = num_kv_blocks(total, free, torch.cuda.memory_allocated(self.device),
                                        stats.get("allocated_bytes.all.peak", 0), stats.get("allocated_bytes.all.current", 0),
                                        cfg.gpu_memory_utilization, cfg.process_gpu_memory_fraction, self.kv_block_bytes())
            if raw <= 0 and cfg.gpu_memory_utilization >= 0.05:                 # base.py:148-153
                import warnings
                warnings.warn(f"KV cache allocation would be 0 pages (gpu_memory_utilization="
                              f"{cfg.gpu_memory_utilization}); using 1 page")
            self.kv_blocks_uncapped = blocks
            # deviation, on purpose: pages beyond what max_num_seqs sequences of max_model_len tokens can address
            # are dead memory (the reference would torch.empty them anyway); the pool here is zero-filled
            cfg.num_kvcache_blocks = max(1, min(blocks, self.max_bs * self.max_blocks))
        self.kv_cache = torch.zeros(2, c.num_hidden_layers, cfg.num_kvcache_blocks, self.block_size,
                                    c.num_key_value_heads, c.head_dim, dtype=BF16, device=self.device)
        ws_bytes = nat.paged_attn_decode_workspace_bytes(1, c.num_attention_heads, c.head_dim,
                                                         self.max_blocks * self.block_size)
        for bs in range(1, self.max_bs + 1):
            ws_bytes = max(ws_bytes, nat.paged_attn_decode_workspace_bytes(
                bs, c.num_attention_heads, c.head_dim, self.max_blocks * self.block_size))
        self.attn_workspace = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=self.device)
        self.fused_attn_workspace = nat.fused_decode_attn_workspace(
            self.max_bs, c.num_attention_heads, c.head_dim, self.max_blocks * self.block_size, self.device)
        layer_id = 0
        for module in self.model.modules():
            if hasattr(module, "k_cache") and hasattr(module, "v_cache"):
                module.k_cache = self.kv_cache[0, layer_id]
                module.v_cache = self.kv_cache[1, layer_id]
                module.workspace = self.attn_workspace
                module.fused_workspace = self.fused_attn_workspace
                layer_id += 1

    # ------------------------------------------------------------------ batch layout
    def prepare_prefill(self, seqs: list[Sequence]):
        """Packed varlen layout (base.py:172-216): positions 0..T-1 per sequence, cu_seqlens, and a
        slot for every prompt token.  Returns (input_embeds [T, H], positions [T])."""
        bs = self.block_size
        lens = [len(s) - s.num_cached_tokens for s in seqs]
        total = sum(lens)
        idx = np.empty(total * 3 + 2 * (len(seqs) + 1), dtype=np.int64)     # positions | slots | cu_q | cu_k
        pos, slots = idx[:total], idx[total:2 * total]
        cu_q = idx[2 * total:2 * total + len(seqs) + 1]
        cu_k = idx[2 * total + len(seqs) + 1:2 * total + 2 * (len(seqs) + 1)]
        cu_q[0] = cu_k[0] = 0
        o = 0
        for i, s in enumerate(seqs):
            n, c = len(s), s.num_cached_tokens
            ar = np.arange(c, n, dtype=np.int64)
            pos[o:o + n - c] = ar
            bt = np.asarray(s.block_table, dtype=np.int64)
            slots[o:o + n - c] = bt[ar // bs] * bs + ar % bs
            o += n - c
            cu_q[i + 1] = cu_q[i] + (n - c)
            cu_k[i + 1] = cu_k[i] + n
        host = torch.from_numpy(idx[:2 * total + 2 * (len(seqs) + 1)]).pin_memory()
        dev = host.to(self.device, non_blocking=True)
        positions = dev[:total]
        slot_mapping = dev[total:2 * total].to(torch.int32)
        cuq = dev[2 * total:2 * total + len(seqs) + 1].to(torch.int32)
        cuk = dev[2 * total + len(seqs) + 1:].to(torch.int32)
        block_tables = None
        if cu_k[-1] > cu_q[-1]:                                              # prefix cache hit (never, see block manager)
            block_tables = self.prepare_block_tables(seqs)
        set_context(True, cuq, cuk, max(lens), max(len(s) for s in seqs), slot_mapping, None, block_tables)
        embeds = torch.cat([s.input_embeds.reshape(-1, s.input_embeds.shape[-1])[s.num_cached_tokens:]
                            for s in seqs], 0).to(self.device, BF16, non_blocking=True)
        return embeds, positions

    def prepare_block_tables(self, seqs: list[Sequence]) -> torch.Tensor:
        w = max(len(s.block_table) for s in seqs)
        t = np.full((len(seqs), w), -1, dtype=np.int32)
        for i, s in enumerate(seqs):
            t[i, :len(s.block_table)] = s.block_table
        return torch.from_numpy(t).pin_memory().to(self.device, non_blocking=True)

    def decode_position(self, seq: Sequence) -> int:
        return len(seq) - 1                                                  # base.py:225

    def fill_decode_layout(self, seqs: list[Sequence], padded: int):
        """Write the decode layout of ``seqs`` into the staging buffer (rows >= len(seqs) up to
        ``padded`` are graph padding: slot -1, context 0) and upload it."""
        h, bs, n = self.layout.begin(), self.block_size, len(seqs)
        # one bulk numpy assignment per field (a per-element store costs ~100 ns each: 7 x 64 of them per step)
        lens = [s.num_tokens for s in seqs]
        h["positions"][:n] = [self.decode_position(s) for s in seqs]
        h["context_lens"][:n] = lens
        h["slot_mapping"][:n] = [s.block_table[-1] * bs + s.last_block_num_tokens - 1 for s in seqs]    # base.py:227
        h["temperatures"][:n] = [s.temperature for s in seqs]
        h["gen_steps"][:n] = [s.generation_steps for s in seqs]
        h["last_tokens"][:n] = [s.last_token if s.last_token is not None else 0 for s in seqs]
        bt = h["block_tables"]
        for i, s in enumerate(seqs):
            nb = len(s.block_table)
            if nb > self.max_blocks:
                raise RuntimeError(f"sequence of {lens[i]} tokens exceeds max_model_len={self.config.max_model_len}")
            if nb == 1:
                bt[i, 0] = s.block_table[0]
            else:
                bt[i, :nb] = s.block_table
        if padded > n:
            h["positions"][n:padded] = 0
            h["context_lens"][n:padded] = 0
            h["slot_mapping"][n:padded] = -1
            h["gen_steps"][n:padded] = 1
            h["last_tokens"][n:padded] = 0
            h["temperatures"][n:padded] = 1.0
        self.layout.upload()

    def set_decode_context(self, rows: int):
        d = self.layout.d
        set_context(False, slot_mapping=d["slot_mapping"][:rows], context_lens=d["context_lens"][:rows],
                    block_tables=d["block_tables"][:rows], max_context_len=self.decode_max_context())

    def decode_max_context(self) -> int:
        """Upper bound of a decode context; only sizes the split-KV grid (any length stays correct)."""
        return self.max_blocks * self.block_size

    def pick_graph(self, n: int) -> int | None:
        if self.enforce_eager:
            return None
        for bs in self.graph_bs:
            if bs >= n:
                return bs
        return None

    # ------------------------------------------------------------------ graphs
    def decode_body(self, rows: int) -> torch.Tensor:
        """The captured region: static inputs -> hidden states [rows, H]."""
        raise NotImplementedError

    @torch.inference_mode()
    def capture_cudagraph(self):
        self.graph_bs = [b for b in self.graph_batch_sizes() if b <= self.max_bs]
        self.graph_pool = None
        self.graph_out: dict[int, torch.Tensor] = {}
        # a warm-up pass outside capture (lazy cuBLAS handles, kernel attribute setup)
        self.set_decode_context(self.graph_bs[0])
        self.decode_body(self.graph_bs[0])
        torch.cuda.synchronize()
        for bs in reversed(self.graph_bs):
            g = torch.cuda.CUDAGraph()
            self.set_decode_context(bs)
            self.decode_body(bs)                                             # warm-up at this size
            with torch.cuda.graph(g, self.graph_pool):
                self.graph_out[bs] = self.decode_body(bs)
            if self.graph_pool is None:
                self.graph_pool = g.pool()
            self.graphs[bs] = g
        torch.cuda.synchronize()
        reset_context()

    def prepare_sample(self, seqs: list[Sequence]) -> torch.Tensor:
        t = torch.tensor([s.temperature for s in seqs], dtype=torch.float32).pin_memory()
        return t.to(self.device, non_blocking=True)
