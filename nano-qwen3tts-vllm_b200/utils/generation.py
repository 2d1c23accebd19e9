"""Prompt-prep embedding assembly on the GPU: the talker's input embeddings, the
trailing text hiddens, the tts pad row and the attention mask, with the reference's call surface
(``prepare_inputs`` / ``generate_icl_prompt`` / ``generate_speaker_prompt``, utils/generation.py:18-379, 412-515,
380-409).

The reference builds each request with ~10 small embedding / projection / cat / add calls; a batch of arrivals (or one
4k-token voice-clone prompt) is therefore hundreds of tiny launches.  Here the whole batch is PLANNED on the host as a
row table - every prompt row is ``text side (+) codec side`` - and executed in a fixed number of launches whatever
the batch size:

  1. one ``text_projection(text_embedding(ids))`` over every text token of every request (the three tts specials
     first) - the projection MLP runs on the native weight-streaming GEMM;
  2. one gather of every codec-tag row, one 16-codebook gather-sum over all ICL reference frames
     (``b2tts_codec_embed_sum``), one copy of the external speaker vectors - all into one codec-side table;
  3. ONE ``b2tts_prompt_assemble`` launch that gathers, adds (bf16, as torch's tensor add) and scatters the rows of all
     prompts and all trailing hiddens into the left-padded / pad-row-padded batch tensors.

There is no CPU path: the callables must produce CUDA tensors."""
from __future__ import annotations

from typing import Callable, List, Optional, Union

import numpy as np
import torch

from .. import _native as nat


class _Plan:
    """Row table of one batch.  Text rows index the projected ids (0 bos, 1 eos, 2 pad, then the requests' tokens);
    codec rows index a table built from codec-token gathers, external vectors and ICL reference-frame sums."""

    def __init__(self, config):
        self.text_ids: list[int] = [config.tts_bos_token_id, config.tts_eos_token_id, config.tts_pad_token_id]
        self.codec_ids: list[int] = []
        self.vectors: list[torch.Tensor] = []          # external rows (speaker x-vectors, rows from a caller's ICL hook)
        self.ref_codes: list[torch.Tensor] = []        # [T, num_code_groups] int64 per ICL request
        self.ref_rows = 0
        self._codec_slot: dict[int, int] = {}
        self.prompts: list[tuple[list[int], list[int]]] = []      # per request (a_idx, b_idx), symbolic b refs
        self.trails: list[tuple[list[int], list[int]]] = []

    BOS, EOS, PAD = 0, 1, 2

    def text(self, ids) -> list[int]:
        start = len(self.text_ids)
        self.text_ids.extend(int(t) for t in ids)
        return list(range(start, len(self.text_ids)))

    def codec(self, token: int):
        slot = self._codec_slot.get(int(token))
        if slot is None:
            slot = self._codec_slot[int(token)] = len(self.codec_ids)
            self.codec_ids.append(int(token))
        return ("c", slot)

    def vector(self, rows: torch.Tensor) -> list:
        rows = rows.reshape(-1, rows.shape[-1])
        start = sum(v.shape[0] for v in self.vectors)
        self.vectors.append(rows)
        return [("v", start + i) for i in range(rows.shape[0])]

    def frames(self, codes: torch.Tensor) -> list:
        start = self.ref_rows
        self.ref_codes.append(codes)
        self.ref_rows += codes.shape[0]
        return [("r", start + i) for i in range(codes.shape[0])]

    def resolve(self, ref) -> int:
        if ref is None:
            return -1
        kind, i = ref
        return i if kind == "c" else len(self.codec_ids) + (i if kind == "v" else sum(v.shape[0] for v in self.vectors) + i)


def _host_ids(t: Optional[torch.Tensor]) -> list[int]:
    return [] if t is None else t.reshape(-1).tolist()


def _plan_request(plan: _Plan, config, input_id, instruct_id, ref_id, language, speaker, clone, spk_vector, ref_code, icl,
                  non_streaming, body_from_hook=False):
    """Rows of one request (generation.py:72-88 instruct, :96-230 tags and role, :232-320 body / ICL).
    ``body_from_hook``: only the instruct / role / tag rows are planned; the caller appends the ICL rows."""
    tc = config.talker_config
    ids = _host_ids(input_id)
    a: list[int] = []
    b: list = []

    def emit(text_rows, codec_refs):
        a.extend(text_rows)
        b.extend(codec_refs)

    if instruct_id is not None:
        rows = plan.text(_host_ids(instruct_id))
        emit(rows, [None] * len(rows))
    spk = None
    if clone:
        if spk_vector is not None:
            spk = plan.vector(spk_vector)[0]
    elif speaker:
        if speaker.lower() not in tc.spk_id:
            raise NotImplementedError(f"Speaker {speaker} not implemented")
        spk = plan.codec(tc.spk_id[speaker.lower()])
    if language.lower() == "auto":
        language_id = None
    else:
        if language.lower() not in tc.codec_language_id:
            raise NotImplementedError(f"Language {language} not implemented")
        language_id = tc.codec_language_id[language.lower()]
    if language.lower() in ("chinese", "auto") and speaker and tc.spk_is_dialect.get(speaker.lower(), False) is not False:
        language_id = tc.codec_language_id[tc.spk_is_dialect[speaker.lower()]]
    think = [tc.codec_nothink_id, tc.codec_think_bos_id, tc.codec_think_eos_id] if language_id is None else \
        [tc.codec_think_id, tc.codec_think_bos_id, language_id, tc.codec_think_eos_id]
    tags = [plan.codec(t) for t in think] + ([spk] if spk is not None else []) + [plan.codec(tc.codec_pad_id),
                                                                                 plan.codec(tc.codec_bos_id)]
    emit(plan.text(ids[:3]), [None] * 3)                                           # the role tokens, text only
    emit([plan.PAD] * (len(tags) - 2) + [plan.BOS], tags[:-1])                     # tts pad ... pad, bos over the tags
    if body_from_hook:
        trail = ([], [])
    elif icl:
        if ref_id is None:
            raise ValueError("ICL mode requires ref_ids (the reference transcript) for every ICL request")
        text = plan.text(_host_ids(ref_id)[3:-2] + ids[3:-5]) + [plan.EOS]
        codec = [plan.codec(tc.codec_bos_id)] + plan.frames(ref_code)
        if non_streaming:
            emit(text, [plan.codec(tc.codec_pad_id)] * len(text))
            emit([plan.PAD] * len(codec), codec)
            trail = ([plan.PAD], [None])
        elif len(text) > len(codec):
            emit(text[:len(codec)], codec)
            rest = text[len(codec):]
            trail = (rest, [None] * len(rest))
        else:
            emit(text + [plan.PAD] * (len(codec) - len(text)), codec)
            trail = ([plan.PAD], [None])
    elif non_streaming:
        body = plan.text(ids[3:-5]) + [plan.EOS]
        emit(body, [plan.codec(tc.codec_pad_id)] * len(body))
        emit([plan.PAD], [plan.codec(tc.codec_bos_id)])
        trail = ([plan.PAD], [None])
    else:
        emit(plan.text(ids[3:4]), [tags[-1]])
        rest = plan.text(ids[4:-5]) + [plan.EOS]
        trail = (rest, [None] * len(rest))
    plan.prompts.append((a, b))
    plan.trails.append(trail)


def _embedding_table(emb) -> Optional[torch.Tensor]:
    w = getattr(emb, "weight", None)
    return w if isinstance(w, torch.Tensor) and w.dim() == 2 else None


def _embed(emb: Callable, ids: torch.Tensor) -> torch.Tensor:
    """Rows of an embedding for a flat id vector: the native gather when ``emb`` exposes a bf16 CUDA table."""
    w = _embedding_table(emb)
    if w is not None and w.is_cuda and w.dtype == torch.bfloat16 and w.is_contiguous():
        return nat.gather_rows(w, ids)
    return emb(ids.view(1, -1)).reshape(ids.numel(), -1)


def prepare_inputs(config, input_ids: list[torch.Tensor], instruct_ids: Optional[list] = None,
                   ref_ids: Optional[list] = None, voice_clone_prompt: Optional[dict] = None,
                   languages: Optional[list[str]] = None, speakers: Optional[list] = None,
                   non_streaming_mode: bool = False, *, text_embedding: Callable, input_embedding: Callable,
                   text_projection: Callable, device: Union[torch.device, str],
                   voice_clone_spk_embeds: Optional[list[torch.Tensor]] = None,
                   generate_speaker_prompt_fn: Optional[Callable] = None, generate_icl_prompt_fn: Optional[Callable] = None,
                   code_predictor_embeddings=None):
    """Same arguments and return value as the reference (utils/generation.py:18-379):
    ``(talker_input_embeds [B, Tmax, H] left-padded with zeros, trailing_text_hiddens [B, Umax, H] padded with the tts
    pad row, tts_pad_embed [1, 1, H], attention_mask [B, Tmax] int64)``.

    ``code_predictor_embeddings`` (extra): the 15 codebook embeddings (modules, or one stacked ``[15, V, H]`` tensor) used
    by ICL requests; when it is absent the reference's ``generate_icl_prompt_fn`` hook is called per ICL request and
    its rows enter the plan as ready-made rows."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("libb200tts: prepare_inputs needs a CUDA device (there is no CPU path)")
    n = len(input_ids)
    if voice_clone_prompt is not None and voice_clone_spk_embeds is None:
        if generate_speaker_prompt_fn is None:
            raise ValueError("voice_clone_prompt is set but neither voice_clone_spk_embeds nor "
                             "generate_speaker_prompt_fn was provided")
        voice_clone_spk_embeds = generate_speaker_prompt_fn(voice_clone_prompt)
    speakers = speakers if speakers is not None else [None] * n
    languages = languages if languages is not None else ["auto"] * n
    tc = config.talker_config
    have_books = code_predictor_embeddings is not None
    plan = _Plan(config)
    hooked: dict[int, tuple] = {}
    for i in range(n):
        vec, code, icl = None, None, False
        if voice_clone_prompt is not None:
            icl = bool(voice_clone_prompt["icl_mode"][i]) and voice_clone_prompt.get("ref_code") is not None
            if voice_clone_prompt["x_vector_only_mode"][i] or voice_clone_prompt["icl_mode"][i]:
                vec = voice_clone_spk_embeds[i]
            if icl:
                code = voice_clone_prompt["ref_code"][i]
                if ref_ids is None or ref_ids[i] is None:
                    raise ValueError(f"ICL mode requires ref_ids, but ref_ids[{i}] is None")
        assert languages[i] is not None
        if icl and not have_books:
            hooked[i] = (code,)
        _plan_request(plan, config, input_ids[i], None if instruct_ids is None else instruct_ids[i],
                      None if ref_ids is None else ref_ids[i], languages[i], speakers[i], voice_clone_prompt is not None,
                      vec, code, icl, non_streaming_mode, body_from_hook=icl and not have_books)

    # ---- 1. every text token of the batch through embedding + projection, once
    ids = torch.tensor(plan.text_ids, dtype=torch.int64).pin_memory().to(device, non_blocking=True)
    text_rows = text_projection(_embed(text_embedding, ids))
    text_rows = text_rows.reshape(ids.numel(), -1)
    hidden = text_rows.shape[1]
    if text_rows.dtype != torch.bfloat16:
        raise RuntimeError(f"libb200tts: prepare_inputs expects bf16 embeddings, got {text_rows.dtype}")

    # ---- requests whose ICL rows come from the caller's hook (no codebook tables given): append them as ready rows
    if hooked:
        if generate_icl_prompt_fn is None:
            raise ValueError("ICL mode requested but neither code_predictor_embeddings nor generate_icl_prompt_fn was provided")
        for i, (code,) in hooked.items():
            extra, trailing = generate_icl_prompt_fn(
                text_id=input_ids[i][:, 3:-5], ref_id=ref_ids[i][:, 3:-2], ref_code=code.to(device),
                tts_pad_embed=text_rows[2].view(1, 1, -1), tts_eos_embed=text_rows[1].view(1, 1, -1),
                non_streaming_mode=non_streaming_mode)
            a, b = plan.prompts[i]
            rows = plan.vector(extra.to(device))
            a.extend([-1] * len(rows)); b.extend(rows)
            rows = plan.vector(trailing.to(device))
            plan.trails[i] = ([-1] * len(rows), rows)

    # ---- 2. the codec-side table: tag rows | external vectors | ICL reference-frame sums
    n_codec, n_vec = len(plan.codec_ids), sum(v.shape[0] for v in plan.vectors)
    table = torch.empty(max(n_codec + n_vec + plan.ref_rows, 1), hidden, dtype=torch.bfloat16, device=device)
    if n_codec:
        cid = torch.tensor(plan.codec_ids, dtype=torch.int64).pin_memory().to(device, non_blocking=True)
        w = _embedding_table(input_embedding)
        if w is not None and w.is_cuda and w.dtype == torch.bfloat16 and w.is_contiguous():
            nat.gather_rows(w, cid, out=table[:n_codec])
        else:
            table[:n_codec] = input_embedding(cid.view(1, -1)).reshape(n_codec, hidden)
    o = n_codec
    for v in plan.vectors:
        table[o:o + v.shape[0]] = v.to(device=device, dtype=torch.bfloat16)
        o += v.shape[0]
    if plan.ref_rows:
        codes = torch.cat([c.reshape(-1, tc.num_code_groups) for c in plan.ref_codes]).to(device=device, dtype=torch.int64)
        books = code_predictor_embeddings
        if not torch.is_tensor(books):
            books = torch.stack([_embedding_table(e) for e in books])
        nat.codec_embed_sum(_embedding_table(input_embedding), books, codes.contiguous(), out=table[o:o + plan.ref_rows])

    # ---- 3. one launch writes every prompt row (left-padded) and every trailing row (pad-row padded)
    tmax = max(len(a) for a, _ in plan.prompts)
    umax = max(len(a) for a, _ in plan.trails)
    a_idx = np.full(n * (tmax + umax), -1, dtype=np.int32)
    b_idx = np.full(n * (tmax + umax), -1, dtype=np.int32)
    mask = np.zeros((n, tmax), dtype=np.int64)
    for i, (a, b) in enumerate(plan.prompts):
        lo = i * tmax + tmax - len(a)
        a_idx[lo:lo + len(a)] = a
        b_idx[lo:lo + len(a)] = [plan.resolve(r) for r in b]
        mask[i, tmax - len(a):] = 1
    base = n * tmax
    for i, (a, b) in enumerate(plan.trails):
        lo = base + i * umax
        a_idx[lo:lo + umax] = plan.PAD
        a_idx[lo:lo + len(a)] = a
        b_idx[lo:lo + len(a)] = [plan.resolve(r) for r in b]
    idx = torch.from_numpy(np.stack([a_idx, b_idx])).pin_memory().to(device, non_blocking=True)
    out = torch.empty(n * (tmax + umax), hidden, dtype=torch.bfloat16, device=device)
    nat.prompt_assemble(text_rows, table, idx[0], idx[1], out)
    prompts = out[:base].view(n, tmax, hidden)
    trailing = out[base:].view(n, umax, hidden)
    return prompts, trailing, text_rows[2].view(1, 1, hidden), torch.from_numpy(mask).to(device)


def generate_speaker_prompt(voice_clone_prompt: dict, device: Union[torch.device, str],
                            dtype: Optional[torch.dtype] = None) -> List[torch.Tensor]:
    """The reference speaker vectors of a voice-clone prompt on ``device`` (utils/generation.py:380-409)."""
    out = []
    for emb in voice_clone_prompt["ref_spk_embedding"]:
        emb = emb.to(device)
        out.append(emb if dtype is None else emb.to(dtype))
    return out


def generate_icl_prompt(text_id, ref_id, ref_code, tts_pad_embed, tts_eos_embed, non_streaming_mode, config,
                        text_embedding, input_embedding, text_projection, code_predictor_embeddings, device):
    """One ICL prompt (utils/generation.py:412-515): ``(icl_input_embed [1, T, H], trailing_text_hidden [1, U, H])``.
    The standalone form of what ``prepare_inputs`` plans for ICL requests; same three launches."""
    device = torch.device(device)
    tc = config.talker_config
    plan = _Plan(config)
    text = plan.text(_host_ids(ref_id) + _host_ids(text_id))
    ids = torch.tensor(plan.text_ids[3:], dtype=torch.int64).pin_memory().to(device, non_blocking=True)
    rows = text_projection(_embed(text_embedding, ids)).reshape(ids.numel(), -1)
    hidden = rows.shape[1]
    # text table: projected tokens | eos | pad
    ttab = torch.cat([rows, tts_eos_embed.reshape(1, hidden).to(device), tts_pad_embed.reshape(1, hidden).to(device)])
    eos, pad = rows.shape[0], rows.shape[0] + 1
    text = [t - 3 for t in text] + [eos]
    t_code = ref_code.shape[0]
    ctab = torch.empty(2 + t_code, hidden, dtype=torch.bfloat16, device=device)
    cid = torch.tensor([tc.codec_bos_id, tc.codec_pad_id], dtype=torch.int64).pin_memory().to(device, non_blocking=True)
    ctab[:2] = _embed(input_embedding, cid)
    books = code_predictor_embeddings
    if not torch.is_tensor(books):
        books = torch.stack([_embedding_table(e) for e in books])
    nat.codec_embed_sum(_embedding_table(input_embedding), books,
                        ref_code.reshape(t_code, tc.num_code_groups).to(device=device, dtype=torch.int64).contiguous(),
                        out=ctab[2:])
    codec = [0] + list(range(2, 2 + t_code))
    if non_streaming_mode:
        a, b, trail = text + [pad] * len(codec), [1] * len(text) + codec, [pad]
    elif len(text) > len(codec):
        a, b, trail = text[:len(codec)], codec, text[len(codec):]
    else:
        a, b, trail = text + [pad] * (len(codec) - len(text)), codec, [pad]
    idx = torch.tensor([a + trail, b + [-1] * len(trail)], dtype=torch.int32).pin_memory().to(device, non_blocking=True)
    out = torch.empty(len(a) + len(trail), hidden, dtype=torch.bfloat16, device=device)
    nat.prompt_assemble(ttab, ctab, idx[0], idx[1], out)
    return out[:len(a)].unsqueeze(0), out[len(a):].unsqueeze(0)
