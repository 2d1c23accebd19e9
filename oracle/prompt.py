"""CPU oracle of the prompt-prep embedding assembly.

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

Restates ``utils/generation.py:18-379`` (``prepare_inputs``) and ``:412-515`` (``generate_icl_prompt``) of the
reference as a ROW TABLE: every row of the talker's prompt is ``text side (+) codec side`` where the text side is a
row of ``text_projection(text_embedding(ids))`` (or one of the three projected tts specials) and the codec side a
row of the talker's codec embedding, an external speaker vector, or the 16-codebook sum of one reference frame.
The adds are bf16 tensor adds, the calls into the three callables are made on the same id slices the reference
uses (the CPU GEMM is not guaranteed row-batch invariant), and ``oracle/pin_against_reference.py`` asserts
bit-equality with the live reference functions for every mode (plain, instruct, speaker, language / dialect,
x-vector, ICL streaming with both length orders, ICL non-streaming, batches of mixed length).
"""
from __future__ import annotations

import torch


def _ids(x: torch.Tensor) -> torch.Tensor:
    return x if x.dim() == 2 else x.unsqueeze(0)


def icl_rows(text_id, ref_id, ref_code, tts_pad, tts_eos, non_streaming, tc, text_embedding, input_embedding,
             text_projection, predictor_embeddings):
    """generation.py:412-515.  Returns (icl rows [1, T, D], trailing [1, T', D])."""
    text = torch.cat([text_projection(text_embedding(torch.cat([ref_id, text_id], dim=-1))), tts_eos], dim=1)   # :452-455
    per_book = []
    for i in range(tc.num_code_groups):                                                                           # :461-473
        emb = input_embedding if (i == 0 or i - 1 >= len(predictor_embeddings)) else predictor_embeddings[i - 1]
        per_book.append(emb(ref_code[:, i:i + 1] if i else ref_code[:, :1]))
    frames = torch.cat(per_book, dim=1).sum(1).unsqueeze(0)                                                      # :477
    bos = input_embedding(torch.tensor([[tc.codec_bos_id]], dtype=text_id.dtype))                                # :480-486
    codec = torch.cat([bos, frames], dim=1)
    nt, ncd = text.shape[1], codec.shape[1]
    if non_streaming:                                                                                             # :493-503
        pad = input_embedding(torch.tensor([[tc.codec_pad_id] * nt], dtype=text_id.dtype))
        return torch.cat([text + pad, codec + tts_pad], dim=1), tts_pad
    if nt > ncd:                                                                                                  # :506-507
        return text[:, :ncd] + codec, text[:, ncd:]
    text = torch.cat([text] + [tts_pad] * (ncd - nt), dim=1)                                                      # :509-510
    return text + codec, tts_pad


def request_rows(config, input_id, instruct_id, ref_id, language, speaker, clone, spk_vector, ref_code, icl, non_streaming,
                 text_embedding, input_embedding, text_projection, predictor_embeddings):
    """One request of generation.py:96-300.  Returns (prompt rows [1, T, D], trailing [1, T', D], tts_pad [1, 1, D])."""
    tc = config.talker_config
    dt = input_id.dtype
    parts = []
    if instruct_id is not None:                                                                                   # :72-88
        parts.append(text_projection(text_embedding(_ids(instruct_id))))
    if clone:                                                                                                     # :118-123
        speaker_embed = spk_vector                    # None unless x-vector / ICL mode
    elif speaker:                                                                                                 # :101-117
        if speaker.lower() not in tc.spk_id:
            raise NotImplementedError(f"Speaker {speaker} not implemented")
        speaker_embed = input_embedding(torch.tensor(tc.spk_id[speaker.lower()], dtype=dt))
    else:
        speaker_embed = None
    if language.lower() == "auto":                                                                                # :126-136
        language_id = None
    else:
        if language.lower() not in tc.codec_language_id:
            raise NotImplementedError(f"Language {language} not implemented")
        language_id = tc.codec_language_id[language.lower()]
    if language.lower() in ("chinese", "auto") and speaker and tc.spk_is_dialect.get(speaker.lower(), False) is not False:
        language_id = tc.codec_language_id[tc.spk_is_dialect[speaker.lower()]]                                    # :138-145
    special = text_projection(text_embedding(torch.tensor(
        [[config.tts_bos_token_id, config.tts_eos_token_id, config.tts_pad_token_id]], dtype=dt)))               # :147-161
    tts_bos, tts_eos, tts_pad = special.chunk(3, dim=1)
    think = [tc.codec_nothink_id, tc.codec_think_bos_id, tc.codec_think_eos_id] if language_id is None else \
        [tc.codec_think_id, tc.codec_think_bos_id, language_id, tc.codec_think_eos_id]                          # :164-181
    tags = [input_embedding(torch.tensor([think], dtype=dt))]
    if speaker_embed is not None:
        tags.append(speaker_embed.view(1, 1, -1))                                                                 # :204-212
    tags.append(input_embedding(torch.tensor([[tc.codec_pad_id, tc.codec_bos_id]], dtype=dt)))                    # :189-199
    codec = torch.cat(tags, dim=1)
    role = text_projection(text_embedding(input_id[:, :3]))                                                       # :214-216
    prefix = torch.cat((tts_pad.expand(-1, codec.shape[1] - 2, -1), tts_bos), dim=1) + codec[:, :-1]             # :218-226
    rows = torch.cat((role, prefix), dim=1)
    if icl:                                                                                                       # :232-258
        extra, trailing = icl_rows(input_id[:, 3:-5], ref_id[:, 3:-2], ref_code, tts_pad, tts_eos, non_streaming, tc,
                                   text_embedding, input_embedding, text_projection, predictor_embeddings)
        rows = torch.cat([rows, extra], dim=1)
    else:
        rows = torch.cat([rows, text_projection(text_embedding(input_id[:, 3:4])) + codec[:, -1:]], dim=1)       # :260-268
        if non_streaming:                                                                                         # :269-309
            body = input_id[:, 3:-5]
            mid = torch.cat((text_projection(text_embedding(body)), tts_eos), dim=1) + input_embedding(
                torch.tensor([[tc.codec_pad_id] * (body.shape[1] + 1)], dtype=dt))
            last = tts_pad + input_embedding(torch.tensor([[tc.codec_bos_id]], dtype=dt))
            rows = torch.cat([rows[:, :-1], mid, last], dim=1)
            trailing = tts_pad
        else:                                                                                                     # :311-320
            trailing = torch.cat((text_projection(text_embedding(input_id[:, 4:-5])), tts_eos), dim=1)
    parts.append(rows)
    return torch.cat(parts, dim=1), trailing, tts_pad


def prepare_inputs(config, input_ids, instruct_ids=None, ref_ids=None, voice_clone_prompt=None, languages=None,
                   speakers=None, non_streaming_mode=False, *, text_embedding, input_embedding, text_projection,
                   predictor_embeddings=()):
    """generation.py:18-379.  Returns (prompts [B, Tmax, D] LEFT-padded with zeros, trailing [B, T'max, D] right-padded
    with the tts pad row, tts_pad [1, 1, D], attention mask [B, Tmax] int64)."""
    n = len(input_ids)
    speakers = speakers or [None] * n
    languages = languages or ["auto"] * n
    prompts, trails, tts_pad = [], [], None
    for i in range(n):
        vec, code, icl = None, None, False
        if voice_clone_prompt is not None:
            icl = bool(voice_clone_prompt["icl_mode"][i]) and voice_clone_prompt.get("ref_code") is not None
            if voice_clone_prompt["x_vector_only_mode"][i] or voice_clone_prompt["icl_mode"][i]:
                vec = voice_clone_prompt["ref_spk_embedding"][i]
            if icl:
                code = voice_clone_prompt["ref_code"][i]
        rows, trailing, tts_pad = request_rows(
            config, input_ids[i], None if instruct_ids is None else instruct_ids[i],
            None if ref_ids is None else ref_ids[i], languages[i], speakers[i], voice_clone_prompt is not None,
            vec, code, icl, non_streaming_mode, text_embedding, input_embedding, text_projection, predictor_embeddings)
        prompts.append(rows.squeeze(0))
        trails.append(trailing.squeeze(0))
    tmax, d = max(p.shape[0] for p in prompts), prompts[0].shape[1]
    batch = torch.zeros(n, tmax, d, dtype=prompts[0].dtype)                                                       # :329-341
    mask = torch.zeros(n, tmax, dtype=torch.int64)
    for i, p in enumerate(prompts):
        batch[i, tmax - p.shape[0]:] = p
        mask[i, tmax - p.shape[0]:] = 1                                                                           # :343-356
    umax = max(t.shape[0] for t in trails)
    trailing = tts_pad.squeeze().expand(n, umax, d).clone()                                                       # :358-377
    for i, t in enumerate(trails):
        trailing[i, :t.shape[0]] = t
    return batch, trailing, tts_pad, mask
