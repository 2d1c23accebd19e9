"""Engine base (reference engine/llm_engine/base.py:15-66): builds ``Config`` from the keyword
arguments it recognises, owns one model runner and one scheduler, exposes ``step`` / ``is_finished``.
Replicas only: no tensor-parallel rank processes are spawned."""
from __future__ import annotations

import atexit
import os
from dataclasses import fields

import torch

from ...config import Config, ModelConfig
from ...sampling_params import SamplingParams
from ..sequence import Sequence


class LLMEngine:
    runner_cls = None
    scheduler_cls = None

    def __init__(self, model: str | ModelConfig, state_dict: dict | None = None, **kwargs):
        names = {f.name for f in fields(Config)}
        cfg_kwargs = {k: v for k, v in kwargs.items() if k in names}          # llm_engine/base.py:18-20
        if isinstance(model, ModelConfig):
            config = Config(model="", hf_config=model, **cfg_kwargs)
        else:
            config = Config(model, **cfg_kwargs)
        self.config = config
        if config.eos < 0 and config.model and os.path.exists(os.path.join(config.model, "tokenizer_config.json")):
            from transformers import AutoTokenizer
            self.tokenizer = AutoTokenizer.from_pretrained(config.model, use_fast=True)
            config.eos = self.tokenizer.eos_token_id                           # llm_engine/base.py:32-33
        self.model_runner = self.runner_cls(config, 0, None, state_dict=state_dict)
        self.scheduler = self.scheduler_cls(config)        # after the runner set config.num_kvcache_blocks
        atexit.register(self.exit)

    def exit(self):
        atexit.unregister(self.exit)
        if getattr(self, "model_runner", None) is not None:
            self.model_runner.call("exit")
            self.model_runner = None

    def _to_device(self, t: torch.Tensor) -> torch.Tensor:
        return t.to(self.model_runner.device, torch.bfloat16, non_blocking=True)

    def is_finished(self):
        return self.scheduler.is_finished()

    def clear_request(self, request_id: str):
        self.scheduler.clear_request(request_id)
