"""``SamplingParams`` with the reference's fields and its greedy-is-forbidden assertion
(sampling_params.py:5-14).  ``top_k`` IS honoured here when a runner is built with
``honor_top_k=True`` (the reference accepts and drops it); the default keeps the
reference behaviour: every row is sampled with top_k=50."""
from dataclasses import dataclass


@dataclass
class SamplingParams:
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False
    do_sample: bool = True
    top_k: int = 50
    top_p: float = 1.0

    def __post_init__(self):
        assert self.temperature > 1e-10, "greedy sampling is not permitted"
