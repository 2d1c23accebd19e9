"""CPU oracle for the Qwen3-TTS decode hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this
directory: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` do, and there only
as the checker / the timed CPU baseline.

Every function restates, op by op and rounding point by rounding point, what
the reference (tsdocode/nano-qwen3tts-vllm, mounted read-only at
``/root/reference`` while this was written) computes on the path named by
``BASELINE.json:north_star``.  Citations are ``file:line`` relative to
``/root/reference/nano-qwen3tts-vllm/``.

Parity pinning: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against the *live reference
modules* imported on CPU in the build container by
``oracle/pin_against_reference.py``; the resulting vectors are committed under
``tests/golden/``.  Two reference ops cannot execute without a GPU
(``store_kvcache`` is a Triton kernel, attention is flash-attn): for those the
oracle is a restatement of ``layers/attention.py:10-30,73-96`` only and the
header of ``oracle/ops.py`` says so ("parity unpinned" for those two; on the
GPU box the tests additionally cross-check against the flash-attn wheel that
the reference itself calls).
"""
