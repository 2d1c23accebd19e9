"""B200-native (sm_100a) Qwen3-TTS decode hot path behind the nano-qwen3tts-vllm operator API.

The arithmetic lives in ``lib/libb200tts.so`` (hand-written CUDA, C ABI in ``include/b200tts.h``)
and is reached through ``_native`` only.  ``layers/``, ``models/`` and ``engine/`` mirror the
reference's Python surfaces (same class names, signatures, state-dict keys) so that the
``TalkerLLM`` / ``PredictorLLM`` step API is a drop-in.
"""
__version__ = "0.1.0"
