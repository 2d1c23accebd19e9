"""Import shim: the package directory is ``nano-qwen3tts-vllm_b200/`` (hyphens, like the
reference's ``nano-qwen3tts-vllm/`` which is mapped to ``nano_qwen3tts_vllm`` by its
pyproject.toml:23-24).  Importing ``nano_qwen3tts_vllm_b200`` makes this module a package whose
search path is that directory and runs its ``__init__.py``."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "nano-qwen3tts-vllm_b200")]
__file__ = _os.path.join(__path__[0], "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
