"""loqa_hub_b200 — B200-native relay-audio ingest for Loqa Hub.

The product is loqa_hub_b200/libgpuaudio.so (hand-written sm_100a CUDA behind the
C ABI of include/gpuaudio.h).  This package only binds it for the Python harness
(tests, bench, smoke).  (The directory is spelled with an underscore because a
hyphen is not a valid Python identifier.)
"""
from . import _lib  # noqa: F401

__all__ = ["_lib", "api", "synth", "sharding"]
