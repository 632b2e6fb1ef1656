"""vibevoice_b200 — B200-native (sm_100a) drop-in for VibeVoice's per-frame acoustic path.

Host side is Python/PyTorch plumbing over the C-ABI library ``libvvb200.so``
(``include/vvb200.h``).  Importing this package does not load the library; the first
operator call does, and raises if it is missing (no fallback).
"""
__version__ = "0.1.0"
