"""Deterministic edge-case signals shared by tests/test_edge_gpu.py and the golden generator
tests/golden/make_edge_golden.py (which runs the unmodified reference on them)."""
import numpy as np


def sig(n, fs, seed, C=2):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / fs
    x = 0.3 * np.sin(2 * np.pi * 440 * t) + 0.1 * np.sin(2 * np.pi * 3520 * t + 1) + 0.02 * rng.standard_normal(n)
    pcm = np.stack([x * (1.0 - 0.3 * c) + 0.004 * rng.standard_normal(n) for c in range(C)])
    return np.clip(np.round(pcm * 32767), -32768, 32767).astype(np.int16)


def cases():
    """name -> (fs, int16 [C, n])."""
    n = 4096
    sq = np.where((np.arange(n) // 37) % 2 == 0, 32767, -32768).astype(np.int16)
    tr = sig(9000, 48000, 7)
    tr[:, ::1500] = 30000
    tr[:, 1::1500] = -30000
    return {
        "silence": (48000, np.zeros((2, n), np.int16)),
        "dc": (48000, np.full((2, n), 12345, np.int16)),
        "square": (48000, np.stack([sq, (-sq.astype(np.int32) - 1).astype(np.int16)])),
        "halfsilent": (48000, np.stack([np.zeros(n, np.int16), sig(n, 48000, 3)[0]])),
        "tiny1": (48000, sig(1, 48000, 11)),
        "tiny1023": (48000, sig(1023, 48000, 12)),
        "tiny1025": (48000, sig(1025, 48000, 13)),
        "transients": (48000, tr),
        "mono44": (44100, sig(5000, 44100, 5, C=1)),
    }
