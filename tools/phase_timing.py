"""Debug: per-phase cycle shares of encode_frame (needs a -DPAC_PHASE_TIMING build).
   PAC_CFLAGS=-DPAC_PHASE_TIMING python -m audio_codec_b200.build --force && python tools/phase_timing.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from audio_codec_b200 import _lib, engine
S, secs = 32, 5.0
n = int(secs * 48000); nb = -(-n // 1024)
pcm = bench.synth(S, n, 422, "cuda")
d = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device="cuda"); d[:, :, :n] = pcm
ctx = _lib.Context(48000, 2, 128)
buf = (C.c_ulonglong * 32)()
names = ["mdct", "hannfft+scale", "band_smr", "bitalloc", "setup", "pvq plan", "pvq leaves", "header(t0)", "merge", "pick_maskers",
         "masked_threshold", "prologue(load)"]
for it in range(2):
    _lib.lib.pac_debug_phases(buf)
    engine.encode_device(ctx, d.reshape(-1), [nb] * S, [n] * S)
    torch.cuda.synchronize()
_lib.lib.pac_debug_phases(buf)
v = np.array(buf[:], dtype=np.float64)
for base, tag in ((0, "long"), (16, "short")):
    tot = v[base:base + 16].sum()
    print(tag, "total Mcycles %.1f" % (tot / 1e6), {nm: round(v[base + i] / tot, 3) for i, nm in enumerate(names)})

# decoder
d_out, off, _ = engine.encode_device(ctx, d.reshape(-1), [nb] * S, [n] * S)
dnames = ["header(t0)", "A parse", "B leaves", "C recombine", "sbr"]
for it in range(2):
    _lib.lib.pac_debug_phases_dec(buf)
    engine.decode_device(ctx, d_out, off)
    torch.cuda.synchronize()
_lib.lib.pac_debug_phases_dec(buf)
v = np.array(buf[:], dtype=np.float64)
for base, tag in ((0, "dec long"), (16, "dec short")):
    tot = v[base:base + 16].sum()
    print(tag, "total Mcycles %.1f (front only; the IMDCT tail is not counted)" % (tot / 1e6), {nm: round(v[base + i] / tot, 3) for i, nm in enumerate(dnames)})
