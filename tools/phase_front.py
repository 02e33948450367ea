"""Debug: fine-grained cycle shares of k_encode_front (needs a -DPAC_PHASE_TIMING build, PAC_LIB_NAME=libpac_phase.so)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from audio_codec_b200 import _lib, engine
S, secs = 48, 5.0
n = int(secs * 48000); nb = -(-n // 1024)
pcm = bench.synth(S, n, 422, "cuda")
d = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device="cuda"); d[:, :, :n] = pcm
ctx = _lib.Context(48000, 2, int(sys.argv[1]) if len(sys.argv) > 1 else 128)
if os.environ.get("PAC_BENCH_VARIANT"):
    ctx.set_option("variant", int(os.environ["PAC_BENCH_VARIANT"]))
buf = (C.c_ulonglong * 64)()
names = {0: "fold", 1: "fft512", 2: "post+max", 3: "hann load", 4: "fft1024+mag", 5: "sbr+scale+intensity", 8: "pk ballot+prefix",
         9: "pk compact", 10: "pk log10/atan", 11: "pk blkmax+links", 12: "smr exact mid", 13: "smr bounds", 14: "smr band scan",
         15: "smr candidates", 16: "hand-off"}
for it in range(2):
    _lib.lib.pac_debug_phases_front(buf)
    engine.encode_device(ctx, d.reshape(-1), [nb] * S, [n] * S)
    torch.cuda.synchronize()
_lib.lib.pac_debug_phases_front(buf)
v = np.array(buf[:], dtype=np.float64)
nblk = S * 2 * (nb + 2)
for base, tag in ((0, "long"), (32, "short")):
    tot = v[base:base + 32].sum()
    print(tag, "total Mcycles %.1f  (%.0f cycles per channel-block of %d)" % (tot / 1e6, tot / nblk, nblk))
    for i, nm in names.items():
        print("   %-22s %5.1f %%  %8.0f cyc/blk" % (nm, 100 * v[base + i] / max(tot, 1), v[base + i] / nblk))
