"""One small encode + decode on cuda:0 (for ncu captures of a single launch per kernel).
usage: small_codec.py [streams] [seconds] [rate]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from audio_codec_b200 import _lib, engine
S = int(sys.argv[1]) if len(sys.argv) > 1 else 48
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 5.0
rate = int(sys.argv[3]) if len(sys.argv) > 3 else 128
n = int(secs * 48000); nb = -(-n // 1024)
pcm = bench.synth(S, n, 422, "cuda")
d = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device="cuda"); d[:, :, :n] = pcm
ctx = _lib.Context(48000, 2, rate)
if os.environ.get("PAC_BENCH_VARIANT"):
    ctx.set_option("variant", int(os.environ["PAC_BENCH_VARIANT"]))
for _ in range(2):
    d_out, off, err = engine.encode_device(ctx, d.reshape(-1), [nb] * S, [n] * S)
    d_pcm, po, err2 = engine.decode_device(ctx, d_out, off)
torch.cuda.synchronize()
print("ok", int(off[-1]), err, err2)
