"""Where the end-to-end path (host buffers through the C ABI) loses time against the device-resident path."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from audio_codec_b200 import _lib, engine
S, secs = 192, 20.0
n = int(secs * 48000); nb = -(-n // 1024)
pcm = bench.synth(S, n, 422, "cuda")
d = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device="cuda"); d[:, :, :n] = pcm
del pcm
ctx = _lib.Context(48000, 2, 128)
if len(sys.argv) > 1: ctx.set_option("host_chunk_mb", float(sys.argv[1]))
nbs, hdr = [nb] * S, [n] * S
h_pcm = d.reshape(-1).cpu().pin_memory()
d_out, off, _ = engine.encode_device(ctx, d.reshape(-1), nbs, hdr)
d_pcm, po, _ = engine.decode_device(ctx, d_out, off)
h_pac = torch.empty(d_out.numel(), dtype=torch.uint8).pin_memory()
h_dec = torch.empty(int(po[-1]), dtype=torch.int16).pin_memory()
def t(fn, k=5):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(k):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))
r = {}
r["enc device"] = t(lambda: engine.encode_device(ctx, d.reshape(-1), nbs, hdr, d_out))
r["dec device"] = t(lambda: engine.decode_device(ctx, d_out, off, d_pcm))
r["enc host"] = t(lambda: engine.encode_host(ctx, h_pcm, nbs, hdr, h_pac))
o, _ = engine.encode_host(ctx, h_pcm, nbs, hdr, h_pac)
r["dec host"] = t(lambda: engine.decode_host(ctx, h_pac, o, h_dec))
dd = d.reshape(-1)
r["h2d pcm alone"] = t(lambda: dd.copy_(h_pcm, non_blocking=True))
r["d2h pcm alone"] = t(lambda: h_dec.copy_(d_pcm[:h_dec.numel()], non_blocking=True))
r["d2h pac alone"] = t(lambda: h_pac[:int(off[-1])].copy_(d_out[:int(off[-1])], non_blocking=True))
for k, v in r.items(): print("%-16s %7.2f ms" % (k, v))
print("enc gap %.2f  dec gap %.2f" % (r["enc host"] - r["enc device"], r["dec host"] - r["dec device"]))
