"""Experiment: do two encode+decode pipelines on two CUDA streams (two contexts, half the streams each) overlap
their kernel tails?  Compares with one context running all streams (the bench's device-resident path)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from audio_codec_b200 import _lib, engine
S, secs = 192, 20.0
n = int(secs * 48000); nb = -(-n // 1024)
pcm = bench.synth(S, n, 422, "cuda")
d = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device="cuda"); d[:, :, :n] = pcm
del pcm

def run(ctx, x, k, stream, out):
    with torch.cuda.stream(stream):
        d_out, off, err = engine.encode_device(ctx, x.reshape(-1), [nb] * k, [n] * k)
        d_pcm, po, err2 = engine.decode_device(ctx, d_out, off)
    out.append((int(off[-1]), err, err2))

def timed(parts):
    ctxs = [_lib.Context(48000, 2, 128) for _ in parts]
    streams = [torch.cuda.Stream() for _ in parts]
    best = 1e9
    for it in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = []
        th = [threading.Thread(target=run, args=(ctxs[i], d[a:b], b - a, streams[i], res)) for i, (a, b) in enumerate(parts)]
        for t in th: t.start()
        for t in th: t.join()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if it >= 2: best = min(best, dt)
    return best, res

for parts in ([(0, S)], [(0, S // 2), (S // 2, S)], [(0, S // 3), (S // 3, 2 * S // 3), (2 * S // 3, S)]):
    t, res = timed(parts)
    print(len(parts), "context(s): %.2f ms per step -> %.0f audio-s/s" % (t * 1e3, S * secs / t), res[:1], flush=True)
