"""BASELINE config 2: batched MDCT / IMDCT analysis-synthesis, N=2048, synthetic stereo frames."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_codec_b200 import _lib, mdct

def run(F=1 << 20, dtype=torch.float32, window=_lib.W_SINE, iters=5):
    g = torch.Generator(device="cuda").manual_seed(422)
    x = (0.25 * torch.randn((2, (F + 1) * 1024), generator=g, device="cuda", dtype=torch.float32)).clamp_(-1, 1).to(dtype)
    ctx = _lib.default_context()
    ctx.set_option("variant", int(os.environ.get("PAC_BENCH_VARIANT", 0)))
    X = mdct.mdct_frames(x, 2048, window, ctx); y = mdct.imdct_ola(X, 2048, window, ctx)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tf = ti = 0.0
    for _ in range(iters):
        e[0].record(); X = mdct.mdct_frames(x, 2048, window, ctx); e[1].record(); y = mdct.imdct_ola(X, 2048, window, ctx); e[2].record()
        torch.cuda.synchronize()
        tf += e[0].elapsed_time(e[1]); ti += e[1].elapsed_time(e[2])
    tf /= iters; ti /= iters
    b = x.element_size()
    algo = 2 * F * 1024 * b * 2            # per direction: read 1024 + write 1024 values per channel-frame
    err = (y[:, 1024:-1024] - x[:, 1024:-1024]).abs().max().item()
    return dict(F=F, dtype=str(dtype), fwd_ms=tf, inv_ms=ti, fwd_gbs=algo / tf / 1e6, inv_gbs=algo / ti / 1e6, tdac_err=err)

if __name__ == "__main__":
    F = int(os.environ.get("MDCT_F", 1 << 20))
    for dt in (torch.float32, torch.float64):
        print(json.dumps(run(F if dt == torch.float32 else F // 2, dt, iters=int(os.environ.get("MDCT_ITERS", 5)))))
