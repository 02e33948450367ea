"""bench.py -- encode+decode throughput of the perceptual-coder hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  torchrun --nproc-per-node N bench.py --gpus N ...

Workload (BASELINE.json metric "encoded audio-sec/sec (enc+dec, 128 kb/s)", config[2]
shape, one chunk of the synthetic corpus per step): S synthetic 48 kHz stereo
streams of D seconds each are ENCODED to .pac (transient detection, block
switching, window+MDCT, psychoacoustic SMRs, bit allocation, PVQ quantisation,
bit packing) at 128 kb/s per channel and DECODED back to 16-bit PCM.
One "audio-second" is one second of a stereo stream.  A step is one pass over the
batch; `value` times it with inputs resident in HBM, `e2e` through the C-ABI
host-buffer calls (pinned host PCM in, .pac out, .pac in, PCM out).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FS = 48000
RATE = 128
N_STREAMS = int(os.environ.get("PAC_BENCH_STREAMS", 192))
SECONDS = float(os.environ.get("PAC_BENCH_SECONDS", 20.0))
ALGO_BYTES_PER_CHBLOCK_IN = 2048          # int16 PCM, SURVEY 8(d)


def synth(n_streams, n_samples, seed, device):
    """SURVEY 8(d) recipe: 8-32 harmonic partials on a random f0 with 1/h roll-off,
    amplitude modulated by Poisson onsets (3/s, 5 ms attack, 200 ms decay), white
    noise at -60 dBFS, peak 0.5, int16; channel 2 = 0.8 x the same partials with
    independent noise.  Returns int16 [S, 2, n]."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    out = torch.empty((n_streams, 2, n_samples), dtype=torch.int16, device=device)
    t = torch.arange(n_samples, device=device, dtype=torch.float32) / FS
    chunk = 16
    for s0 in range(0, n_streams, chunk):
        S = min(chunk, n_streams - s0)
        f0 = 55.0 * (32.0 ** torch.rand(S, 1, generator=g, device=device))
        nh = torch.randint(8, 33, (S, 1), generator=g, device=device)
        x = torch.zeros((S, n_samples), device=device)
        for h in range(1, 33):
            amp = (nh >= h).float() / h
            x += amp * torch.sin(2 * np.pi * h * f0 * t[None, :] + 0.37 * h)
        K = int(3 * n_samples / FS * 2) + 4
        gaps = -torch.log(torch.rand(S, K, generator=g, device=device).clamp_min(1e-6)) / 3.0
        onsets = torch.cumsum(gaps, dim=1)
        idx = torch.searchsorted(onsets, t[None, :].expand(S, -1).contiguous()) - 1
        prev = torch.gather(onsets, 1, idx.clamp_min(0))
        dt = torch.where(idx >= 0, t[None, :] - prev, t[None, :] + 0.05)
        env = (1 - torch.exp(-dt / 0.005)) * torch.exp(-dt / 0.2) + 0.02
        x = x * env
        x = 0.5 * x / x.abs().amax(dim=1, keepdim=True).clamp_min(1e-9)
        n1 = 1e-3 * torch.randn(S, n_samples, generator=g, device=device)
        n2 = 1e-3 * torch.randn(S, n_samples, generator=g, device=device)
        out[s0:s0 + S, 0] = torch.round((x + n1).clamp(-1, 1) * 32767).to(torch.int16)
        out[s0:s0 + S, 1] = torch.round((0.8 * x + n2).clamp(-1, 1) * 32767).to(torch.int16)
    return out


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop, self.rows = index, threading.Event(), []

    def run(self):
        while not self.stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                self.rows.append([c.strip() for c in o.stdout.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons}


def oracle_rate(pcm_list, procs):
    """audio-seconds per second of the oracle (NumPy port of the reference) doing
    encode+decode of the given int16 [2, n] streams, `procs` worker processes."""
    t0 = time.time()
    if procs <= 1:
        for p in pcm_list:
            _oracle_one(p)
    else:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(procs) as pool:
            pool.map(_oracle_one, pcm_list)
    dt = time.time() - t0
    return sum(p.shape[1] for p in pcm_list) / FS / dt, dt


def _oracle_one(pcm):
    from oracle import pac as opac
    buf = opac.encode_pcm(pcm, FS, RATE)
    opac.decode_pac(buf)
    return len(buf)


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores (the oracle, a
    NumPy restatement pinned byte-for-byte to the reference; the Python reference
    itself cannot travel to the GPU box), all cores, bounded sample per step."""
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    secs = float(os.environ.get("PAC_REF_SECONDS", 2.0))
    n = int(secs * FS)
    pcm = synth(cores, n, 422, "cpu").numpy()
    streams = [pcm[i] for i in range(cores)]
    for _ in range(max(0, min(args.warmup, 1))):
        oracle_rate(streams[:cores], cores)
    t0 = time.time()
    for _ in range(args.steps):
        oracle_rate(streams, cores)
    dt = time.time() - t0
    val = args.steps * cores * secs / dt
    line = {"impl": "reference", "metric": "encoded audio-sec/sec (enc+dec, 128 kb/s)", "value": val,
            "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"enc+dec 128 kb/s/ch, synthetic 48 kHz stereo, {cores} streams x {secs:g} s per step"},
            "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": cores, "kind": "port",
                             "sample": f"{cores} streams x {secs:g} s per step, {args.steps} steps, multiprocessing.Pool({cores})"},
            "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def mdct_leg(ctx, peak):
    """BASELINE config 2 / the metric's "MDCT HBM GB/s": batched window+MDCT and IMDCT+window+OLA,
    N = 2048, sine window, 1 Mi synthetic stereo frames (fp32), timed with CUDA events; algorithmic
    bytes = one read + one write of 1024 values per channel-frame and direction."""
    import torch
    from audio_codec_b200 import _lib, mdct
    F = int(os.environ.get("PAC_MDCT_FRAMES", 1 << 20))
    out = {"workload": f"{F} stereo frames, N=2048, sine window, inputs {2 * (F + 1) * 4096 / 1e9:.1f} GB (larger than L2)"}
    for name, dt, f in (("f32", torch.float32, F), ("f64", torch.float64, F // 2)):
        g = torch.Generator(device="cuda").manual_seed(422)
        x = (0.25 * torch.randn((2, (f + 1) * 1024), generator=g, device="cuda", dtype=torch.float32)).clamp_(-1, 1).to(dt)
        X = mdct.mdct_frames(x, 2048, _lib.W_SINE, ctx)
        y = mdct.imdct_ola(X, 2048, _lib.W_SINE, ctx)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        tfs, tis = [], []
        for _ in range(7):
            ev[0].record()
            X = mdct.mdct_frames(x, 2048, _lib.W_SINE, ctx)
            ev[1].record()
            y = mdct.imdct_ola(X, 2048, _lib.W_SINE, ctx)
            ev[2].record()
            torch.cuda.synchronize()
            tfs.append(ev[0].elapsed_time(ev[1]))
            tis.append(ev[1].elapsed_time(ev[2]))
        tf, ti = float(np.median(tfs)), float(np.median(tis))      # median of 7: allocator hiccups out
        algo = 2 * f * 1024 * x.element_size() * 2
        err = float((y[:, 1024:-1024] - x[:, 1024:-1024]).abs().max())
        out[name] = {"fwd_ms": tf, "inv_ms": ti, "fwd_gbs": algo / tf / 1e6, "inv_gbs": algo / ti / 1e6,
                     "fwd_frac": algo / tf / 1e6 / peak, "inv_frac": algo / ti / 1e6 / peak, "tdac_max_err": err,
                     "fwd_ms_all": [round(v, 3) for v in tfs], "inv_ms_all": [round(v, 3) for v in tis]}
        del x, X, y
    return out


def smr_leg(ctx, peak):
    """The metric's "SMR HBM GB/s": psychoac.CalcSMRs batched over synthetic frames (N = 2048; one CTA
    per frame: Hann FFT, tonal maskers, masked threshold, per-band SMR).  Buffers are fp64 as the
    reference's arrays are (16 384 B samples + 8 192 B lines + 4 B scale in, 8 B x bands out per
    frame); the arithmetic is fp64 or fp32.  Compute bound (>= 250 flop/B): the fraction of HBM peak
    is reported for the record next to frames/s."""
    import ctypes as C
    import torch
    from audio_codec_b200 import _lib, mdct
    F = int(os.environ.get("PAC_SMR_FRAMES", 1 << 17))
    g = torch.Generator(device="cuda").manual_seed(422)
    t = torch.arange(2048, device="cuda", dtype=torch.float64) / FS
    f0 = 55.0 * 2 ** (5 * torch.rand((F, 1), generator=g, device="cuda", dtype=torch.float64))
    x = 0.25 * torch.sin(2 * np.pi * f0 * t) + 0.1 * torch.sin(2 * np.pi * 3.1 * f0 * t + 1.0)
    x += 0.01 * torch.randn((F, 2048), generator=g, device="cuda", dtype=torch.float64)
    X = mdct.mdct_frames(x.reshape(1, -1)[:, :(F + 1) * 1024].contiguous(), 2048, _lib.W_SINE, ctx)
    lines = (X.reshape(-1, 1024)[:F] * 4.0).contiguous()      # hop-1024 frames of the same signal, scaled by 2^2
    assert lines.shape[0] == F
    scale = torch.full((F,), 2, dtype=torch.int32, device="cuda")
    nl, lo = ctx.bands()
    nB = len(nl)
    smr = torch.empty((F, nB), dtype=torch.float64, device="cuda")
    lo_p, nl_p = (C.c_int * nB)(*[int(v) for v in lo]), (C.c_int * nB)(*[int(v) for v in nl])
    out = {"workload": f"{F} frames, N=2048, {nB} bands, fp64 buffers ({F * (16384 + 8192) / 1e9:.1f} GB in, larger than L2)"}
    algo = F * (2048 * 8 + 1024 * 8 + 4 + 8 * nB)
    for name, dt in (("f32", 0), ("f64", 1)):
        def run():
            _lib.check(_lib.lib.pac_calc_smr(ctx.h, dt, C.c_void_p(x.data_ptr()), C.c_void_p(lines.data_ptr()),
                                             C.c_void_p(scale.data_ptr()), F, 2048, nB, lo_p, nl_p,
                                             C.c_void_p(smr.data_ptr()), None, _lib.stream_ptr()))
        run()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        out[name] = {"ms": ms, "frames_per_s": F / (ms * 1e-3), "gbs": algo / ms / 1e6, "frac": algo / ms / 1e6 / peak}
    return out


def variant_leg(engine, _lib, local, d_pcm, nbs, hdr, d_out64, off64, audio_s, steps, rate=None, precision="f64"):
    """The fp32 fast path (precision="f32": transforms + psychoacoustic model in fp32) on the same
    workload, device resident, plus its code-mismatch rate against the fp64 encoding just made
    (fraction of (block, channel) records whose bytes differ, first 8 streams)."""
    import torch
    ctx = _lib.Context(sample_rate=FS, n_channels=2, rate_kbps=rate or RATE, device=local, precision=precision)
    d_out = torch.empty_like(d_out64)
    d_dec = None
    for _ in range(2):
        _, off, err = engine.encode_device(ctx, d_pcm, nbs, hdr, d_out)
        d_dec, po, err2 = engine.decode_device(ctx, d_out, off, d_dec)
    assert (err | err2) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    enc_ms, dec_ms = [], []
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        _, off, _ = engine.encode_device(ctx, d_pcm, nbs, hdr, d_out)
        engine.decode_device(ctx, d_out, off, d_dec)
        a, b = ctx.last_kernel_ms()
        enc_ms.append(a)
        dec_ms.append(b)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    res = {"value": audio_s * steps / (ms * 1e-3), "unit": "audio-s/s (this rank)", "k_encode_ms": float(np.mean(enc_ms)),
           "k_decode_ms": float(np.mean(dec_ms))}
    if precision == "f64":                      # another bit rate: no comparison with the 128 kb/s encoding
        res["config"] = f"{rate or RATE} kb/s/ch, SBR {'on' if (rate or RATE) < 128 else 'off'}, fp64, same streams"
        return res
    k = min(8, len(nbs))
    a = d_out[:int(off[k])].cpu().numpy().tobytes()
    b = d_out64[:int(off64[k])].cpu().numpy().tobytes()
    diff = tot = 0
    for s in range(k):
        ra = engine.block_records(a[int(off[s]):int(off[s + 1])])
        rb = engine.block_records(b[int(off64[s]):int(off64[s + 1])])
        tot += max(len(ra), len(rb))
        diff += max(len(ra), len(rb)) - sum(1 for x, y in zip(ra, rb) if x == y)
    res.update({"record_mismatch_rate": diff / max(tot, 1),
                "mismatch_sample": f"{tot} (block, channel) records of the first {k} streams vs the fp64 encoding",
                "tolerance": "MDCT/SMR 1e-5 rel, PCM <= 1 LSB (tests/test_fp32_gpu.py)"})
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-mdct", action="store_true", help="skip the MDCT/IMDCT and SMR bandwidth legs (BASELINE config 2)")
    ap.add_argument("--no-fp32", action="store_true", help="skip the fp32 fast-path and 96 kb/s legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    import torch.distributed as dist
    from audio_codec_b200 import _lib, engine
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(args.warmup, 3)
    n = int(SECONDS * FS)
    nb = -(-n // 1024)
    S = N_STREAMS
    pcm = synth(S, n, 422 + 1000 * rank, dev)                    # [S, 2, n] int16, seed per rank
    d_pcm = torch.zeros((S, 2, nb * 1024), dtype=torch.int16, device=dev)
    d_pcm[:, :, :n] = pcm
    d_pcm = d_pcm.reshape(-1)
    nbs = [nb] * S
    hdr = [n] * S
    ctx = _lib.Context(sample_rate=FS, n_channels=2, rate_kbps=RATE, device=local)
    if os.environ.get("PAC_BENCH_VARIANT"):                      # A/B timing of earlier kernel forms (same results)
        ctx.set_option("variant", int(os.environ["PAC_BENCH_VARIANT"]))
    cap = _lib.lib.pac_encode_bound(ctx.h, (__import__("ctypes").c_int * S)(*nbs), S)
    d_out = torch.empty(int(cap), dtype=torch.uint8, device=dev)
    d_dec = None
    audio_s = S * n / FS
    n_chblocks = S * (nb + 2) * 2

    def step():
        nonlocal d_dec
        _, off, err = engine.encode_device(ctx, d_pcm, nbs, hdr, d_out)
        d_dec, po, err2 = engine.decode_device(ctx, d_out, off, d_dec)
        return off, po, err | err2

    for _ in range(W):
        off, po, err = step()
    assert err == 0, f"soft error flags {err:#x}"
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    enc_ms, dec_ms = [], []
    kern = {k: [] for k in ("k_encode", "k_pvq", "k_dequant", "k_decode")}
    kern_n = {}
    launches0 = ctx.kernel_launches()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step()
        a, b = ctx.last_kernel_ms()
        enc_ms.append(a)
        dec_ms.append(b)
        for kind, name in enumerate(kern):          # events around every launch of the four main kernels
            ms, nl = ctx.kernel_stats(kind)
            kern[name].append(ms)
            kern_n[name] = nl
    e1.record()
    launches = ctx.kernel_launches() - launches0
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1)
    sampler.stop.set()
    sampler.join()
    # ---- end to end through the host-buffer C ABI
    h_pcm = torch.empty(d_pcm.numel(), dtype=torch.int16).pin_memory()
    h_pcm.copy_(d_pcm)
    pac_bytes = int(off[-1])
    h_pac = torch.empty(pac_bytes + 4096, dtype=torch.uint8).pin_memory()
    h_dec = torch.empty(int(po[-1]), dtype=torch.int16).pin_memory()

    def step_host():
        o, _ = engine.encode_host(ctx, h_pcm, nbs, hdr, h_pac)
        engine.decode_host(ctx, h_pac, o, h_dec)
        return o

    step_host()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ke = max(2, args.steps // 2)
    for _ in range(ke):
        step_host()
    torch.cuda.synchronize()
    host_ms = 1000 * (time.perf_counter() - t0)
    tt = torch.tensor([dev_ms, host_ms / ke * args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.barrier()
    dev_ms, host_ms_scaled = float(tt[0]), float(tt[1])
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # dominant kernel: k_encode (window + MDCT, Hann FFT, psychoacoustic model), one launch per batch of
        # channel-blocks; its launches of a step together process every channel-block once
        ke_ms = float(np.mean(kern["k_encode"]))
        ke_n = max(1, kern_n["k_encode"])
        algo = n_chblocks * ALGO_BYTES_PER_CHBLOCK_IN + pac_bytes
        ach = algo / (ke_ms * 1e-3) / 1e9
        line = {"metric": "encoded audio-sec/sec (enc+dec, 128 kb/s)", "value": world * audio_s * args.steps / (dev_ms * 1e-3),
                "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": W,
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"enc+dec 128 kb/s/ch (VQ, block switching), {S} synthetic 48 kHz stereo streams x {SECONDS:g} s per GPU per step",
                           "l2": "inputs (%.0f MB PCM/step) larger than L2" % (d_pcm.numel() * 2 / 1e6),
                           "parallelism": f"streams sharded over {world} GPU(s), no collective on the data path"},
                "e2e": {"value": world * audio_s * args.steps / (host_ms_scaled * 1e-3), "unit": "audio-s/s",
                        "h2d_bytes_per_step": d_pcm.numel() * 2 + pac_bytes, "d2h_bytes_per_step": pac_bytes + int(po[-1]) * 2},
                # counted by the library (pac_kernel_launches): per step k_transient, per batch k_encode / k_bitalloc /
                # k_pvq / k_pack, k_sizes, scan, k_gather, k_headers; 2 x k_walk, per batch k_hdr / k_dequant / k_decode, k_ola
                "gpu_launches": int(launches),
                "clocks": sampler.summary(),
                "roofline": {"bound": "hbm", "kernel": "k_encode", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": None,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                             "launches_per_step": ke_n, "avg_launch_ms": ke_ms / ke_n,
                             "algorithmic_bytes_per_launch": algo / ke_n,
                             "kernels_ms_per_step": {k: float(np.mean(v)) for k, v in kern.items()},
                             "kernel_launches_per_step": dict(kern_n),
                             "k_encode_ms": float(np.mean(enc_ms)), "k_decode_ms": float(np.mean(dec_ms)),
                             "stage_note": "k_encode_ms / k_decode_ms are the whole encode / decode stages (events around all their kernels)",
                             "note": "k_encode is compute bound (fp64 psychoacoustic model: ~3 transcendentals per line), see DESIGN.md"}}
        try:        # dram bytes per k_encode launch from the committed ncu capture of this same configuration
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tr.get("streams") == S and tr.get("seconds") == SECONDS:
                line["roofline"]["traffic"] = tr["k_encode_dram_bytes"]
                line["roofline"]["traffic_other_kernels"] = {k: v for k, v in tr.items() if k.endswith("_dram_bytes") and k != "k_encode_dram_bytes"}
        except Exception:
            pass
        if not args.no_mdct:
            line["mdct"] = mdct_leg(ctx, peak)
            line["smr"] = smr_leg(ctx, peak)
        if not args.no_fp32:
            line["fp32"] = variant_leg(engine, _lib, local, d_pcm, nbs, hdr, d_out, off, audio_s, args.steps, precision="f32")
            line["rate96"] = variant_leg(engine, _lib, local, d_pcm, nbs, hdr, d_out, off, audio_s, args.steps, rate=96)
        if not args.no_cpu:
            cs = float(os.environ.get("PAC_CPU_SECONDS", 12.0))
            sample = pcm[0, :, :int(cs * FS)].cpu().numpy()
            v, dt = oracle_rate([sample], 1)
            line["cpu_baseline"] = {"value": v, "unit": "audio-s/s", "cores": 1, "kind": "port",
                                    "sample": f"stream 0, first {cs:g} s, encode+decode, 1 process ({dt:.1f} s wall)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
