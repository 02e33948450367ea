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


def _oracle_codec(job):
    """(int16 [C, n], fs, rate) -> (.pac bytes, decoded int16 [C, m]) by the oracle (worker process)."""
    from oracle import pac as opac
    pcm, fs, rate = job
    buf = opac.encode_pcm(pcm, fs, rate)
    _, y = opac.decode_pac(buf)
    return buf, y


def _compare(engine, gpu_pacs, gpu_pcm, ora):
    """Records (block, channel) that differ between the GPU's and the oracle's encodings of the same streams, and
    the largest |difference| (LSB) between the GPU's decode of ITS stream and the oracle's decode of its own."""
    rec = bad = 0
    lsb = 0
    for g, y, (o, yo) in zip(gpu_pacs, gpu_pcm, ora):
        rg, ro = engine.block_records(g), engine.block_records(o)
        n = max(len(rg), len(ro))
        rec += n
        bad += n - sum(1 for a, b in zip(rg, ro) if a == b)
        hl = 30 + 2 * int.from_bytes(g[26:30], "little")
        if g[:hl] != o[:hl]:
            bad += 1                                    # stream header
        m = min(y.shape[1], yo.shape[1])
        d = int(np.abs(y[:, :m].astype(np.int32) - yo[:, :m].astype(np.int32)).max()) if m else 0
        lsb = max(lsb, d, 0 if y.shape == yo.shape else 65535)
    return rec, bad, lsb


def parity_leg(engine, _lib, local, pcm, budget_s):
    """GPU (fp64 path, through the C ABI) against the oracle on a sample of the bench corpus, at 128 and 96 kb/s:
    the first T seconds of as many streams as the host cores code in ~budget_s seconds.  Bit-exactness is the bar:
    `mismatching` is the number of (block, channel) records whose bytes differ, `pcm_max_lsb` the largest
    difference between the decoded PCM.  The ill-conditioned DC case (tests/edge_signals.py "dc": FFT rounding
    noise gets quantised) is reported separately as a count, not folded into the sample."""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    T = float(os.environ.get("PAC_PARITY_STREAM_SECONDS", 2.0))
    per_core_rate = 0.5                                  # audio-s/s of one oracle process, encode + decode
    per_rate = max(2, min(pcm.shape[0], int(budget_s * cores * per_core_rate / (2 * T))))
    n = int(T * FS)
    sample = [np.ascontiguousarray(pcm[i, :, :n].cpu().numpy()) for i in range(per_rate)]
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import edge_signals
    dc_fs, dc_pcm = edge_signals.cases()["dc"]
    jobs = [(p, FS, r) for r in (128, 96) for p in sample] + [(dc_pcm, dc_fs, r) for r in (128, 96)]
    t0 = time.time()
    with mp.get_context("fork").Pool(min(cores, len(jobs))) as pool:
        ora = pool.map(_oracle_codec, jobs, chunksize=1)
    wall = time.time() - t0
    out = {"records": 0, "mismatching": 0, "pcm_max_lsb": 0, "by_rate": {}}
    for k, r in enumerate((128, 96)):
        ctx = _lib.Context(sample_rate=FS, n_channels=2, rate_kbps=r, device=local)
        g = engine.encode_streams(ctx, sample)
        y = engine.decode_streams(g, ctx=ctx)
        rec, bad, lsb = _compare(engine, g, y, ora[k * per_rate:(k + 1) * per_rate])
        out["by_rate"][str(r)] = {"records": rec, "mismatching": bad, "pcm_max_lsb": lsb}
        out["records"] += rec
        out["mismatching"] += bad
        out["pcm_max_lsb"] = max(out["pcm_max_lsb"], lsb)
        ctx.close()
    dc = {}
    for k, r in enumerate((128, 96)):
        ctx = _lib.Context(sample_rate=dc_fs, n_channels=dc_pcm.shape[0], rate_kbps=r, device=local)
        g = engine.encode_streams(ctx, [dc_pcm])
        y = engine.decode_streams(g, ctx=ctx)
        rec, bad, lsb = _compare(engine, g, y, ora[2 * per_rate + k:2 * per_rate + k + 1])
        dc[str(r)] = {"records": rec, "mismatching": bad, "pcm_max_lsb": lsb}
        ctx.close()
    out["ill_conditioned_dc"] = dc
    out["sample"] = (f"first {T:g} s of streams 0..{per_rate - 1} of the bench corpus at 128 and 96 kb/s/ch, GPU fp64 path vs the "
                     f"oracle on {min(cores, len(jobs))} host processes ({wall:.1f} s wall); records = (block, channel) payloads")
    return out


def copy_roofline(dev, h2d_bytes, d2h_bytes, world):
    """What the host side can do at best: bare pinned H2D + D2H copies of one step's bytes, all ranks at once
    (barrier before), no kernels.  GB/s per rank: the slowest rank's; e2e cannot beat bytes / this."""
    import torch
    import torch.distributed as dist
    n_up, n_dn = int(h2d_bytes), int(d2h_bytes)
    h_up = torch.empty(n_up, dtype=torch.uint8).pin_memory()
    h_dn = torch.empty(n_dn, dtype=torch.uint8).pin_memory()
    d_up = torch.empty(n_up, dtype=torch.uint8, device=dev)
    d_dn = torch.empty(n_dn, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for name in ("h2d", "d2h", "both"):
        ts = []
        for it in range(4):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            if name in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_up.copy_(h_up, non_blocking=True)
            if name in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_dn.copy_(d_dn, non_blocking=True)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        t = min(ts[1:])
        b = (n_up if name != "d2h" else 0) + (n_dn if name != "h2d" else 0)
        res[name] = b / t / 1e9
    v = torch.tensor([res["h2d"], res["d2h"], res["both"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(v, op=dist.ReduceOp.MIN)
    return {"ranks": world, "h2d_gbs": float(v[0]), "d2h_gbs": float(v[1]), "both_directions_gbs": float(v[2]),
            "bytes": {"h2d": n_up, "d2h": n_dn}, "note": "pinned host memory, min over ranks, all ranks copying at once"}


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
    N = 2048, sine and Kaiser ("KBD", coder/window.py:45-58) windows, 1 Mi synthetic stereo frames (fp32), timed
    with CUDA events; algorithmic bytes = one read + one write of 1024 values per channel-frame and direction."""
    import torch
    from audio_codec_b200 import _lib, mdct
    F = int(os.environ.get("PAC_MDCT_FRAMES", 1 << 20))
    out = {"workload": f"{F} stereo frames, N=2048, sine and Kaiser windows, inputs {2 * (F + 1) * 4096 / 1e9:.1f} GB (larger than L2)"}
    for name, dt, f in (("f32", torch.float32, F), ("f64", torch.float64, F // 2)):
        g = torch.Generator(device="cuda").manual_seed(422)
        x = (0.25 * torch.randn((2, (f + 1) * 1024), generator=g, device="cuda", dtype=torch.float32)).clamp_(-1, 1).to(dt)
        algo = 2 * f * 1024 * x.element_size() * 2
        for wname, wid in (("", _lib.W_SINE), ("kaiser_", _lib.W_KAISER)):
            X = mdct.mdct_frames(x, 2048, wid, ctx)
            y = mdct.imdct_ola(X, 2048, wid, ctx)
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            tfs, tis = [], []
            for _ in range(7 if not wname else 5):
                ev[0].record()
                X = mdct.mdct_frames(x, 2048, wid, ctx)
                ev[1].record()
                y = mdct.imdct_ola(X, 2048, wid, ctx)
                ev[2].record()
                torch.cuda.synchronize()
                tfs.append(ev[0].elapsed_time(ev[1]))
                tis.append(ev[1].elapsed_time(ev[2]))
            tf, ti = float(np.median(tfs)), float(np.median(tis))      # median: allocator hiccups out
            r = out.setdefault(name, {})
            r.update({wname + "fwd_ms": tf, wname + "inv_ms": ti, wname + "fwd_gbs": algo / tf / 1e6,
                      wname + "inv_gbs": algo / ti / 1e6, wname + "fwd_frac": algo / tf / 1e6 / peak,
                      wname + "inv_frac": algo / ti / 1e6 / peak})
            if not wname:       # sine: perfect reconstruction (the Kaiser window of the reference is not power complementary)
                r["tdac_max_err"] = float((y[:, 1024:-1024] - x[:, 1024:-1024]).abs().max())
                r["fwd_ms_all"] = [round(v, 3) for v in tfs]
                r["inv_ms_all"] = [round(v, 3) for v in tis]
            del X, y
        del x
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


def cfg4_leg(engine, _lib, local, steps):
    """BASELINE config 4: gain-shape / PVQ + SBR encode and Decode_SBR at 96 kb/s on the test signals batched
    x1000 (coder/codec.py:426-555, :95-222).  The committed excerpts of the four signals whose .wav ships with the
    reference (tests/golden/excerpts.npz, ~0.9 s each) are tiled x1000 along the stream axis; replica 0 of each
    is checked byte for byte against the reference-written golden .pac and its decode against the reference's PCM
    inside the leg.  The fifth signal (glockenspiel, decode only: its .wav is not in the checkout) is the
    reference's own 96 kb/s .pac decoded in x16 replicas, checked against the reference decoder's PCM (md5)."""
    import hashlib
    import torch
    g = os.path.join(ROOT, "tests", "golden")
    ex = np.load(os.path.join(g, "excerpts.npz"))
    dec = np.load(os.path.join(g, "decoded.npz"))
    names = ("castanet", "harpsichord", "quar48_1", "spmg")
    REP = int(os.environ.get("PAC_CFG4_REPLICAS", 1000))
    fs = int(ex["castanet_fs"])
    ctx = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=96, device=local)
    streams = [ex[n] for n in names]
    buf1, nb1 = engine.pack_pcm(streams, 2)
    # stream order: replica-major (replica r holds the four signals), so replica 0 = streams 0..3
    h_pcm = torch.from_numpy(np.tile(buf1, REP)).pin_memory()
    nbs = [int(v) for v in nb1] * REP
    hdr = [s.shape[1] for s in streams] * REP
    d_pcm = h_pcm.cuda()
    cap = _lib.lib.pac_encode_bound(ctx.h, (__import__("ctypes").c_int * len(nbs))(*nbs), len(nbs))
    d_out = torch.empty(int(cap), dtype=torch.uint8, device=d_pcm.device)
    d_dec = None
    for _ in range(2):
        _, off, err = engine.encode_device(ctx, d_pcm, nbs, hdr, d_out)
        d_dec, po, err2 = engine.decode_device(ctx, d_out, off, d_dec)
    assert (err | err2) == 0
    host = d_out[:int(off[4])].cpu().numpy().tobytes()
    pcm0 = d_dec[:int(po[4])].cpu().numpy()
    ok = True
    for k, n in enumerate(names):
        ok &= host[int(off[k]):int(off[k + 1])] == open(os.path.join(g, f"{n}_96.pac"), "rb").read()
        want = dec[f"{n}_96"]
        y = pcm0[int(po[k]):int(po[k + 1])].reshape(2, -1)
        ok &= bool(np.array_equal(y[:, :want.shape[1]], want))
    last = d_out[int(off[-5]):int(off[-1])].cpu().numpy().tobytes()      # the last replica equals the first
    ok &= last == host
    audio_s = sum(s.shape[1] for s in streams) / fs * REP
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        _, off, _ = engine.encode_device(ctx, d_pcm, nbs, hdr, d_out)
        engine.decode_device(ctx, d_out, off, d_dec)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    h_pac = torch.empty(int(off[-1]) + 4096, dtype=torch.uint8).pin_memory()
    h_dec = torch.empty(int(po[-1]), dtype=torch.int16).pin_memory()
    o, _ = engine.encode_host(ctx, h_pcm, nbs, hdr, h_pac)
    engine.decode_host(ctx, h_pac, o, h_dec)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(max(2, steps // 2)):
        o, _ = engine.encode_host(ctx, h_pcm, nbs, hdr, h_pac)
        engine.decode_host(ctx, h_pac, o, h_dec)
    torch.cuda.synchronize()
    hms = 1000 * (time.perf_counter() - t0) / max(2, steps // 2)
    ok &= h_pac[:int(off[4])].numpy().tobytes() == host
    res = {"workload": f"4 test-signal excerpts x {REP} replicas = {4 * REP} streams ({audio_s:.0f} audio-s), {fs} Hz stereo, "
                       "96 kb/s/ch, VQ + SBR, encode + decode",
           "value": audio_s / (ms * 1e-3), "e2e": audio_s / (hms * 1e-3), "unit": "audio-s/s (this rank)", "ms_per_step": ms,
           "replica0_identical_to_reference_goldens": bool(ok)}
    del d_pcm, d_out, d_dec, h_pcm, h_pac, h_dec
    # glockenspiel: decode only
    md5 = json.load(open(os.path.join(g, "full_md5.json")))
    gl = open(os.path.join(g, "glockenspiel_coded_96.pac"), "rb").read()
    GR = int(os.environ.get("PAC_CFG4_GLOCK_REPLICAS", 16))
    off = np.arange(GR + 1, dtype=np.int64) * len(gl)
    d_pac = torch.from_numpy(np.frombuffer(gl * GR, dtype=np.uint8).copy()).cuda()
    d_pcm, po, err = engine.decode_device(ctx, d_pac, off)
    y = d_pcm[int(po[GR - 1]):int(po[GR])].cpu().numpy().reshape(2, -1)
    gok = err == 0 and hashlib.md5(y.astype("<i2").tobytes()).hexdigest() == md5["oracle:glockenspiel_96.pcm"]
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        engine.decode_device(ctx, d_pac, off, d_pcm)
    e1.record()
    torch.cuda.synchronize()
    gms = e0.elapsed_time(e1) / steps
    gs = y.shape[1] / fs * GR
    res["glockenspiel_decode"] = {"workload": f"the reference's glockenspiel_coded_96.pac x {GR} replicas ({gs:.0f} audio-s), decode only",
                                  "value": gs / (gms * 1e-3), "unit": "audio-s/s (this rank)", "ms_per_step": gms,
                                  "pcm_identical_to_reference_decoder": bool(gok)}
    ctx.close()
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
    ap.add_argument("--no-parity", action="store_true", help="skip the GPU-vs-oracle parity sample (~30 s of host work)")
    ap.add_argument("--no-cfg4", action="store_true", help="skip the test-signals x1000 leg (BASELINE config 4)")
    ap.add_argument("--no-copy", action="store_true", help="skip the bare host<->device copy measurement")
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
    if os.environ.get("PAC_BENCH_BATCH_CB"):                     # channel-blocks per launch of the thread-per-band kernels
        ctx.set_option("batch_cb", int(os.environ["PAC_BENCH_BATCH_CB"]))
    if os.environ.get("PAC_BENCH_HOST_CHUNK_MB"):                # smallest chunk of the host-buffer pipeline (MB of PCM)
        ctx.set_option("host_chunk_mb", float(os.environ["PAC_BENCH_HOST_CHUNK_MB"]))
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
    # ---- BASELINE config 5: decode only (.pac unpack + dequantise + SBR / IMDCT + overlap-add), every rank
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0.record()
    for _ in range(args.steps):
        engine.decode_device(ctx, d_out, off, d_dec)
    e1.record()
    torch.cuda.synchronize()
    dec_only_ms = e0.elapsed_time(e1)
    engine.decode_host(ctx, h_pac, off, h_dec)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(ke):
        engine.decode_host(ctx, h_pac, off, h_dec)
    torch.cuda.synchronize()
    dec_host_ms = 1000 * (time.perf_counter() - t0) / ke * args.steps
    tt = torch.tensor([dev_ms, host_ms / ke * args.steps, dec_only_ms, dec_host_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.barrier()
    dev_ms, host_ms_scaled, dec_only_ms, dec_host_ms = (float(v) for v in tt)
    h2d_step, d2h_step = d_pcm.numel() * 2 + pac_bytes, pac_bytes + int(po[-1]) * 2
    copies = None if args.no_copy else copy_roofline(dev, h2d_step, d2h_step, world)
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
                        "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step},
                # BASELINE config 5: decode only, device resident and through pac_decode_host (host .pac in, host PCM out)
                "decode_only": {"value": world * audio_s * args.steps / (dec_only_ms * 1e-3), "unit": "audio-s/s",
                                "ms_per_step": dec_only_ms / args.steps,
                                "e2e": {"value": world * audio_s * args.steps / (dec_host_ms * 1e-3), "unit": "audio-s/s",
                                        "h2d_bytes_per_step": pac_bytes, "d2h_bytes_per_step": int(po[-1]) * 2},
                                "workload": "the .pac streams of this step's encode: k_walk, k_hdr, k_dequant, k_decode, k_ola"},
                # counted by the library (pac_kernel_launches): per step k_transient, per batch k_encode / k_bitalloc /
                # k_pvq / k_pack, k_sizes, scan, k_gather, k_headers; 2 x k_walk, per batch k_hdr / k_dequant / k_decode, k_ola
                "gpu_launches": int(launches),
                "clocks": sampler.summary(),
                "roofline": {"bound": "hbm", "kernel": "k_encode (k_encode_front)", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": None,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                             "launches_per_step": ke_n, "avg_launch_ms": ke_ms / ke_n,
                             "algorithmic_bytes_per_launch": algo / ke_n,
                             "kernels_ms_per_step": {k: float(np.mean(v)) for k, v in kern.items()},
                             "kernel_launches_per_step": dict(kern_n),
                             "k_encode_ms": float(np.mean(enc_ms)), "k_decode_ms": float(np.mean(dec_ms)),
                             "stage_note": "k_encode_ms / k_decode_ms are the whole encode / decode stages (events around all their kernels)",
                             "note": "k_encode is latency / fp64 bound, not HBM bound (see roofline.compute and DESIGN.md 4.1): the HBM fraction is small by construction"}}
        if copies:
            line["copy_roofline_gbs"] = copies
            # the step's copies at that rate, against the e2e step: how much of e2e is the host link
            best = max(h2d_step / copies["h2d_gbs"], d2h_step / copies["d2h_gbs"]) / 1e6
            line["e2e"]["copy_bound_ms_per_step"] = best
            line["e2e"]["ms_per_step"] = host_ms_scaled / args.steps
        try:        # per-step ncu figures of this same configuration, from the committed capture (replayed, not measured here)
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tr.get("streams") == S and tr.get("seconds") == SECONDS:
                line["roofline"]["traffic"] = tr["k_encode_dram_bytes"] / ke_n
                line["roofline"]["traffic_source"] = tr.get("source")
                line["roofline"]["traffic_per_step_by_kernel"] = {k: v for k, v in tr.items() if k.endswith("_dram_bytes")}
                if tr.get("k_encode_fp64_warp_inst"):
                    # compute side: fp64 pipe.  64 fp64 lanes per SM (an fp64 warp instruction occupies the pipe for 2 cycles).
                    clk = (line["clocks"].get("sm_mhz") or 1965.0) * 1e6
                    peak_ops = 148 * 64 * clk
                    ach_ops = tr["k_encode_fp64_warp_inst"] * 32 / (ke_ms * 1e-3)
                    line["roofline"]["compute"] = {
                        "bound": "fp64", "kernel": "k_encode", "achieved": ach_ops / 1e12, "peak": peak_ops / 1e12,
                        "unit": "T fp64 lane-ops/s (FMA = 1)", "frac": ach_ops / peak_ops,
                        "fp64_warp_instructions_per_step": tr["k_encode_fp64_warp_inst"],
                        "ncu_fp64_pipe_pct": tr.get("k_encode_fp64_pipe_pct"), "ncu_issue_active_pct": tr.get("k_encode_issue_active_pct"),
                        "source": "instruction count from " + str(tr.get("source")) + "; duration and clock measured live"}
        except Exception:
            pass
        if not args.no_parity:
            line["parity"] = parity_leg(engine, _lib, local, pcm, float(os.environ.get("PAC_PARITY_SECONDS", 30.0)))
        if not args.no_cfg4:
            line["cfg4"] = cfg4_leg(engine, _lib, local, max(2, args.steps // 2))
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
