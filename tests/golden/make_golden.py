"""Generate the committed golden vectors by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py [--files] [--kats]

Outputs (committed, small):
  tests/golden/excerpts.npz      int16 PCM excerpts of the four test signals
  tests/golden/<name>_<rate>.pac reference .pac of each excerpt (VQ; SBR @96)
  tests/golden/decoded.npz       reference-decoded int16 PCM of those .pac
  tests/golden/<name>_bfp128.pac reference .pac with useVQ=False,useSBR=False
  tests/golden/kats.npz          per-function known answers (MDCT, windows,
                                 ScaleFactor, CalcSMRs, BitAlloc, PVQ, BFP...)
  tests/golden/full_md5.json     md5 of the reference's own shipped fixtures
                                 test_decoded_full/* (not copied; hashes only)
The reference is driven through its own public entry points
(pacfile.encode_decode_test, PACFile/PCMFile, codec.*).
"""
import argparse
import hashlib
import json
import os
import struct
import sys
import wave
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
from oracle import refload  # noqa: E402

# (signal, first block, number of samples)
EXCERPTS = {
    "castanet": (0, 40 * 1024 + 300),
    "harpsichord": (0, 40 * 1024),
    "quar48_1": (5, 36 * 1024 + 17),
    "spmg": (0, 36 * 1024),
}
RATES = (96, 128)


def read_wav(path):
    w = wave.open(str(path))
    n = w.getnframes()
    x = np.frombuffer(w.readframes(n), dtype="<i2").reshape(-1, w.getnchannels())
    return w.getframerate(), x.T.copy()


def write_wav(path, fs, pcm):
    w = wave.open(str(path), "wb")
    w.setnchannels(pcm.shape[0])
    w.setsampwidth(2)
    w.setframerate(fs)
    w.writeframes(pcm.T.astype("<i2").tobytes())
    w.close()


def gen_files(tmp):
    import pacfile
    from pacfile import PACFile, PCMFile, parTransientDetect
    pacfile.output_dir = tmp / "out"
    os.makedirs(pacfile.output_dir, exist_ok=True)
    exc, dec = {}, {}
    for name, (b0, ns) in EXCERPTS.items():
        fs, x = read_wav(Path(refload.REF_ROOT) / "test_signals" / f"{name}.wav")
        pcm = x[:, b0 * 1024:b0 * 1024 + ns]
        exc[name] = pcm
        exc[name + "_fs"] = np.int64(fs)
        src = tmp / f"{name}.wav"
        write_wav(src, fs, pcm)
        for rate in RATES:
            pacfile.encode_decode_test(src, rate)
            pac = (pacfile.output_dir / f"{name}_coded_{rate}.pac").read_bytes()
            (HERE / f"{name}_{rate}.pac").write_bytes(pac)
            _, y = read_wav(pacfile.output_dir / f"{name}_{rate}.wav")
            dec[f"{name}_{rate}"] = y
        # block-floating-point variant: same driver loop as
        # pacfile.encode_decode_test but useVQ=False, useSBR=False
        if name in ("castanet", "harpsichord"):
            out = tmp / f"{name}_bfp128.pac"
            inF, outF = PCMFile(str(src)), PACFile(str(out))
            cp = inF.OpenForReading()
            cp.nMDCTLines = 1024
            cp.nScaleBits = 4
            cp.nMantSizeBits = 12
            cp.targetBitsPerSample = 128 / (cp.sampleRate / 1000)
            cp.useSBR = False
            cp.useVQ = False
            cp.nSamplesPerBlock = 1024
            outF.OpenForWriting(cp)
            look = np.zeros((cp.nChannels, 2048))
            cur = last = False
            while True:
                data = inF.ReadDataBlock(cp)
                if not data:
                    nxt = False
                else:
                    look = np.concatenate((np.copy(data), look[:, 1024:]), axis=1)
                    nxt = parTransientDetect(look)
                outF.WriteDataBlock(look[:, :1024], cp, lastTrans=last,
                                    curTrans=cur, nextTrans=nxt)
                last, cur = cur, nxt
                if not data:
                    break
            inF.Close(cp)
            outF.Close(cp)
            (HERE / f"{name}_bfp128.pac").write_bytes(out.read_bytes())
            inF, outF = PACFile(str(out)), PCMFile(str(tmp / "bfp.wav"))
            cp = inF.OpenForReading()
            cp.bitsPerSample = 16
            outF.OpenForWriting(cp)
            while True:
                data = inF.ReadDataBlock(cp)
                if not data:
                    break
                outF.WriteDataBlock(data, cp)
            inF.Close(cp)
            outF.Close(cp)
            _, y = read_wav(tmp / "bfp.wav")
            dec[f"{name}_bfp128"] = y
    np.savez_compressed(HERE / "excerpts.npz", **exc)
    np.savez_compressed(HERE / "decoded.npz", **dec)
    md5 = {}
    for p in sorted((Path(refload.REF_ROOT) / "test_decoded_full").iterdir()):
        md5[p.name] = hashlib.md5(p.read_bytes()).hexdigest()
    (HERE / "full_md5.json").write_text(json.dumps(md5, indent=1))


class _Bag:
    pass


def gen_kats():
    """Per-function known answers straight from the reference modules."""
    import mdct as r_mdct
    import window as r_window
    import quantize as r_q
    import psychoac as r_psy
    import bitalloc as r_ba
    import gain_shape_quantize as r_gs
    import codec as r_codec
    import sbr as r_sbr
    from bitpack import PackedBits
    from detect_transients import parTransientDetect
    k = {}
    rng = np.random.default_rng(422)
    # the reference's own self-test vectors (mdct.py:86-107, quantize.py:283-319,
    # bitpack.py:184-197, pvq_exploration.ipynb cell 5)
    k["mdct_kat_in"] = np.array([0, 0, 0, 0, 0, 1, 2, 3], float)
    k["mdct_kat_out"] = r_mdct.MDCT(k["mdct_kat_in"], 4, 4)
    qin = np.array([-0.99, -0.38, -0.10, -0.01, -0.001, 0.0, 0.05, 0.28, 0.65, 0.97, 1.0])
    k["q_in"] = qin
    k["q_u8"] = np.array([r_q.QuantizeUniform(v, 8) for v in qin])
    k["q_u12"] = np.array([r_q.QuantizeUniform(v, 12) for v in qin])
    k["q_vu12"] = r_q.vQuantizeUniform(qin, 12)
    k["q_du12"] = r_q.vDequantizeUniform(k["q_vu12"], 12)
    k["q_scale"] = np.array([r_q.ScaleFactor(v, 3, 5) for v in qin])
    k["q_mant"] = np.array([r_q.Mantissa(v, s, 3, 5) for v, s in zip(qin, k["q_scale"])])
    k["q_deq"] = np.array([r_q.Dequantize(s, m, 3, 5) for s, m in zip(k["q_scale"], k["q_mant"])])
    pb = PackedBits()
    pb.Size(2)
    for v, n in zip((3, 5, 11, 3, 1), (4, 3, 5, 3, 1)):
        pb.WriteBits(v, n)
    k["bitpack_kat"] = np.frombuffer(pb.GetPackedData(), np.uint8)
    # windows and transforms on random data
    for N in (2048, 256):
        x = rng.uniform(-1, 1, N)
        k[f"x{N}"] = x
        k[f"sine{N}"] = r_window.SineWindow(x)
        k[f"hann{N}"] = r_window.HanningWindow(x)
        k[f"kbd{N}"] = r_window.KBDWindow(x)
        k[f"mdct{N}"] = r_mdct.MDCT(r_window.SineWindow(x), N // 2, N // 2)
        k[f"imdct{N}"] = r_mdct.IMDCT(k[f"mdct{N}"], N // 2, N // 2)
    x = k["x2048"]
    k["start2048"] = r_window.StartWindow(x, 2048, 256)
    k["stop2048"] = r_window.StopWindow(x, 2048, 256)
    k["startstop2048"] = r_window.StartStopWindow(x, 2048, 256)
    # band tables
    for fs in (44100, 48000):
        for n in (1024, 128):
            sf = r_psy.ScaleFactorBands(r_psy.AssignMDCTLinesFromFreqLimits(n, fs))
            k[f"bands_{fs}_{n}"] = np.asarray(sf.nLines)
            if n == 1024:
                k[f"omit_{fs}"] = np.asarray(r_sbr.omitted_bands(sf))
    # real frames: SMR / BitAlloc / quantisers, long and short, both rates
    fs, pcm = read_wav(Path(refload.REF_ROOT) / "test_signals" / "harpsichord.wav")
    sig = np.sign(pcm) * 2.0 * np.abs(pcm.astype(np.int64)) / 65535.0
    sig[pcm == -32768] = 0.0
    frames, smrs, mdcts, scales, bas, budgets = [], [], [], [], [], []
    sfL = r_psy.ScaleFactorBands(r_psy.AssignMDCTLinesFromFreqLimits(1024, fs))
    sfS = r_psy.ScaleFactorBands(r_psy.AssignMDCTLinesFromFreqLimits(128, fs))
    for i, b in enumerate((3, 9, 17, 30, 44, 58)):
        fr = sig[i % 2, b * 1024:b * 1024 + 2048].copy()
        frames.append(fr)
        ml = r_mdct.MDCT(r_window.SineWindow(fr), 1024, 1024)[:1024]
        sc = r_q.ScaleFactor(np.max(np.abs(ml)), 4)
        ml = ml * (1 << sc)
        smr = r_psy.CalcSMRs(fr.copy(), ml.copy(), sc, fs, sfL)
        budget = (128 / (fs / 1000)) * 1024 - 4 - 12 * sfL.nBands
        ba = r_ba.BitAlloc(budget, 16, sfL.nBands, sfL.nLines, smr)
        mdcts.append(ml), scales.append(sc), smrs.append(smr), bas.append(ba)
        budgets.append(budget)
    k["fs"] = np.int64(fs)
    k["frames"], k["frame_mdct"], k["frame_scale"] = map(np.array, (frames, mdcts, scales))
    k["frame_smr"], k["frame_ba"], k["frame_budget"] = map(np.array, (smrs, bas, budgets))
    k["frame_mask0"] = r_psy.getMaskedThreshold(frames[0].copy(), mdcts[0].copy(),
                                                int(scales[0]), fs, sfL)
    sh_smr, sh_ba, sh_fr = [], [], []
    for b in (3, 30):
        fr = sig[0, b * 1024 + 448:b * 1024 + 448 + 256].copy()
        ml = r_mdct.MDCT(r_window.SineWindow(fr), 128, 128)[:128]
        sc = r_q.ScaleFactor(np.max(np.abs(ml)), 4)
        ml = ml * (1 << sc)
        smr = r_psy.CalcSMRs(fr.copy(), ml.copy(), sc, fs, sfS)
        budget = (128 / (fs / 1000)) * int(1.45 * 128) - 4 - 12 * sfS.nBands
        sh_fr.append(fr), sh_smr.append(smr)
        sh_ba.append(r_ba.BitAlloc(budget, 16, sfS.nBands, sfS.nLines, smr))
    k["short_frames"], k["short_smr"], k["short_ba"] = map(np.array, (sh_fr, sh_smr, sh_ba))
    # gain-shape VQ: every band of frame 0 at its allocated budget, flattened
    gs_idx, gs_bits, gs_off, gs_deq = [], [], [0], []
    for ib in range(sfL.nBands):
        lo, hi = sfL.lowerLine[ib], sfL.upperLine[ib] + 1
        xb = mdcts[0][lo:hi]
        R = int(bas[0][ib] * sfL.nLines[ib])
        if R == 0:
            gs_off.append(gs_off[-1])
            continue
        idx, bits = r_gs.quantize_gain_shape(xb, R, 0)
        gs_idx += [int(v) for v in idx]
        gs_bits += [int(v) for v in bits]
        gs_off.append(len(gs_idx))
        pb = PackedBits()
        pb.Size((sum(bits) + 7) // 8 + 1)
        for v, n in zip(idx, bits):
            pb.WriteBits(int(v), int(n))
        pb.ResetPointers()
        gs_deq.append(r_gs.dequantize_gain_shape(pb, R, len(xb), 0))
    k["gs_idx"] = np.array(gs_idx, np.uint64)
    k["gs_bits"], k["gs_off"] = np.array(gs_bits), np.array(gs_off)
    k["gs_deq"] = np.concatenate(gs_deq)
    # PVQ codebook facts
    k["pvq_N_6_12"] = r_gs.pvq_codebook_size(6, 12)
    kk = [(L, b) + tuple(int(v) for v in r_gs.pvq_compute_k_for_R(L, b)[:2])
          for L in (3, 4, 5, 7, 13, 21, 41, 82, 163) for b in (1, 5, 12, 20, 27, 32)]
    k["pvq_k_for_R"] = np.array(kk)
    N22 = r_gs.pvq_codebook_size(2, 2)
    vecs = [(2, 0), (-2, 0), (1, 1), (1, -1), (-1, 1), (-1, -1), (0, 2), (0, -2)]
    k["pvq_s22"] = np.array([r_gs.encode_pvq_vector(np.array(v), 2, N22) for v in vecs])
    # BFP on frame 0
    sfs, mants = [], []
    for ib in range(sfL.nBands):
        lo, hi = sfL.lowerLine[ib], sfL.upperLine[ib] + 1
        bb = int(bas[0][ib])
        s = r_q.ScaleFactor(np.max(np.abs(mdcts[0][lo:hi])), 4, bb)
        sfs.append(s)
        if bb:
            mants.append(r_q.vMantissa(mdcts[0][lo:hi], s, 4, bb))
    k["bfp_sf"], k["bfp_mant"] = np.array(sfs), np.concatenate(mants)
    # transient detector
    blk = np.stack([sig[:, b * 1024:(b + 1) * 1024] for b in range(64)])
    k["trans_flags"] = np.array([
        bool(parTransientDetect(np.concatenate([bk, np.zeros((2, 1024))], axis=1)))
        for bk in blk])
    k["trans_pcm"] = pcm[:, :64 * 1024]
    np.savez_compressed(HERE / "kats.npz", **k)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", action="store_true")
    ap.add_argument("--kats", action="store_true")
    a = ap.parse_args()
    refload.install()
    tmp = Path("/tmp/pac_ref_scratch/golden")
    os.makedirs(tmp, exist_ok=True)
    if a.kats or not a.files:
        gen_kats()
    if a.files or not a.kats:
        gen_files(tmp)
