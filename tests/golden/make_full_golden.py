"""Whole-file goldens (north_star acceptance: the five test_signals at 96 and 128 kb/s).

Run in the build container only (needs /root/reference):
    python tests/golden/make_full_golden.py [--skip-reference-decode]

Outputs (committed):
  tests/golden/full_pcm.npz       the four test_signals as the reference's PCMFile reads them
                                  (int16 [C, n] incl. the trailing-chunk quirk) + header sample counts
  tests/golden/full_records.npz   per (signal, rate):
      <tag>_fix_crc / _fix_len    crc32 / length of every (block, channel) record of the reference's own
                                  shipped fixture test_decoded_full/<signal>_coded_<rate>.pac
      <tag>_fix_hdr               the fixture's header bytes
      <tag>_orc_crc / _orc_len    the same for the oracle's whole-file encoding made here
      <tag>_fix_pcm_crc           crc32 per (output block, channel) of the fixture's decoded .wav
      <tag>_orc_pcm_crc           the same for the oracle's decode of the oracle's .pac
    glockenspiel (decode only: its .wav is not in the checkout, .MISSING_LARGE_BLOBS):
      glockenspiel_<rate>_ref_pcm_crc   crc32 per (block, channel) of the UNMODIFIED reference's decode
      glockenspiel_<rate>_ref_pcm_n     samples per channel its WAV writer kept (it stops at numSamples)
      glockenspiel_<rate>_orc_pcm_crc   the oracle's decode (must equal the reference's)
  tests/golden/glockenspiel_coded_{96,128}.pac   the reference's shipped fixtures (test vectors)
  tests/golden/full_md5.json      md5 of fixtures (+ "oracle:<tag>.pac", "oracle:<tag>.pcm", "ref:glockenspiel_<rate>.pcm")
"""
import hashlib
import json
import multiprocessing as mp
import os
import struct
import sys
import wave
import zlib
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
from oracle import pac as opac, refload  # noqa: E402

REF = Path(refload.REF_ROOT)
SIGNALS = ("castanet", "harpsichord", "quar48_1", "spmg")
RATES = (96, 128)


def records(buf):
    """(header bytes, list of record payloads) of one .pac image (coder/pacfile.py:239-256)."""
    nb = struct.unpack_from("<L", buf, 26)[0]
    pos = 30 + 2 * nb
    hdr, out = buf[:pos], []
    while pos + 4 <= len(buf):
        n = struct.unpack_from("<L", buf, pos)[0]
        out.append(buf[pos + 4:pos + 4 + n])
        pos += 4 + n
    return hdr, out


def rec_crc(buf):
    hdr, rec = records(buf)
    return (np.frombuffer(hdr, np.uint8).copy(), np.array([zlib.crc32(r) for r in rec], np.uint32),
            np.array([len(r) for r in rec], np.int32))


def pcm_crc(pcm):
    """crc32 of every 1024-sample block of every channel: [C, ceil(n / 1024)]."""
    C, n = pcm.shape
    nb = -(-n // 1024)
    out = np.zeros((C, nb), np.uint32)
    for c in range(C):
        for b in range(nb):
            out[c, b] = zlib.crc32(np.ascontiguousarray(pcm[c, b * 1024:(b + 1) * 1024]).astype("<i2").tobytes())
    return out


def read_wav(path):
    w = wave.open(str(path))
    x = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, w.getnchannels())
    return w.getframerate(), x.T.copy()


def _oracle_job(args):
    name, rate = args
    fs, pcm, nh = opac.read_wav_like_reference(REF / "test_signals" / f"{name}.wav")
    buf = opac.encode_pcm(pcm, fs, rate, header_samples=nh)
    _, y = opac.decode_pac(buf)
    return name, rate, buf, y


def _oracle_glock(rate):
    buf = (REF / "test_decoded_full" / f"glockenspiel_coded_{rate}.pac").read_bytes()
    _, y = opac.decode_pac(buf)
    return rate, y


def _reference_glock(rate):
    """The unmodified reference's decoder loop (coder/pacfile.py:758-776) on the shipped fixture."""
    refload.install(scratch=f"/tmp/pac_ref_scratch/glock{rate}")
    from pacfile import PACFile, PCMFile
    src = REF / "test_decoded_full" / f"glockenspiel_coded_{rate}.pac"
    dst = Path(f"/tmp/pac_ref_scratch/glock{rate}/glockenspiel_{rate}.wav")
    inF, outF = PACFile(str(src)), PCMFile(str(dst))
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
    return rate, read_wav(dst)[1]


def main():
    skip_ref = "--skip-reference-decode" in sys.argv
    md5 = {}
    for p in sorted((REF / "test_decoded_full").iterdir()):
        md5[p.name] = hashlib.md5(p.read_bytes()).hexdigest()
    pcm_out, rec = {}, {}
    for name in SIGNALS:
        fs, pcm, nh = opac.read_wav_like_reference(REF / "test_signals" / f"{name}.wav")
        pcm_out[name] = pcm
        pcm_out[name + "_fs"] = np.int64(fs)
        pcm_out[name + "_nh"] = np.int64(nh)
        for rate in RATES:
            tag = f"{name}_{rate}"
            fix = (REF / "test_decoded_full" / f"{name}_coded_{rate}.pac").read_bytes()
            rec[tag + "_fix_hdr"], rec[tag + "_fix_crc"], rec[tag + "_fix_len"] = rec_crc(fix)
            _, y = read_wav(REF / "test_decoded_full" / f"{name}_{rate}.wav")
            rec[tag + "_fix_pcm_crc"] = pcm_crc(y)
            rec[tag + "_fix_pcm_n"] = np.int64(y.shape[1])
    np.savez_compressed(HERE / "full_pcm.npz", **pcm_out)
    jobs = [(n, r) for n in SIGNALS for r in RATES]
    with mp.get_context("fork").Pool(min(8, os.cpu_count() or 1)) as pool:
        res = pool.map(_oracle_job, jobs, chunksize=1)
        glock_o = pool.map(_oracle_glock, RATES, chunksize=1)
    for name, rate, buf, y in res:
        tag = f"{name}_{rate}"
        _, rec[tag + "_orc_crc"], rec[tag + "_orc_len"] = rec_crc(buf)
        rec[tag + "_orc_pcm_crc"] = pcm_crc(y)
        rec[tag + "_orc_pcm_n"] = np.int64(y.shape[1])
        md5[f"oracle:{tag}.pac"] = hashlib.md5(buf).hexdigest()
        md5[f"oracle:{tag}.pcm"] = hashlib.md5(y.astype("<i2").tobytes()).hexdigest()
        d = int(np.sum(rec[tag + "_orc_crc"] != rec[tag + "_fix_crc"])) if len(rec[tag + "_orc_crc"]) == len(rec[tag + "_fix_crc"]) else -1
        print(f"{tag}: oracle vs shipped fixture: {d} of {len(rec[tag + '_fix_crc'])} records differ"
              f" {np.nonzero(rec[tag + '_orc_crc'] != rec[tag + '_fix_crc'])[0].tolist() if d > 0 else ''}", flush=True)
    for rate, y in glock_o:
        rec[f"glockenspiel_{rate}_orc_pcm_crc"] = pcm_crc(y)
        rec[f"glockenspiel_{rate}_pcm_n"] = np.int64(y.shape[1])
        md5[f"oracle:glockenspiel_{rate}.pcm"] = hashlib.md5(y.astype("<i2").tobytes()).hexdigest()
        (HERE / f"glockenspiel_coded_{rate}.pac").write_bytes(
            (REF / "test_decoded_full" / f"glockenspiel_coded_{rate}.pac").read_bytes())
    if not skip_ref:
        with mp.get_context("spawn").Pool(2) as pool:
            for rate, y in pool.map(_reference_glock, RATES, chunksize=1):
                rec[f"glockenspiel_{rate}_ref_pcm_crc"] = pcm_crc(y)
                rec[f"glockenspiel_{rate}_ref_pcm_n"] = np.int64(y.shape[1])
                md5[f"ref:glockenspiel_{rate}.pcm"] = hashlib.md5(y.astype("<i2").tobytes()).hexdigest()
                o = rec[f"glockenspiel_{rate}_orc_pcm_crc"]
                r = rec[f"glockenspiel_{rate}_ref_pcm_crc"]
                n = min(o.shape[1], r.shape[1])
                print(f"glockenspiel_{rate}: reference decode {y.shape}, oracle blocks {o.shape}: "
                      f"{int(np.sum(o[:, :n - 1] != r[:, :n - 1]))} blocks differ", flush=True)
    np.savez_compressed(HERE / "full_records.npz", **rec)
    (HERE / "full_md5.json").write_text(json.dumps(md5, indent=1))


if __name__ == "__main__":
    main()
