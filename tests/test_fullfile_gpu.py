"""GPU: north_star's acceptance line - whole files, all five test signals, 96 and 128 kb/s.

The four test_signals (committed as the reference's PCMFile reads them, tests/golden/full_pcm.npz) are encoded
in ONE batch per rate through the C ABI and compared record by record with
  * the reference's own shipped fixtures test_decoded_full/<signal>_coded_<rate>.pac (crc32 per record, md5),
  * the oracle's whole-file encoding made in the build container (tests/golden/make_full_golden.py);
then decoded and compared block by block with the fixture .wav files and the oracle's decode.
glockenspiel (decode only: its .wav is not in the checkout) is decoded from the shipped .pac and compared with
the UNMODIFIED reference decoder's PCM.

quar48_1: 4 of 1978 records (978, 979, 1026, 1027: written blocks 489 and 513, both channels) are ill conditioned -
they lie in a stretch of near-silence (samples in {-1, 0}, some blocks -1 throughout): one real MDCT line and ~1000
lines of FFT rounding noise that the reference still quantises.  The reference does not reproduce its own shipped fixture there when run
in the build container, and the oracle does not even reproduce itself from run to run (pocketfft's SIMD path depends
on buffer alignment; tests/test_oracle_golden.py).  Every other record must be bit-identical to the fixture AND to
the oracle; those four may differ and are REPORTED (count + indices in gpurun_out/fullfile_parity_<rate>.txt).
"""
import hashlib
import json
import os
import struct
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
SIGNALS = ("castanet", "harpsichord", "quar48_1", "spmg")
ILL_CONDITIONED = {"quar48_1": [978, 979, 1026, 1027]}         # record index = block * 2 + channel


def _records(buf):
    nb = struct.unpack_from("<L", buf, 26)[0]
    pos = 30 + 2 * nb
    hdr, out = buf[:pos], []
    while pos + 4 <= len(buf):
        n = struct.unpack_from("<L", buf, pos)[0]
        out.append(buf[pos + 4:pos + 4 + n])
        pos += 4 + n
    return hdr, out


def _pcm_crc(pcm):
    C, n = pcm.shape
    nb = -(-n // 1024)
    out = np.zeros((C, nb), np.uint32)
    for c in range(C):
        for b in range(nb):
            out[c, b] = zlib.crc32(np.ascontiguousarray(pcm[c, b * 1024:(b + 1) * 1024]).astype("<i2").tobytes())
    return out


@pytest.fixture(scope="module")
def gold():
    return (np.load(os.path.join(GOLDEN, "full_pcm.npz")), np.load(os.path.join(GOLDEN, "full_records.npz")),
            json.load(open(os.path.join(GOLDEN, "full_md5.json"))))


@pytest.mark.parametrize("rate", [96, 128])
def test_whole_files_encode_decode_identical_to_reference(rate, gold):
    from audio_codec_b200 import _lib, engine
    pcm, rec, md5 = gold
    fs = int(pcm["castanet_fs"])
    assert all(int(pcm[n + "_fs"]) == fs for n in SIGNALS)
    ctx = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=rate)
    pacs = engine.encode_streams(ctx, [pcm[n] for n in SIGNALS], hdr_samples=[int(pcm[n + "_nh"]) for n in SIGNALS])
    report = {}
    for name, buf in zip(SIGNALS, pacs):
        tag = f"{name}_{rate}"
        hdr, recs = _records(buf)
        crc = np.array([zlib.crc32(r) for r in recs], np.uint32)
        assert hdr == rec[tag + "_fix_hdr"].tobytes(), f"{tag}: header differs from the shipped fixture"
        ill = ILL_CONDITIONED.get(name, [])
        # the oracle (== the reference as run in the build container)
        orc = rec[tag + "_orc_crc"]
        assert len(crc) == len(orc)
        diff_o = np.nonzero(crc != orc)[0].tolist()
        assert set(diff_o) <= set(ill), f"{tag}: records {diff_o[:8]} differ from the oracle"
        # the reference's shipped fixture
        fix = rec[tag + "_fix_crc"]
        assert len(crc) == len(fix)
        diff = np.nonzero(crc != fix)[0].tolist()
        report[tag] = (len(crc), diff, diff_o)
        assert set(diff) <= set(ill), f"{tag}: records {diff[:8]} differ from the shipped fixture"
        if not diff:
            assert hashlib.md5(buf).hexdigest() == md5[f"{name}_coded_{rate}.pac"]
        if not diff_o:
            assert hashlib.md5(buf).hexdigest() == md5[f"oracle:{tag}.pac"]
    # decode the GPU's own streams in one batch
    ys = engine.decode_streams(pacs)
    for name, y in zip(SIGNALS, ys):
        tag = f"{name}_{rate}"
        assert y.shape[1] == int(rec[tag + "_orc_pcm_n"])
        # a record that differs touches its own output block and the next one (overlap-add)
        allowed = {b for r in ILL_CONDITIONED.get(name, []) for b in (r // 2, r // 2 + 1)}
        bad_o = set(np.nonzero(_pcm_crc(y) != rec[tag + "_orc_pcm_crc"])[1].tolist())
        assert bad_o <= allowed, f"{tag}: decoded blocks {sorted(bad_o)} differ from the oracle"
        if not bad_o:
            assert hashlib.md5(y.astype("<i2").tobytes()).hexdigest() == md5[f"oracle:{tag}.pcm"]
        n = int(rec[tag + "_fix_pcm_n"])                    # the reference's WAV writer stops at numSamples
        bad = set(np.nonzero(_pcm_crc(y[:, :n]) != rec[tag + "_fix_pcm_crc"])[1].tolist())
        assert bad <= allowed, f"{tag}: decoded blocks {sorted(bad)} differ from the shipped .wav"
        report[tag] += (sorted(bad),)
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"fullfile_parity_{rate}.txt"), "w") as f:
        for tag, (n, diff, diff_o, bad) in report.items():
            f.write(f"{tag}: {n} records; {len(diff)} differ from the shipped fixture {diff}; {len(diff_o)} differ from "
                    f"the oracle {diff_o}; decoded blocks differing from the shipped .wav: {bad}\n")


@pytest.mark.parametrize("rate", [96, 128])
def test_glockenspiel_decode_identical_to_reference_decoder(rate, gold):
    """The fifth test signal: test_decoded_full/glockenspiel_coded_<rate>.pac (a reference output) decoded on the
    GPU == the unmodified reference's decoder run on the same file (crc32 per 1024-sample block, md5)."""
    from audio_codec_b200 import engine
    _, rec, md5 = gold
    buf = open(os.path.join(GOLDEN, f"glockenspiel_coded_{rate}.pac"), "rb").read()
    assert hashlib.md5(buf).hexdigest() == md5[f"glockenspiel_coded_{rate}.pac"]
    y = engine.decode_streams([buf])[0]
    key = f"glockenspiel_{rate}_ref_pcm_crc"
    assert np.array_equal(_pcm_crc(y), rec[f"glockenspiel_{rate}_orc_pcm_crc"])
    assert hashlib.md5(y.astype("<i2").tobytes()).hexdigest() == md5[f"oracle:glockenspiel_{rate}.pcm"]
    if key in rec.files:                                       # written by the reference's own decoder loop
        want = rec[key]
        n = int(rec[f"glockenspiel_{rate}_ref_pcm_n"]) if f"glockenspiel_{rate}_ref_pcm_n" in rec.files else want.shape[1] * 1024
        assert np.array_equal(_pcm_crc(y[:, :n]), want), "GPU decode differs from the reference decoder"
        assert hashlib.md5(y[:, :n].astype("<i2").tobytes()).hexdigest() == md5[f"ref:glockenspiel_{rate}.pcm"]
    else:
        pytest.fail("golden file lacks the reference decoder's PCM for glockenspiel")
