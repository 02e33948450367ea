"""GPU parity: the device decoder vs the reference-decoded goldens (int16 PCM)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("tag", ["128", "96", "bfp128"])
def test_decode_matches_reference_pcm(tag):
    from audio_codec_b200 import engine
    dec = np.load(os.path.join(GOLDEN, "decoded.npz"))
    names = [n for n in ("castanet", "harpsichord", "quar48_1", "spmg")
             if os.path.exists(os.path.join(GOLDEN, f"{n}_{tag}.pac"))]
    pacs = [open(os.path.join(GOLDEN, f"{n}_{tag}.pac"), "rb").read() for n in names]
    got = engine.decode_streams(pacs)
    msgs, ok = [], True
    for n, y in zip(names, got):
        g = dec[f"{n}_{tag}"]
        m = min(g.shape[1], y.shape[1])
        d = np.abs(y[:, :m].astype(np.int32) - g[:, :m].astype(np.int32))
        msgs.append(f"{n}_{tag}: samples {m} (decoded {y.shape[1]}), max |diff| {d.max()} LSB, "
                    f"mismatching {np.count_nonzero(d)}")
        ok &= d.max() == 0 and y.shape[1] >= g.shape[1]
    print("\n".join(msgs))
    assert ok, "\n".join(msgs)


def test_encode_decode_roundtrip_48k_identical_to_oracle():
    """48 kHz (BASELINE config) has no reference fixture: compare with the oracle."""
    from audio_codec_b200 import _lib, engine
    from oracle import pac as opac
    rng = np.random.default_rng(422)
    n = 12 * 1024 + 77
    t = np.arange(n) / 48000.0
    env = np.where((t * 7) % 1 < 0.1, 1.0, 0.05)
    sig = 0.4 * env * np.sin(2 * np.pi * 440 * t) + 0.01 * rng.standard_normal(n)
    pcm = np.stack([sig, 0.8 * sig + 0.005 * rng.standard_normal(n)])
    pcm = np.clip(np.round(pcm * 32767), -32768, 32767).astype(np.int16)
    for rate in (128, 96):
        ctx = _lib.Context(sample_rate=48000, n_channels=2, rate_kbps=rate)
        got = engine.encode_streams(ctx, [pcm])[0]
        ref = opac.encode_pcm(pcm, 48000, rate)
        assert got == ref, f"48k {rate}: .pac differs from oracle"
        y = engine.decode_streams([got])[0]
        _, yr = opac.decode_pac(ref)
        assert y.shape == yr.shape
        assert np.array_equal(y, yr), f"48k {rate}: max diff {np.abs(y.astype(int) - yr.astype(int)).max()}"


def test_mixed_configurations_in_one_call_are_refused():
    """One decode call shares one context (k_walk skips a fixed header size, the band tables are the context's):
    streams of different sample rate / SBR flag in one call raise instead of decoding to garbage; a header whose
    band table differs from the context's is refused too."""
    from audio_codec_b200 import _lib, engine
    a = open(os.path.join(GOLDEN, "castanet_128.pac"), "rb").read()
    b = open(os.path.join(GOLDEN, "castanet_96.pac"), "rb").read()        # SBR stream
    with pytest.raises(_lib.PacError, match="use_sbr"):
        engine.decode_streams([a, b])
    bad = bytearray(a)
    bad[30] ^= 1                                                           # first band's line count
    with pytest.raises(_lib.PacError, match="n_lines|band"):
        engine.decode_streams([a, bytes(bad)])
    with pytest.raises(_lib.PacError, match="band"):
        engine.decode_streams([bytes(bad)])
    bad = bytearray(a)
    bad[18] = 6                                                            # nScaleBits = 6: unsupported by the device quantisers
    with pytest.raises(_lib.PacError, match="unsupported"):
        engine.decode_streams([bytes(bad)])


def test_legacy_seven_field_generation_decodes_to_the_reference_wav(tmp_path):
    """the reference ships files of older bitstream generations (README.md:4-5).  The 7-field
    header generation (test_decoded/: no useVQ field; block floating point with block switching, nMantSizeBits = 16)
    decodes bit-exactly to the .wav that sits beside it in the reference: first 60 blocks of
    test_decoded/castanet_coded_128.pac against the first 61 440 samples of test_decoded/castanet_128.wav, through
    the batched decoder and through the PACFile block API.  The 6-field generation (25 unmerged bands) is refused."""
    from audio_codec_b200 import _lib, engine, pacfile
    buf = open(os.path.join(GOLDEN, "legacy7_castanet_128.pac"), "rb").read()
    want = np.load(os.path.join(GOLDEN, "legacy7_castanet_128_pcm.npz"))["pcm"]
    h = engine.parse_header(buf)
    assert h["layout"] == 7 and not h["use_vq"] and not h["use_sbr"] and h["n_mant_size_bits"] == 16 and len(h["n_lines"]) == 18
    y = engine.decode_streams([buf])[0]
    assert y.shape[1] >= want.shape[1] and np.array_equal(y[:, :want.shape[1]], want)
    # block API: the reference's decode loop (coder/pacfile.py:758-776) on the drop-in classes
    from audio_codec_b200.pcmfile import PCMFile
    src = tmp_path / "legacy.pac"
    src.write_bytes(buf)
    inF, outF = pacfile.PACFile(str(src)), PCMFile(str(tmp_path / "legacy.wav"))
    cp = inF.OpenForReading()
    cp.bitsPerSample = 16
    cp.numSamples = want.shape[1]
    outF.OpenForWriting(cp)
    while True:
        data = inF.ReadDataBlock(cp)
        if not data:
            break
        outF.WriteDataBlock(data, cp)
    inF.Close(cp)
    outF.Close(cp)
    import wave
    with wave.open(str(tmp_path / "legacy.wav")) as w:
        z = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, 2).T
    assert np.array_equal(z[:, :want.shape[1]], want)
    six = buf[:22] + (25).to_bytes(4, "little") + b"".join(int(v).to_bytes(2, "little") for v in
          (5, 4, 5, 5, 5, 5, 7, 7, 7, 9, 10, 11, 13, 15, 17, 21, 26, 32, 42, 51, 61, 83, 116, 163, 304)) + b"\0" * 64
    assert engine.parse_header(six)["layout"] == 6
    with pytest.raises(_lib.PacError, match="6-field"):
        engine.decode_streams([six])
