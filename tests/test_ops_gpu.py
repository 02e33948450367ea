"""GPU parity of the per-function drop-ins (audio_codec_b200.{window, detect_transients,
psychoac, bitalloc, quantize, gain_shape_quantize, bitpack, codec}) against the oracle on
seeded inputs.  Integer results must be identical; floating point within the stated tolerance."""
import os

import numpy as np
import pytest

from oracle import bits as obits
from oracle import codec as ocodec
from oracle import dsp, psycho, pvq

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _frames(n, N, seed, fs=48000):
    """A mix of noise, tones and onsets, in (-1, 1)."""
    rng = np.random.default_rng(seed)
    t = np.arange(N) / fs
    out = []
    for i in range(n):
        f0 = 110.0 * 2 ** rng.uniform(0, 6)
        x = 0.3 * np.sin(2 * np.pi * f0 * t + rng.uniform(0, 6)) + 0.1 * np.sin(2 * np.pi * 2.7 * f0 * t)
        x += 10 ** rng.uniform(-4, -1) * rng.standard_normal(N)
        if i % 3 == 0:
            x[N // 2:] *= 0.01
        out.append(np.clip(x, -0.99, 0.99))
    return np.array(out)


# ------------------------------------------------------------------ window.*
def _same_window(got, want):
    """Window tables are built on the host with long-double sin/cos/I0 and rounded once (correctly
    rounded); NumPy's SIMD sin/cos are within 1 ulp of that (3 of 2048 sine entries differ on the
    build machine), so a windowed sample may differ by one rounding: relative 4.5e-16."""
    np.testing.assert_allclose(got, want, rtol=4.5e-16, atol=0)
    assert np.count_nonzero(got != want) <= max(4, got.size // 256)


@pytest.mark.parametrize("N", [2048, 256])
def test_windows_match_oracle(N):
    from audio_codec_b200 import window
    x = _frames(3, N, 1)
    _same_window(window.SineWindow(x[0]), dsp.sine_window(N) * x[0])
    _same_window(window.HanningWindow(x[1]), dsp.hann_window(N) * x[1])
    np.testing.assert_allclose(window.KBDWindow(x[2]), x[2] * dsp.kaiser_window(N), rtol=1e-14, atol=0)
    _same_window(window.SineWindow(x), dsp.sine_window(N) * x)            # batched
    if N == 2048:
        _same_window(window.StartWindow(x[0], 2048, 256), dsp.start_window(2048) * x[0])
        _same_window(window.StopWindow(x[0], 2048, 256), dsp.stop_window(2048) * x[0])
        _same_window(window.StartStopWindow(x[0], 2048, 256), dsp.startstop_window(2048) * x[0])
        with pytest.raises(ValueError):
            window.StartWindow(x[0][:100], 2048, 256)


def test_windows_any_length_and_fp32():
    import torch
    from audio_codec_b200 import window
    rng = np.random.default_rng(2)
    for N in (12, 100, 1000):
        x = rng.uniform(-1, 1, N)
        np.testing.assert_allclose(window.SineWindow(x), dsp.sine_window(N) * x, rtol=0, atol=2e-16)
        np.testing.assert_allclose(window.KBDWindow(x, 3.0), x * dsp.kaiser_window(N, 3.0), rtol=1e-14, atol=1e-17)
    x32 = torch.rand(4, 2048, device="cuda") - 0.5
    y = window.HanningWindow(x32)
    assert y.dtype == torch.float32 and y.is_cuda
    np.testing.assert_allclose(y.cpu().numpy(), dsp.hann_window(2048) * x32.cpu().numpy().astype(np.float64), atol=1e-7)


# ------------------------------------------------------------------ parTransientDetect
def test_transient_detect_matches_oracle():
    import torch
    from audio_codec_b200 import detect_transients as dt
    rng = np.random.default_rng(3)
    blocks = []
    for i in range(64):
        b = 0.01 * rng.standard_normal((2, 2048))
        b[:, 1024:] = 0
        if i % 2:
            k = rng.integers(0, 1024)
            b[i % 2, k:k + 20] += rng.uniform(0.02, 0.5)
        blocks.append(b)
    blocks.append(np.zeros((2, 2048)))
    want = [dsp.transient_detect(b) for b in blocks]
    got = [dt.parTransientDetect(b) for b in blocks]
    assert [bool(w) for w in want] == [bool(g) for g in got]
    assert got[-1] == 0 and got[-1] is not False          # silent block returns 0, as the reference does
    batch = dt.parTransientDetect(torch.from_numpy(np.array(blocks)).cuda())
    assert [bool(v) for v in batch.cpu().numpy()] == [bool(w) for w in want]
    assert 8 < sum(bool(w) for w in want) < 60


# ------------------------------------------------------------------ psychoac.*
def test_psy_scalar_maps_and_band_tables(kats):
    from audio_codec_b200 import psychoac as pa, sbr
    f = np.array([5.0, 10.0, 50.0, 440.0, 1000.0, 3300.0, 12000.0, 20000.0])
    np.testing.assert_allclose(pa.Bark(f), psycho.bark(f.copy()), rtol=1e-14)
    np.testing.assert_allclose(pa.Thresh(f), psycho.thresh(f.copy()), rtol=1e-13)
    I = np.array([0.0, 1e-12, 1e-6, 0.5, 3.0])
    np.testing.assert_allclose(pa.SPL(I), psycho.spl(I.copy()), rtol=1e-14)
    np.testing.assert_allclose(pa.Intensity(np.array([-30.0, 0.0, 60.0, 96.0])), 10 ** ((np.array([-30.0, 0.0, 60.0, 96.0]) - 96) / 10),
                               rtol=1e-13)
    for fs in (44100, 48000):
        for n in (1024, 128):
            b = pa.ScaleFactorBands(pa.AssignMDCTLinesFromFreqLimits(n, fs))
            assert np.array_equal(b.nLines, kats[f"bands_{fs}_{n}"])
            o = psycho.Bands(psycho.assign_lines(n, fs))
            assert np.array_equal(b.lowerLine, o.lowerLine) and np.array_equal(b.upperLine, o.upperLine)
        b = pa.ScaleFactorBands(pa.AssignMDCTLinesFromFreqLimits(1024, fs))
        assert np.array_equal(sbr.omitted_bands(b), kats[f"omit_{fs}"])


@pytest.mark.parametrize("N,fs", [(2048, 48000), (256, 48000), (2048, 44100)])
def test_calc_smrs_matches_oracle(N, fs):
    """fp64 device path vs NumPy fp64: SMRs (dB) within 1e-9 absolute; north star asks 1e-5 relative."""
    from audio_codec_b200 import psychoac as pa
    x = _frames(12, N, 4, fs)
    bands_o = psycho.Bands(psycho.assign_lines(N // 2, fs))
    bands = pa.ScaleFactorBands(pa.AssignMDCTLinesFromFreqLimits(N // 2, fs))
    lines, scales, want, wmask = [], [], [], []
    for fr in x:
        X = dsp.mdct(dsp.sine_window(N) * fr, N // 2, N // 2)[:N // 2]
        s = dsp.scale_factor(np.max(np.abs(X)), 4)
        lines.append(X * (1 << s))
        scales.append(s)
        want.append(psycho.calc_smrs(fr.copy(), lines[-1].copy(), s, fs, bands_o))
        wmask.append(psycho.masked_threshold(fr.copy(), N // 2, fs))
    got = pa.CalcSMRs(x, np.array(lines), np.array(scales), fs, bands)          # batched
    np.testing.assert_allclose(got, np.array(want), rtol=0, atol=1e-9)
    one = pa.CalcSMRs(x[3], lines[3], scales[3], fs, bands)                      # reference call shape
    np.testing.assert_allclose(one, want[3], rtol=0, atol=1e-9)
    mask = pa.getMaskedThreshold(x, None, None, fs, bands)
    np.testing.assert_allclose(mask, np.array(wmask), rtol=0, atol=1e-9)


# ------------------------------------------------------------------ bitalloc.*
def test_bit_alloc_identical_to_oracle():
    from audio_codec_b200 import bitalloc
    rng = np.random.default_rng(5)
    for fs, n in ((48000, 1024), (48000, 128), (44100, 1024)):
        b = psycho.Bands(psycho.assign_lines(n, fs))
        nB = b.nBands
        smr = rng.uniform(-30, 40, (200, nB))
        smr[::7] -= 60                          # some frames with nothing above the mask
        budgets = rng.uniform(50, 3500, 200)
        want = np.array([psycho.bit_alloc(budgets[i], 16, nB, b.nLines, smr[i]) for i in range(200)])
        got = bitalloc.BitAlloc(budgets, 16, nB, b.nLines, smr)
        assert np.array_equal(got, want)
        one = bitalloc.BitAlloc(float(budgets[0]), 16, nB, b.nLines, smr[0])
        assert np.array_equal(one, want[0])
    b = psycho.Bands(psycho.assign_lines(1024, 48000))
    nl = b.nLines.copy()
    om = psycho.omitted_bands(b)
    got = bitalloc.BitAlloc_SBR(2000.0, 16, b.nBands, nl, smr[0][:b.nBands] if smr.shape[1] >= b.nBands else np.zeros(b.nBands), om)
    assert all(nl[i] == 1 for i in om)          # in-place edit of the caller's array, as in the reference
    assert got.shape == (b.nBands,)


# ------------------------------------------------------------------ quantize.*
def test_quantizers_identical_to_oracle(kats):
    from audio_codec_b200 import quantize as q
    v = kats["q_in"]
    assert [q.QuantizeUniform(a, 8) for a in v] == list(kats["q_u8"])
    assert np.array_equal(q.vQuantizeUniform(v, 12), kats["q_vu12"])
    assert np.array_equal(q.vDequantizeUniform(kats["q_vu12"], 12), kats["q_du12"])
    sc = [q.ScaleFactor(a, 3, 5) for a in v]
    assert sc == list(kats["q_scale"])
    assert [q.Mantissa(a, s, 3, 5) for a, s in zip(v, sc)] == list(kats["q_mant"])
    assert [q.Dequantize(s, int(m), 3, 5) for s, m in zip(sc, kats["q_mant"])] == list(kats["q_deq"])
    rng = np.random.default_rng(6)
    x = np.concatenate([rng.uniform(-1.2, 1.2, 4000), [0.0, 1.0, -1.0, 1e-9, -1e-9]])
    for nb in (1, 2, 5, 12, 16, 24, 31):
        c = q.vQuantizeUniform(x, nb)
        assert np.array_equal(c, dsp.vquantize_uniform(x, nb))
        assert np.array_equal(q.vDequantizeUniform(c, nb), dsp.vdequantize_uniform(c, nb))
    for nm in (2, 5, 9, 16):
        band = rng.uniform(-1, 1, 300) * 10 ** rng.uniform(-5, 0)
        s = q.ScaleFactor(float(np.max(np.abs(band))), 4, nm)
        assert s == dsp.scale_factor(np.max(np.abs(band)), 4, nm)
        m = q.vMantissa(band, s, 4, nm)
        assert np.array_equal(m, dsp.vmantissa(band, s, 4, nm))
        assert np.array_equal(q.vDequantize(s, m, 4, nm), dsp.vdequantize(s, m, 4, nm))
    assert q.MantissaVector is q.vMantissa


# ------------------------------------------------------------------ bitpack.PackedBits
def test_packed_bits_kat_and_roundtrip(kats):
    from audio_codec_b200.bitpack import PackedBits
    pb = PackedBits()
    pb.Size(2)
    for v, n in zip((3, 5, 11, 3, 1), (4, 3, 5, 3, 1)):
        pb.WriteBits(v, n)
    assert pb.GetPackedData() == bytes(kats["bitpack_kat"]) == b"\x3a\xb7"
    pb.ResetPointers()
    assert [pb.ReadBits(n) for n in (4, 3, 5, 3, 1)] == [3, 5, 11, 3, 1]
    rng = np.random.default_rng(7)
    widths = [int(w) for w in rng.integers(1, 80, 300)]
    vals = [int(rng.integers(0, 1 << 62)) & ((1 << w) - 1) if w <= 62 else (int(rng.integers(0, 1 << 62)) << (w - 62)) | 5
            for w in widths]
    nbytes = (sum(widths) + 7) // 8
    pb = PackedBits()
    pb.Size(nbytes)
    ow = obits.BitWriter(nbytes)
    for i, (v, w) in enumerate(zip(vals, widths)):
        pb.WriteBits(v, w)
        ow.WriteBits(v, w)
        if i == 100:
            _ = pb.data                       # force a flush in the middle (unaligned continuation)
    assert pb.GetPackedData() == ow.bytes()
    rd = PackedBits()
    rd.SetPackedData(ow.bytes())
    assert [rd.ReadBits(w) for w in widths[:40]] == vals[:40]
    assert rd.ReadFields([w for w in widths[40:] if w <= 64][:50]) is not None


# ------------------------------------------------------------------ gain_shape_quantize.*
@pytest.mark.parametrize("L,bits", [(13, 40), (19, 78), (38, 200), (107, 450), (363, 726), (363, 2500), (7, 9), (1, 6),
                                    (45, 580)])
def test_gain_shape_identical_to_oracle(L, bits):
    from audio_codec_b200 import gain_shape_quantize as gs
    from audio_codec_b200.bitpack import PackedBits
    rng = np.random.default_rng(L * 1000 + bits)
    for trial in range(6):
        x = rng.standard_normal(L) * 10 ** rng.uniform(-3, 0)
        if trial == 1 and L > 4:
            x[L // 2:] *= 1e-3
        if trial == 2 and L > 4:
            x[1:] = 0                           # a single pulse
        wi, wb = pvq.quantize_gain_shape(x.copy(), bits, 0)
        gi, gb = gs.quantize_gain_shape(x, bits)
        assert gb == [int(b) for b in wb], (L, bits, trial)
        assert gi == [int(i) for i in wi], (L, bits, trial)
        nbytes = (sum(gb) + 7) // 8 + 1
        pb = PackedBits()
        pb.Size(nbytes)
        ow = obits.BitWriter(nbytes)
        for v, w in zip(gi, gb):
            pb.WriteBits(v, w)
            ow.WriteBits(v, w)
        pb.ResetPointers()
        y = gs.dequantize_gain_shape(pb, bits, L)
        yo = pvq.dequantize_gain_shape(obits.BitReader(ow.bytes()), bits, L, 0)
        np.testing.assert_allclose(y, yo, rtol=1e-12, atol=1e-15)
        assert pb.iByte * 8 + pb.iBit == sum(gb)
    assert gs.quantize_gain_shape(np.zeros(L), bits) == ([0], [0])


# ------------------------------------------------------------------ codec.*
def _params(fs, rate, use_vq=True, use_sbr=None, short=False):
    from audio_codec_b200 import psychoac as pa, sbr
    from audio_codec_b200.audiofile import CodingParams
    p = CodingParams()
    p.sampleRate, p.nChannels = fs, 2
    p.nMDCTLines = 128 if short else 1024
    p.nScaleBits, p.nMantSizeBits = 4, 12
    p.targetBitsPerSample = rate / (fs / 1000)
    p.useVQ = use_vq
    p.useSBR = (rate < 128) if use_sbr is None else use_sbr
    p.sfBands = pa.ScaleFactorBands(pa.AssignMDCTLinesFromFreqLimits(1024, fs))
    p.sfBandsShort = pa.ScaleFactorBands(pa.AssignMDCTLinesFromFreqLimits(128, fs))
    p.omittedBands = sbr.omitted_bands(p.sfBands) if p.useSBR else []
    return p


@pytest.mark.parametrize("fs,rate,vq,sbr,short", [(48000, 128, True, False, False), (48000, 128, True, False, True),
                                                   (44100, 96, True, True, False), (48000, 128, False, False, False),
                                                   (48000, 128, False, False, True)])
def test_encode_decode_single_channel_matches_oracle(fs, rate, vq, sbr, short):
    from audio_codec_b200 import codec
    from audio_codec_b200.bitpack import PackedBits
    N = 256 if short else 2048
    x = _frames(5, N, 8, fs)
    po = ocodec.Params(fs, rate, 2, vq, sbr)
    flagsets = [(False, short, False), (True, short, False), (False, short, True), (True, short, True), (False, short, False)]
    for fr, (last, cur, nxt) in zip(x, flagsets):
        p = _params(fs, rate, vq, sbr, short)
        want = ocodec.encode_channel(fr.copy(), po, last, cur, nxt, sbr=sbr)
        fn = codec.EncodeSingleChannel_SBR if sbr else codec.EncodeSingleChannel
        got = fn(fr, p, last, cur, nxt)
        half = N // 2
        if vq:
            ba, idx, wid, scale = got
            assert scale == want["scale"]
            assert np.array_equal(ba, want["bitAlloc"])
            assert [list(zip(i, w)) for i, w in zip(idx, wid)] == [[(int(v), int(w)) for v, w in f] for f in want["fields"]]
            if sbr:
                assert all(p.sfBands.nLines[b] == 1 for b in p.omittedBands)
            # decode: band payloads in a PackedBits, as PACFile.getDecodedBlock hands them over; an SBR stream
            # carries only the LAST field of an omitted band (coder/pacfile.py:385-393)
            coded = [b for b in range(len(ba)) if ba[b]]
            om = set(int(b) for b in po.omittedBands) if sbr else set()
            sent = [list(zip(i, w))[-1:] if b in om else list(zip(i, w)) for b, i, w in zip(coded, idx, wid)]
            tot = sum(n for f in sent for _, n in f)
            pb = PackedBits()
            pb.Size((tot + 7) // 8 + 2)
            ow = obits.BitWriter((tot + 7) // 8 + 2)
            pb.WriteBits(5, 3)                                 # cursor not byte aligned
            ow.WriteBits(5, 3)
            for f in sent:
                for v, n in f:
                    pb.WriteBits(v, n)
                    ow.WriteBits(v, n)
            pb.ResetPointers()
            pb.ReadBits(3)
            rd = obits.BitReader(ow.bytes())
            rd.ReadBits(3)
            if sbr:                                            # coder/codec.py:95-222
                y = codec.Decode_SBR(None, ba, None, scale, pb, _params(fs, rate, vq, sbr, short), last, cur, nxt)
                yo = ocodec.decode_channel_sbr(rd, ba, scale, po, half, last, cur, nxt)
            else:
                y = codec.Decode(None, ba, None, scale, pb, _params(fs, rate, vq, sbr, short), last, cur, nxt)
                yo = ocodec.decode_channel_vq(rd, ba, scale, po, half, last, cur, nxt)
            np.testing.assert_allclose(y, yo, rtol=0, atol=1e-12)
            assert pb.iByte * 8 + pb.iBit == 3 + tot
        else:
            sf, ba, mant, scale = got
            assert scale == want["scale"]
            assert np.array_equal(ba, want["bitAlloc"])
            assert np.array_equal(sf, want["sf"])
            wm = np.concatenate(want["mant"]) if want["mant"] else np.empty(0)
            assert np.array_equal(mant, wm)
            bands = po.sfBandsShort if short else po.sfBands
            full = np.zeros(half, dtype=np.int64)
            k = 0
            for b in range(bands.nBands):
                if ba[b]:
                    n = int(bands.nLines[b])
                    full[bands.lowerLine[b]:bands.lowerLine[b] + n] = mant[k:k + n]
                    k += n
            y = codec.Decode(sf, ba, mant, scale, None, _params(fs, rate, vq, sbr, short), last, cur, nxt)
            yo = ocodec.decode_channel_bfp(ba, sf, full, scale, po, half, last, cur, nxt)
            np.testing.assert_allclose(y, yo, rtol=0, atol=1e-12)


def test_pcm_mapping_exact_for_every_code():
    """PCMFile.ReadDataBlock's int16 -> fraction mapping (coder/pcmfile.py:89-98) as the stream encoder applies
    it (multiply + one Newton correction instead of the division): all 65 536 codes equal the oracle's division."""
    import ctypes as C
    import torch
    from audio_codec_b200 import _lib
    codes = np.arange(-32768, 32768, dtype=np.int16)
    d = torch.from_numpy(codes).cuda()
    out = torch.empty(codes.size, dtype=torch.float64, device="cuda")
    ctx = _lib.default_context(48000)
    _lib.check(_lib.lib.pac_pcm_to_float(ctx.h, C.c_void_p(d.data_ptr()), codes.size, C.c_void_p(out.data_ptr()),
                                         _lib.stream_ptr()))
    assert np.array_equal(out.cpu().numpy(), dsp.pcm_to_float(codes))


def test_straight_line_transcendentals_are_bit_identical_to_the_library_calls():
    """The stream encoder evaluates log10 / atan through straight-line transcriptions of the library's main paths
    (so that several evaluations interleave).  They must give the library's bits: SPL and Bark through both forms
    (pac_psy_map ops 0 / 4 and 2 / 5) over 2 M inputs spanning the whole range the codec produces, plus edge values."""
    import ctypes as C
    import torch
    from audio_codec_b200 import _lib
    rng = np.random.default_rng(99)
    inten = np.concatenate([10.0 ** rng.uniform(-30, 6, 1 << 20), rng.uniform(0, 4, 1 << 16),
                            [0.0, 1e-8, 1.0, 2.0, 0.5, 1e-300, 4.0, np.sqrt(2.0), np.nextafter(np.sqrt(2.0), 0), 1e-16]])
    freq = np.concatenate([rng.uniform(0, 24000, 1 << 20), 10.0 ** rng.uniform(-3, 4.4, 1 << 16),
                           [0.0, 1315.789, 1315.79, 7500.0, 7499.999999, 24000.0, 1000.0 / 0.76]])
    ctx = _lib.default_context(48000)
    for x, a, b in ((inten, _lib.PSY_SPL, 4), (freq, _lib.PSY_BARK, 5)):
        d = torch.from_numpy(x).cuda()
        o1, o2 = torch.empty_like(d), torch.empty_like(d)
        for op, o in ((a, o1), (b, o2)):
            _lib.check(_lib.lib.pac_psy_map(ctx.h, op, C.c_void_p(d.data_ptr()), d.numel(), C.c_void_p(o.data_ptr()),
                                            _lib.stream_ptr()))
        assert torch.equal(o1.view(torch.int64), o2.view(torch.int64)), int((o1 != o2).sum())


def test_shared_reciprocal_division_equals_ieee_division():
    """pac_div (one correctly rounded reciprocal per divisor + residual correction, csrc/pac_common.cuh) is what
    the thread-per-band quantisers use for x / norm and x / sqrt(2): it must be the IEEE quotient bit for bit
    (Markstein).  2^27 pseudo-random pairs in four magnitude classes, two seeds."""
    from audio_codec_b200 import _lib
    ctx = _lib.default_context(48000)
    for seed in (1, 0xC0FFEE):
        bad = _lib.lib.pac_selftest_division(ctx.h, 1 << 26, seed, _lib.stream_ptr())
        assert bad == 0, f"{bad} quotients differ from the division instruction"
