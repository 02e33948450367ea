"""GPU parity: fused window+MDCT / IMDCT+OLA kernels vs the oracle (coder/mdct.py)."""
import numpy as np
import pytest

from oracle import dsp

pytestmark = pytest.mark.gpu

TOL = {np.float64: 1e-12, np.float32: 1e-5}   # relative to max|X| per frame


@pytest.mark.parametrize("N", [2048, 256])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_mdct_single_frame_matches_oracle(N, dt, kats):
    from audio_codec_b200 import mdct
    x = kats[f"x{N}"]
    xw = dsp.sine_window(N) * x
    ref = kats[f"mdct{N}"]
    got = mdct.MDCT(xw.astype(dt), N // 2, N // 2)
    assert got.dtype == dt
    assert np.max(np.abs(got - ref)) <= TOL[dt] * np.max(np.abs(ref))
    back = mdct.IMDCT(ref.astype(dt), N // 2, N // 2)
    refb = kats[f"imdct{N}"]
    assert np.max(np.abs(back - refb)) <= TOL[dt] * np.max(np.abs(refb))


def test_mdct_reference_kat_sizes_rejected():
    from audio_codec_b200 import mdct
    with pytest.raises(ValueError):
        mdct.MDCT(np.zeros(8), 4, 4)


@pytest.mark.parametrize("N", [2048, 256])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("win", ["sine", "kaiser"])
def test_batched_frames_match_oracle(N, dt, win):
    from audio_codec_b200 import _lib, mdct
    rng = np.random.default_rng(7)
    M = N // 2
    rows, F = 3, 11                      # ragged vs. frames-per-CTA (4 / 16)
    x = rng.uniform(-1, 1, (rows, (F + 1) * M))
    w = dsp.sine_window(N) if win == "sine" else dsp.kaiser_window(N)
    wid = _lib.W_SINE if win == "sine" else _lib.W_KAISER
    got = mdct.mdct_frames(x.astype(dt), N, wid)
    assert got.shape == (rows, F, M)
    for r in range(rows):
        for f in range(F):
            ref = dsp.mdct(w * x[r, f * M:f * M + N], M, M)
            assert np.max(np.abs(got[r, f] - ref)) <= TOL[dt] * np.max(np.abs(ref)), (r, f)


@pytest.mark.parametrize("N", [2048, 256])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_tdac_roundtrip_sine(N, dt):
    """Sine window analysis -> synthesis -> overlap-add reconstructs the interior."""
    from audio_codec_b200 import _lib, mdct
    rng = np.random.default_rng(3)
    M = N // 2
    rows, F = 2, 9
    x = rng.uniform(-1, 1, (rows, (F + 1) * M)).astype(dt)
    X = mdct.mdct_frames(x, N, _lib.W_SINE)
    y = mdct.imdct_ola(X, N, _lib.W_SINE)
    assert y.shape == x.shape
    err = np.max(np.abs(y[:, M:F * M] - x[:, M:F * M]))
    assert err <= (1e-12 if dt == np.float64 else 2e-6)


@pytest.mark.parametrize("shape", ["start", "stop", "startstop"])
def test_edler_windows(shape):
    from audio_codec_b200 import _lib, mdct
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (2, 4 * 1024))
    w = {"start": dsp.start_window, "stop": dsp.stop_window, "startstop": dsp.startstop_window}[shape](2048)
    wid = {"start": _lib.W_START, "stop": _lib.W_STOP, "startstop": _lib.W_STARTSTOP}[shape]
    got = mdct.mdct_frames(x, 2048, wid)
    for r in range(2):
        for f in range(3):
            ref = dsp.mdct(w * x[r, f * 1024:f * 1024 + 2048], 1024, 1024)
            assert np.max(np.abs(got[r, f] - ref)) <= 1e-12 * np.max(np.abs(ref))


def test_empty_batch():
    from audio_codec_b200 import mdct
    out = mdct.MDCT(np.zeros((0, 2048)), 1024, 1024)
    assert out.shape == (0, 1024)
