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


def test_reference_self_test_runs_on_the_drop_in(kats):
    """The reference's own self-test (coder/mdct.py:86-107), verbatim in structure, on the drop-in:
    hop-4 frames of N = 8 through MDCTslow / IMDCT reconstruct the signal (TDAC), MDCT == MDCTslow on the
    20-sample signal (a = b = 10), and the reference's known answer for N = 8 (kats: mdct_kat_*)."""
    from audio_codec_b200 import mdct
    x = np.array([0, 1, 2, 3, 4, 4, 4, 4, 3, 1, -1, -3], dtype=float)
    frame_size, noverlap = 8, 4
    x = np.concatenate([np.zeros(noverlap), x, np.zeros(noverlap)])
    x_hat = np.zeros_like(x)
    for i in range(0, len(x) - noverlap, noverlap):
        data = x[i:i + frame_size]
        X = mdct.MDCTslow(data, 4, 4, isInverse=False)
        assert np.allclose(X, dsp.mdct(data, 4, 4), rtol=0, atol=1e-14)
        x_hat[i:i + frame_size] += mdct.MDCTslow(X, 4, 4, isInverse=True) / 2
    assert np.allclose(x, x_hat)
    a = b = len(x) // 2
    Xa = mdct.MDCTslow(x, a, b, isInverse=False)
    Xb = mdct.MDCT(x, a, b, isInverse=False)
    assert np.allclose(Xa, Xb) and np.allclose(Xa, dsp.mdct(x, a, b), rtol=0, atol=1e-14)
    assert np.allclose(mdct.MDCTslow(Xa, a, b, isInverse=True), mdct.MDCT(Xb, a, b, isInverse=True))
    assert np.allclose(mdct.IMDCT(Xb, a, b), dsp.imdct(Xb, a, b), rtol=0, atol=1e-13)
    assert np.allclose(mdct.MDCT(kats["mdct_kat_in"], 4, 4), kats["mdct_kat_out"], rtol=0, atol=1e-14)


@pytest.mark.parametrize("a,b", [(4, 4), (10, 10), (1024, 128), (128, 1024), (6, 2), (33, 31), (500, 500)])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_any_block_split_matches_oracle(a, b, dt):
    """mdct.MDCT(data, a, b) for splits the FFT kernels do not cover (a != b, other N): direct-sum kernel."""
    from audio_codec_b200 import mdct
    rng = np.random.default_rng(a * 1000 + b)
    N = a + b
    x = rng.uniform(-1, 1, (3, N))
    ref = np.stack([dsp.mdct(r, a, b) for r in x])
    got = mdct.MDCT(x.astype(dt), a, b)
    assert got.shape == (3, N // 2) and got.dtype == dt
    tol = 1e-12 if dt == np.float64 else 1e-5
    assert np.max(np.abs(got - ref)) <= tol * np.max(np.abs(ref))
    back = mdct.IMDCT(ref.astype(dt), a, b)
    refb = np.stack([dsp.imdct(r, a, b) for r in ref])
    assert np.max(np.abs(back - refb)) <= tol * np.max(np.abs(refb))


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
