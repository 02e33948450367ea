"""GPU: the fp32 fast path (precision="f32") against the fp64 reference-exact path.

Tolerances (north star): MDCT lines and SMRs within 1e-5 relative (of the signal's / the SMR
range's scale), decoded PCM within 1 LSB; the .pac written by the fp32 encoder must be a valid
stream for the fp64 reference decoder (oracle), and the record mismatch rate is reported."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
NAMES = ("castanet", "harpsichord", "quar48_1", "spmg")
SMR_RTOL = 1e-5
PCM_LSB = 1


def test_smr_fp32_within_tolerance():
    import torch
    from audio_codec_b200 import _lib
    k = np.load(os.path.join(GOLDEN, "kats.npz"))
    fs = int(k["fs"])
    frames, mdct, scale = k["frames"], k["frame_mdct"], k["frame_scale"]
    lo = np.concatenate([[0], np.cumsum(k[f"bands_{fs}_1024"])[:-1]]).astype(np.int32)
    nl = k[f"bands_{fs}_1024"].astype(np.int32)
    ctx = _lib.default_context(fs, n_channels=2)
    d = torch.from_numpy(frames).cuda()
    m = torch.from_numpy(mdct).cuda()
    s = torch.from_numpy(scale.astype(np.int32)).cuda()
    out = {}
    import ctypes as C
    for name, dt in (("f32", 0), ("f64", 1)):
        smr = torch.empty(len(frames), len(nl), dtype=torch.float64, device="cuda")
        _lib.check(_lib.lib.pac_calc_smr(ctx.h, dt, C.c_void_p(d.data_ptr()), C.c_void_p(m.data_ptr()),
                                         C.c_void_p(s.data_ptr()), len(frames), 2048, len(nl),
                                         (C.c_int * len(nl))(*lo), (C.c_int * len(nl))(*nl),
                                         C.c_void_p(smr.data_ptr()), None, _lib.stream_ptr()))
        out[name] = smr.cpu().numpy()
    assert np.array_equal(out["f64"], k["frame_smr"]) or np.allclose(out["f64"], k["frame_smr"], rtol=0, atol=1e-9)
    span = np.abs(k["frame_smr"]).max()
    err = np.abs(out["f32"] - k["frame_smr"]).max()
    print(f"SMR fp32: max abs err {err:.3e} dB over a {span:.1f} dB range (rel {err / span:.2e})")
    os.makedirs("gpurun_out", exist_ok=True)            # kept with the run's evidence (copied to profiles/)
    with open(os.path.join("gpurun_out", "smr_fp32_error.txt"), "w") as f:
        f.write(f"SMR fp32 vs reference kats: max abs err {err:.3e} dB, range {span:.1f} dB, rel {err / span:.2e}, "
                f"tolerance {SMR_RTOL:g}\n")
    assert err <= SMR_RTOL * span, f"fp32 SMR error {err} dB exceeds {SMR_RTOL} x {span} dB"


@pytest.mark.parametrize("tag", ["128", "96", "bfp128"])
def test_decode_fp32_within_one_lsb(tag):
    from audio_codec_b200 import engine
    dec = np.load(os.path.join(GOLDEN, "decoded.npz"))
    names = [n for n in NAMES if os.path.exists(os.path.join(GOLDEN, f"{n}_{tag}.pac"))]
    pacs = [open(os.path.join(GOLDEN, f"{n}_{tag}.pac"), "rb").read() for n in names]
    got = engine.decode_streams(pacs, precision="f32")
    for n, y in zip(names, got):
        g = dec[f"{n}_{tag}"]
        assert y.shape[1] >= g.shape[1]
        d = np.abs(y[:, :g.shape[1]].astype(np.int32) - g.astype(np.int32))
        print(f"{n}_{tag}: fp32 decode max |diff| {d.max()} LSB, differing {np.count_nonzero(d) / d.size:.2%}")
        assert d.max() <= PCM_LSB


@pytest.mark.parametrize("rate", [128, 96])
def test_encode_fp32_is_valid_stream_and_close(rate):
    """fp32-encoded .pac: same size class, decodable by the fp64 oracle, audio close to the fp64 encoding."""
    from audio_codec_b200 import _lib, engine
    from oracle import pac as opac
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    rates = []
    for n in NAMES:
        fs = int(ex[f"{n}_fs"])
        pcm = ex[n]
        c32 = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=rate, precision="f32")
        c64 = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=rate)
        a = engine.encode_streams(c32, [pcm])[0]
        b = engine.encode_streams(c64, [pcm])[0]
        r = engine.record_mismatch_rate(a, b)
        rates.append(r)
        assert abs(len(a) - len(b)) <= 0.01 * len(b) + 64
        _, ya = opac.decode_pac(a)          # the reference decoder accepts the fp32 encoder's stream
        yb = engine.decode_streams([b])[0]
        m = min(ya.shape[1], yb.shape[1])
        assert ya.shape == yb.shape
        # both are lossy encodings of pcm: compare each one's coding noise, not sample equality
        na = np.mean((ya[:, :m].astype(float) - _aligned(pcm, m)) ** 2)
        nb = np.mean((yb[:, :m].astype(float) - _aligned(pcm, m)) ** 2)
        print(f"{n}@{rate}: record mismatch {r:.2%}, noise power fp32 {na:.1f} vs fp64 {nb:.1f}")
        assert na <= 1.05 * nb + 1.0
    print(f"mean record mismatch rate @{rate} kb/s: {np.mean(rates):.2%}")


def _aligned(pcm, m):
    """The decoder's output is the input delayed by nothing (the codec pads a block up front and
    drops it); pad the input with zeros to m samples."""
    out = np.zeros((pcm.shape[0], m))
    k = min(m, pcm.shape[1])
    out[:, :k] = pcm[:, :k]
    return out
