"""GPU: the earlier forms of a stage that pac_set_option(ctx, PAC_OPT_VARIANT, bits) keeps selectable for A/B timing
give the SAME bytes and samples as the current kernels (include/pac_b200.h: bit 0 the 256-thread front end, bit 1 the
exhaustive masked threshold, bit 3 the LDG / STG streaming MDCT, bit 4 the CTA-per-block decoder back end)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _streams():
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    names = ("castanet", "harpsichord", "quar48_1", "spmg")
    return int(ex["castanet_fs"]), [ex[n] for n in names] + [np.ascontiguousarray(ex["castanet"][:, :5000])]


@pytest.mark.parametrize("rate", [128, 96])
@pytest.mark.parametrize("bits", [1, 2, 3, 16, 19])
def test_codec_variants_give_identical_streams_and_pcm(rate, bits):
    from audio_codec_b200 import _lib, engine
    fs, streams = _streams()
    ctx = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=rate)
    ref_pac = engine.encode_streams(ctx, streams)
    ref_pcm = engine.decode_streams(ref_pac, ctx=ctx)
    ctx.set_option("variant", bits)
    got_pac = engine.encode_streams(ctx, streams)
    got_pcm = engine.decode_streams(got_pac, ctx=ctx)
    ctx.set_option("variant", 0)
    for i, (a, b) in enumerate(zip(got_pac, ref_pac)):
        assert a == b, f"stream {i}: .pac differs with variant bits {bits}"
    for i, (a, b) in enumerate(zip(got_pcm, ref_pcm)):
        assert np.array_equal(a, b), f"stream {i}: PCM differs with variant bits {bits}"


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_streaming_mdct_variants_agree(dtype):
    """bulk-copy pipeline (mdct_tma.cuh, packed fp32 arithmetic, factors folded into the window copy) against the LDG /
    STG pipeline (mdct_stream.cuh): the same butterflies; the fold's products may contract into FMAs differently, so
    the comparison is to a few ulps of the largest coefficient, not bit for bit."""
    import torch
    from audio_codec_b200 import _lib, mdct
    g = torch.Generator(device="cuda").manual_seed(7)
    x = (0.25 * torch.randn((3, 41 * 1024), generator=g, device="cuda", dtype=torch.float32)).to(getattr(torch, dtype))
    tol = 4e-6 if dtype == "float32" else 1e-14
    out = {}
    for bits in (0, 8):
        ctx = _lib.default_context()
        ctx.set_option("variant", bits)
        for w in (_lib.W_SINE, _lib.W_KAISER):
            X = mdct.mdct_frames(x, 2048, w, ctx)
            y = mdct.imdct_ola(X, 2048, w, ctx)
            out[(bits, w)] = (X.double().cpu().numpy(), y.double().cpu().numpy())
    for w in (_lib.W_SINE, _lib.W_KAISER):
        for k, what in ((0, "forward MDCT"), (1, "IMDCT + overlap-add")):
            a, b = out[(0, w)][k], out[(8, w)][k]
            assert np.abs(a - b).max() <= tol * np.abs(b).max(), f"{what} differs between the two pipelines (window {w})"
