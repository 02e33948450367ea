"""GPU: the batched pipeline (k_encode / k_bitalloc / k_pvq / k_pack and k_hdr / k_dequant / k_decode run
per batch of channel-blocks, scratch reused batch after batch) gives the same bytes and the same PCM
whatever the batch size, including batches that are not a multiple of 32 channel-blocks (partly filled
lane groups) and streams cut across batch boundaries.  The default-batch results are the ones the other
tests pin to the oracle and the reference goldens."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _streams():
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    names = [n for n in ("castanet", "harpsichord", "quar48_1", "spmg")]
    fs = int(ex[names[0] + "_fs"])
    base = [ex[n] for n in names]
    # ragged copies so that batch boundaries fall at different places of different streams
    rng = np.random.default_rng(7)
    out = list(base)
    for i in range(8):
        s = base[i % 4]
        n = int(rng.integers(3000, s.shape[1]))
        out.append(np.ascontiguousarray(s[:, :n]))
    return fs, out


@pytest.mark.parametrize("rate", [128, 96])
@pytest.mark.parametrize("batch", [40, 97, 1024])
def test_batch_size_invariance(rate, batch):
    from audio_codec_b200 import _lib, engine
    fs, streams = _streams()
    ctx = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=rate)
    ref_pac = engine.encode_streams(ctx, streams)
    ref_pcm = engine.decode_streams(ref_pac, ctx=ctx)
    ctx.set_option("batch_cb", batch)                       # pac_set_option(PAC_OPT_BATCH_CB)
    got_pac = engine.encode_streams(ctx, streams)
    got_pcm = engine.decode_streams(got_pac, ctx=ctx)       # the small-batch encoding through the small-batch decoder
    ctx.set_option("batch_cb", 0)
    for i, (a, b) in enumerate(zip(got_pac, ref_pac)):
        assert a == b, f"stream {i}: .pac differs with batches of {batch} channel-blocks"
    for i, (a, b) in enumerate(zip(got_pcm, ref_pcm)):
        assert np.array_equal(a, b), f"stream {i}: PCM differs with batches of {batch} channel-blocks"


def test_kernel_stats_and_launch_counter():
    from audio_codec_b200 import _lib, engine
    fs, streams = _streams()
    ctx = _lib.Context(sample_rate=fs, n_channels=2, rate_kbps=128)
    n0 = ctx.kernel_launches()
    pac = engine.encode_streams(ctx, streams[:4])
    n1 = ctx.kernel_launches()
    assert n1 - n0 >= 10                     # k_transient, k_encode, k_bitalloc, k_pvq, k_pack, sizes / scan / gather / headers
    ms, n = ctx.kernel_stats(0)
    assert n >= 1 and ms > 0.0               # k_encode ran and was timed
    ms, n = ctx.kernel_stats(1)
    assert n >= 1 and ms > 0.0               # k_pvq
    assert len(pac) == 4
