"""GPU robustness: corrupted and truncated .pac streams must neither hang nor fault the decoder, and
must not disturb intact streams decoded in the same batch (one CTA per channel-block, no shared state)."""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _decode(engine, ctx, bufs):
    import torch
    off = np.concatenate([[0], np.cumsum([len(b) for b in bufs])])
    d = torch.from_numpy(np.frombuffer(b"".join(bufs), dtype=np.uint8).copy()).cuda()
    d_pcm, po, err = engine.decode_device(ctx, d, off)
    torch.cuda.synchronize()
    pcm = d_pcm.cpu().numpy()
    return [pcm[int(po[s]):int(po[s + 1])] for s in range(len(bufs))], err


@pytest.mark.parametrize("tag", ["128", "96", "bfp128"])
def test_corrupted_streams_are_contained(tag):
    from audio_codec_b200 import engine
    good = open(os.path.join(GOLDEN, f"castanet_{tag}.pac"), "rb").read()
    h = engine.parse_header(good)
    ctx = engine.context_for_header(h)
    hdr = 30 + 2 * len(h["n_lines"])
    ref, err = _decode(engine, ctx, [good])
    assert err == 0
    rng = np.random.default_rng(20251104)
    bad = []
    for i in range(24):
        b = bytearray(good)
        kind = i % 4
        if kind == 0:                                     # random bytes in the payload
            for p in rng.integers(hdr, len(b), size=1 + i):
                b[p] = int(rng.integers(0, 256))
        elif kind == 1:                                   # truncation at an arbitrary byte
            b = b[:int(rng.integers(hdr + 1, len(b)))]
        elif kind == 2:                                   # a length prefix blown up / zeroed
            p = hdr
            for _ in range(int(rng.integers(0, 20))):
                n = int.from_bytes(b[p:p + 4], "little")
                if p + 4 + n + 4 > len(b):
                    break
                p += 4 + n
            b[p:p + 4] = (0xFFFFFFFF if i % 8 == 2 else 0).to_bytes(4, "little")
        else:                                             # all-ones / all-zero payload runs
            p = int(rng.integers(hdr, len(b) - 300))
            b[p:p + 256] = bytes([0xFF if i % 8 == 3 else 0]) * 256
        bad.append(bytes(b))
    batch = [good] + bad + [good]
    out, err = _decode(engine, ctx, batch)                # must return; soft error flags are fine
    assert len(out) == len(batch)
    assert np.array_equal(out[0], ref[0]) and np.array_equal(out[-1], ref[0]), "intact streams disturbed by their neighbours"
    for o in out:
        assert o.dtype == np.int16 and o.size % (1024 * h["n_channels"]) == 0
    # and the context is still usable afterwards
    again, err2 = _decode(engine, ctx, [good])
    assert err2 == 0 and np.array_equal(again[0], ref[0])
