"""GPU parity: the fused device encoder vs the oracle / reference goldens (.pac bytes)."""
import os
import struct

import numpy as np
import pytest

from oracle import pac as opac

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _blocks(buf):
    nb = struct.unpack_from("<L", buf, 26)[0]
    pos = 30 + 2 * nb
    out = []
    while pos < len(buf):
        n = struct.unpack_from("<L", buf, pos)[0]
        out.append(buf[pos:pos + 4 + n])
        pos += 4 + n
    return buf[:30 + 2 * nb], out


def _report(name, got, ref):
    hg, bg = _blocks(got)
    hr, br = _blocks(ref)
    bad = [i for i, (x, y) in enumerate(zip(bg, br)) if x != y]
    return (f"{name}: header_eq={hg == hr} blocks {len(bg)} vs {len(br)}; mismatching channel-blocks "
            f"{len(bad)}/{len(br)} first={bad[:8]}")


@pytest.mark.parametrize("tag,kw", [("128", dict(rate_kbps=128)), ("96", dict(rate_kbps=96)),
                                    ("bfp128", dict(rate_kbps=128, use_vq=False, use_sbr=False))])
def test_excerpts_byte_identical_to_reference(tag, kw):
    from audio_codec_b200 import _lib, engine
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    names = [n for n in ("castanet", "harpsichord", "quar48_1", "spmg")
             if os.path.exists(os.path.join(GOLDEN, f"{n}_{tag}.pac"))]
    fs = int(ex[names[0] + "_fs"])
    ctx = _lib.Context(sample_rate=fs, n_channels=2, **kw)
    got = engine.encode_streams(ctx, [ex[n] for n in names])
    msgs, ok = [], True
    for n, g in zip(names, got):
        ref = open(os.path.join(GOLDEN, f"{n}_{tag}.pac"), "rb").read()
        msgs.append(_report(n, g, ref))
        ok &= g == ref
    print("\n".join(msgs))
    assert ok, "\n".join(msgs)
