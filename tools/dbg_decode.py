import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import test_ops_gpu as t
from audio_codec_b200 import codec
from oracle import codec as ocodec, dsp as omdct
fs, rate, vq, sbr, short = 48000, 128, False, False, False
N = 2048; half = 1024
x = t._frames(5, N, 8, fs)
po = ocodec.Params(fs, rate, 2, vq, sbr)
flagsets = [(False, short, False), (True, short, False), (False, short, True), (True, short, True), (False, short, False)]
for fr, (last, cur, nxt) in zip(x, flagsets):
    p = t._params(fs, rate, vq, sbr, short)
    sf, ba, mant, scale = codec.EncodeSingleChannel(fr, p, last, cur, nxt)
    bands = po.sfBands
    full = np.zeros(half, dtype=np.int64); k = 0
    for b in range(bands.nBands):
        if ba[b]:
            n = int(bands.nLines[b]); full[bands.lowerLine[b]:bands.lowerLine[b] + n] = mant[k:k + n]; k += n
    y = codec.Decode(sf, ba, mant, scale, None, t._params(fs, rate, vq, sbr, short), last, cur, nxt)
    yo = ocodec.decode_channel_bfp(ba, sf, full, scale, po, half, last, cur, nxt)
    d = np.abs(y - yo)
    print("flags", last, cur, nxt, "scale", scale, "max diff", d.max(), "ratio y/yo med", np.median(y[500:1500] / yo[500:1500]))
    Y = omdct.mdct(y, half, half) ; YO = omdct.mdct(yo, half, half)
    dd = np.abs(Y - YO); bad = np.nonzero(dd > 1e-9 * np.abs(YO).max())[0]
    print("  mdct-domain differing lines:", len(bad), bad[:10], bad[-5:] if len(bad) else "")
