import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import test_ops_gpu as t
from audio_codec_b200 import codec, psychoac, mdct as gm
from oracle import codec as ocodec, psycho as opsy, dsp
fs, rate = 48000, 128
po = ocodec.Params(fs, rate, 2, True, False)
dc = 12345 / 32768.0
frames = [np.concatenate([np.zeros(1024), np.full(1024, dc)]), np.full(2048, dc), np.concatenate([np.full(1024, dc), np.zeros(1024)])]
for fi, fr in enumerate(frames):
    p = t._params(fs, rate, True, False, False)
    want = ocodec.encode_channel(fr.copy(), po, False, False, False, sbr=False)
    ba, idx, wid, scale = codec.EncodeSingleChannel(fr, p, False, False, False)
    print("frame", fi, "scale", scale, want["scale"], "ba equal", np.array_equal(ba, want["bitAlloc"]))
    if not np.array_equal(ba, want["bitAlloc"]):
        print("  ba gpu", list(ba)); print("  ba ref", list(want["bitAlloc"]))
    # SMR stage
    win = dsp.sine_window(2048) if hasattr(dsp, "sine_window") else np.sin(np.pi * (np.arange(2048) + 0.5) / 2048)
    X = dsp.mdct(win * fr, 1024, 1024)
    sc = want["scale"]
    smr_o = opsy.calc_smrs(fr, X * (1 << sc), sc, fs, po.sfBands)
    smr_g = psychoac.CalcSMRs(fr, X * (1 << sc), sc, fs, p.sfBands)
    d = np.abs(np.asarray(smr_g) - np.asarray(smr_o))
    print("  smr max diff", d.max(), "at band", int(d.argmax()), "gpu", np.asarray(smr_g)[int(d.argmax())], "ref", np.asarray(smr_o)[int(d.argmax())])
    f_g = [list(zip(i, w)) for i, w in zip(idx, wid)]
    f_o = [[(int(v), int(w)) for v, w in f] for f in want["fields"]]
    bad = [b for b, (a, c) in enumerate(zip(f_g, f_o)) if a != c]
    print("  bands with differing fields:", bad)
