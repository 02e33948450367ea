import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
import test_ops_gpu as t
from audio_codec_b200 import codec, bitpack
from audio_codec_b200.bitpack import PackedBits
from oracle import codec as ocodec, bits as obits
fs, rate, vq, sbr, short = 48000, 128, True, False, False
N = 2048; half = 1024
x = t._frames(5, N, 8, fs)
po = ocodec.Params(fs, rate, 2, vq, sbr)
flagsets = [(False, short, False), (True, short, False), (False, short, True), (True, short, True), (False, short, False)]
for fr, (last, cur, nxt) in zip(x, flagsets):
    p = t._params(fs, rate, vq, sbr, short)
    ba, idx, wid, scale = codec.EncodeSingleChannel(fr, p, last, cur, nxt)
    tot = sum(sum(w) for w in wid)
    pb = PackedBits(); pb.Size((tot + 7) // 8 + 2)
    ow = obits.BitWriter((tot + 7) // 8 + 2)
    pb.WriteBits(5, 3); ow.WriteBits(5, 3)
    for i, w in zip(idx, wid):
        for v, n in zip(i, w):
            pb.WriteBits(v, n); ow.WriteBits(v, n)
    pb.ResetPointers(); pb.ReadBits(3)
    data = pb.device_data().cpu().numpy()
    ref = np.frombuffer(ow.bytes(), dtype=np.uint8)
    print("packed bytes equal:", np.array_equal(data[:len(ref)], ref), len(data), len(ref))
    # unpack check
    start = 3; avail = len(data) * 8 - start; n64 = (avail + 63) // 64
    ch = bitpack.unpack_rows(pb.device_data().reshape(1, -1), [start], [[64] * n64], [n64])[0].cpu().numpy().astype(np.uint64)
    bitsr = np.unpackbits(np.concatenate([data, np.zeros(16, np.uint8)]))[start:start + 64 * n64].reshape(n64, 64)
    want = np.array([int("".join(map(str, r)), 2) for r in bitsr], dtype=np.uint64)
    print("unpack chunks equal:", np.array_equal(ch[:n64], want), np.nonzero(ch[:n64] != want)[0][:5])
    rd = obits.BitReader(ow.bytes()); rd.ReadBits(3)
    y = codec.Decode(None, ba, None, scale, pb, t._params(fs, rate, vq, sbr, short), last, cur, nxt)
    yo = ocodec.decode_channel_vq(rd, ba, scale, po, half, last, cur, nxt)
    print("flags", last, cur, nxt, "max diff", np.abs(y - yo).max(), "cursor", pb.iByte * 8 + pb.iBit, 3 + tot)
