"""GPU parity of the file-level drop-ins: the reference's own driver loop
(coder/pacfile.py:674-756) run with audio_codec_b200.{pcmfile, pacfile, detect_transients}
in place of the reference modules must write the reference's .pac bytes and decode to the
reference's PCM (goldens written by the unmodified reference, tests/golden/make_golden.py)."""
import os
import wave

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _write_wav(path, fs, pcm):
    w = wave.open(str(path), "wb")
    w.setnchannels(pcm.shape[0])
    w.setsampwidth(2)
    w.setframerate(fs)
    w.writeframes(pcm.T.astype("<i2").tobytes())
    w.close()


def _drive(in_path, out_path, direction, data_rate=None, use_vq=True, use_sbr=None):
    """The body of encode_decode_test, verbatim in structure, on the drop-in classes."""
    from audio_codec_b200.detect_transients import parTransientDetect
    from audio_codec_b200.pacfile import PACFile
    from audio_codec_b200.pcmfile import PCMFile
    if direction == "Encode":
        inFile, outFile = PCMFile(str(in_path)), PACFile(str(out_path))
    else:
        inFile, outFile = PACFile(str(in_path)), PCMFile(str(out_path))
    codingParams = inFile.OpenForReading()
    if direction == "Encode":
        codingParams.nMDCTLines = 1024
        codingParams.nScaleBits = 4
        codingParams.nMantSizeBits = 12
        codingParams.targetBitsPerSample = data_rate / (codingParams.sampleRate / 1000)
        codingParams.useSBR = (data_rate < 128) if use_sbr is None else use_sbr
        codingParams.useVQ = use_vq
        codingParams.nSamplesPerBlock = codingParams.nMDCTLines
    else:
        codingParams.bitsPerSample = 16
    outFile.OpenForWriting(codingParams)
    lookahead = np.zeros((codingParams.nChannels, 2 * codingParams.nSamplesPerBlock))
    cur = last = False
    while True:
        if direction == "Encode":
            data = inFile.ReadDataBlock(codingParams)
            if not data:
                nxt = False
            else:
                lookahead = np.concatenate((np.copy(data), lookahead[:, codingParams.nSamplesPerBlock:]), axis=1)
                nxt = parTransientDetect(lookahead)
            outFile.WriteDataBlock(lookahead[:, :codingParams.nSamplesPerBlock], codingParams, lastTrans=last,
                                   curTrans=cur, nextTrans=nxt)
            last, cur = cur, nxt
            if not data:
                break
        else:
            data = inFile.ReadDataBlock(codingParams)
            if not data:
                break
            outFile.WriteDataBlock(data, codingParams)
    inFile.Close(codingParams)
    outFile.Close(codingParams)


@pytest.mark.parametrize("name,tag,kw", [("castanet", "128", dict(data_rate=128)),
                                         ("harpsichord", "96", dict(data_rate=96)),
                                         ("castanet", "bfp128", dict(data_rate=128, use_vq=False, use_sbr=False))])
def test_reference_driver_on_dropin_classes(tmp_path, name, tag, kw):
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    fs = int(ex[name + "_fs"])
    ref_pac = open(os.path.join(GOLDEN, f"{name}_{tag}.pac"), "rb").read()
    src = tmp_path / f"{name}.wav"
    _write_wav(src, fs, ex[name])
    pac = tmp_path / f"{name}.pac"
    _drive(src, pac, "Encode", **kw)
    got = open(pac, "rb").read()
    assert got == ref_pac, f"{name}_{tag}: .pac written through the drop-in PACFile differs from the reference's"
    wav = tmp_path / f"{name}_dec.wav"
    _drive(pac, wav, "Decode")
    w = wave.open(str(wav))
    y = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, 2).T
    g = np.load(os.path.join(GOLDEN, "decoded.npz"))[f"{name}_{tag}"]
    assert y.shape == g.shape and np.array_equal(y, g), f"{name}_{tag}: decoded PCM differs from the reference's"


def test_batched_file_api_matches_block_api(tmp_path):
    from audio_codec_b200 import pacfile
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    fs = int(ex["spmg_fs"])
    src = tmp_path / "spmg.wav"
    _write_wav(src, fs, ex["spmg"])
    for rate in (96, 128):
        pac, wav = pacfile.encode_decode_test(src, rate, str(tmp_path))
        assert open(pac, "rb").read() == open(os.path.join(GOLDEN, f"spmg_{rate}.pac"), "rb").read()
        w = wave.open(str(wav))
        y = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, 2).T
        g = np.load(os.path.join(GOLDEN, "decoded.npz"))[f"spmg_{rate}"]
        assert np.array_equal(y, g)


def test_batch_cli_encodes_directory_and_resumes(tmp_path):
    """python -m audio_codec_b200.batch: directory -> .pac files identical to the goldens, per-rank
    manifest, finished outputs skipped on a second run."""
    import json
    import wave
    from audio_codec_b200 import batch
    ex = np.load(os.path.join(GOLDEN, "excerpts.npz"))
    src, out = tmp_path / "in", tmp_path / "out"
    src.mkdir()
    names = ["castanet", "harpsichord", "spmg"]
    for n in names:
        with wave.open(str(src / f"{n}.wav"), "wb") as w:
            w.setnchannels(2); w.setsampwidth(2); w.setframerate(int(ex[f"{n}_fs"]))
            w.writeframes(np.ascontiguousarray(ex[n].T.astype("<i2")).tobytes())
    assert batch.main(["encode", str(src), str(out), "--rate", "128", "--decode"]) == 0
    dec = np.load(os.path.join(GOLDEN, "decoded.npz"))
    for n in names:
        assert (out / f"{n}_coded_128.pac").read_bytes() == open(os.path.join(GOLDEN, f"{n}_128.pac"), "rb").read()
        with wave.open(str(out / f"{n}_128.wav")) as w:
            y = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, 2).T
        g = dec[f"{n}_128"]
        assert np.array_equal(y[:, :g.shape[1]], g)
    rows = [json.loads(ln) for ln in open(out / "manifest.rank0.jsonl")]
    assert sorted(r["file"] for r in rows) == sorted(f"{n}.wav" for n in names) and all(r["err"] == 0 for r in rows)
    assert batch.main(["encode", str(src), str(out), "--rate", "128"]) == 0          # resume: nothing left to do
    assert len(open(out / "manifest.rank0.jsonl").read().splitlines()) == len(names)
    assert len(batch.merge_manifest(str(out))) == len(names)
