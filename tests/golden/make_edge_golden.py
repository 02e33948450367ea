"""Edge-case goldens from the UNMODIFIED reference (build container only, needs /root/reference):
    python tests/golden/make_edge_golden.py
For every signal of tests/edge_signals.py the reference's own pacfile.encode_decode_test runs at 128
and 96 kb/s; outputs tests/golden/edge.npz: '<name>_<rate>_pac' (uint8) and '<name>_<rate>_pcm' (int16).
Cases the reference itself cannot process are recorded as '<name>_<rate>_error' (message)."""
import os
import sys
import tempfile
import wave
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
sys.path.insert(0, str(HERE.parent))
from oracle import refload  # noqa: E402
import edge_signals  # noqa: E402


def write_wav(path, fs, pcm):
    w = wave.open(str(path), "wb")
    w.setnchannels(pcm.shape[0]); w.setsampwidth(2); w.setframerate(fs)
    w.writeframes(pcm.T.astype("<i2").tobytes()); w.close()


def read_wav(path):
    w = wave.open(str(path))
    x = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").reshape(-1, w.getnchannels())
    return x.T.copy()


def main():
    refload.install()
    import pacfile
    out = {}
    with tempfile.TemporaryDirectory() as td:
        tmp = Path(td)
        pacfile.output_dir = tmp / "out"
        os.makedirs(pacfile.output_dir, exist_ok=True)
        for name, (fs, pcm) in edge_signals.cases().items():
            src = tmp / f"{name}.wav"
            write_wav(src, fs, pcm)
            for rate in (128, 96):
                try:
                    pacfile.encode_decode_test(src, rate)
                    out[f"{name}_{rate}_pac"] = np.frombuffer((pacfile.output_dir / f"{name}_coded_{rate}.pac").read_bytes(), dtype=np.uint8)
                    out[f"{name}_{rate}_pcm"] = read_wav(pacfile.output_dir / f"{name}_{rate}.wav")
                    print(name, rate, "ok", len(out[f"{name}_{rate}_pac"]))
                except Exception as e:      # noqa: BLE001
                    out[f"{name}_{rate}_error"] = np.array(repr(e))
                    print(name, rate, "reference failed:", repr(e)[:200])
    np.savez_compressed(HERE / "edge.npz", **out)


if __name__ == "__main__":
    main()
