"""Batch driver: a directory of WAV files -> .pac files (and back), sharded over the GPUs of a node.

Replaces the reference's `Pool(8).map(encode_decode_test, ...)` loop (coder/pacfile.py:765-781):

    python -m audio_codec_b200.batch encode IN_DIR OUT_DIR --rate 128 [--decode]
    torchrun --nproc-per-node 8 -m audio_codec_b200.batch encode IN_DIR OUT_DIR --rate 96

One process per GPU (RANK / WORLD_SIZE / LOCAL_RANK from the environment).  Files are dealt to the
ranks longest-first (engine.shard_streams); each rank codes its files in batches that share a
configuration (sample rate, channels) through the host-buffer entry points, writes
<stem>_coded_<rate>.pac (and <stem>_<rate>.wav with --decode), and appends one line per file to
OUT_DIR/manifest.rank<r>.jsonl: {"file", "bytes", "samples", "seconds", "err"}.  Files whose output
already exists are skipped (resume).  No collective is used: the manifests are per-rank files;
`manifest` merges them.
"""
import argparse
import json
import os
import sys
import wave

import numpy as np


def _wav_info(path):
    with wave.open(path) as w:
        return w.getframerate(), w.getnchannels(), w.getnframes(), w.getsampwidth()


def _read_wav(path):
    """Through the drop-in PCMFile reader, so the sample count follows the reference's header rule."""
    from .pacfile import _read_wav as rd
    _, pcm, n_header = rd(path)
    return pcm, n_header


def _write_wav(path, fs, pcm):
    with wave.open(path, "wb") as w:
        w.setnchannels(pcm.shape[0])
        w.setsampwidth(2)
        w.setframerate(fs)
        w.writeframes(np.ascontiguousarray(pcm.T.astype("<i2")).tobytes())


def plan(in_dir, out_dir, rate, world):
    """-> per-rank lists of (path, fs, channels, frames), skipping finished outputs and non-16-bit files."""
    from . import engine
    todo = []
    for name in sorted(os.listdir(in_dir)):
        if not name.lower().endswith(".wav"):
            continue
        path = os.path.join(in_dir, name)
        stem = os.path.splitext(name)[0]
        if os.path.exists(os.path.join(out_dir, f"{stem}_coded_{rate}.pac")):
            continue
        fs, c, n, sw = _wav_info(path)
        if sw != 2 or n == 0:
            print(f"skip {name}: {8 * sw}-bit / empty", file=sys.stderr)
            continue
        todo.append((path, fs, c, n))
    shards = engine.shard_streams([t[3] / t[1] for t in todo], world)
    return [[todo[i] for i in idx] for idx in shards]


def encode_rank(files, out_dir, rate, decode, rank, batch_seconds=4000.0):
    import torch
    from . import _lib, engine
    man = open(os.path.join(out_dir, f"manifest.rank{rank}.jsonl"), "a")
    groups = {}
    for f in files:
        groups.setdefault((f[1], f[2]), []).append(f)
    done = 0
    for (fs, c), fl in groups.items():
        ctx = _lib.Context(sample_rate=fs, n_channels=c, rate_kbps=rate)
        i = 0
        while i < len(fl):
            j, secs = i, 0.0
            while j < len(fl) and (j == i or secs + fl[j][3] / fs <= batch_seconds):
                secs += fl[j][3] / fs
                j += 1
            rd = [_read_wav(f[0]) for f in fl[i:j]]
            pcm, n_hdr = [r[0] for r in rd], [r[1] for r in rd]
            buf, nbs = engine.pack_pcm(pcm, c)
            h_pcm = torch.from_numpy(buf).pin_memory()
            cap = _lib.lib.pac_encode_bound(ctx.h, (_lib.C.c_int * len(nbs))(*[int(v) for v in nbs]), len(nbs))
            h_out = torch.empty(int(cap), dtype=torch.uint8).pin_memory()
            off, err = engine.encode_host(ctx, h_pcm, nbs, n_hdr, h_out)
            out = h_out.numpy()
            if decode:
                po, _ = engine.decode_host(ctx, h_out, off, None)
                h_dec = torch.empty(int(po[-1]), dtype=torch.int16).pin_memory()
                po, derr = engine.decode_host(ctx, h_out, off, h_dec)
                err |= derr
            for k, f in enumerate(fl[i:j]):
                stem = os.path.splitext(os.path.basename(f[0]))[0]
                pac = os.path.join(out_dir, f"{stem}_coded_{rate}.pac")
                with open(pac + ".part", "wb") as fp:
                    fp.write(out[int(off[k]):int(off[k + 1])].tobytes())
                os.replace(pac + ".part", pac)               # a finished name is a finished file (resume)
                if decode:
                    y = h_dec.numpy()[int(po[k]):int(po[k + 1])].reshape(c, -1)
                    _write_wav(os.path.join(out_dir, f"{stem}_{rate}.wav"), fs, y)
                man.write(json.dumps({"file": os.path.basename(f[0]), "bytes": int(off[k + 1] - off[k]),
                                      "samples": int(f[3]), "seconds": f[3] / fs, "err": int(err)}) + "\n")
                done += 1
            man.flush()
            i = j
    man.close()
    return done


def merge_manifest(out_dir):
    rows = []
    for name in sorted(os.listdir(out_dir)):
        if name.startswith("manifest.rank") and name.endswith(".jsonl"):
            rows += [json.loads(ln) for ln in open(os.path.join(out_dir, name)) if ln.strip()]
    rows.sort(key=lambda r: r["file"])
    with open(os.path.join(out_dir, "manifest.json"), "w") as fp:
        json.dump({"files": rows, "seconds": sum(r["seconds"] for r in rows), "bytes": sum(r["bytes"] for r in rows)}, fp, indent=1)
    return rows


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m audio_codec_b200.batch")
    sub = ap.add_subparsers(dest="cmd", required=True)
    e = sub.add_parser("encode")
    e.add_argument("in_dir")
    e.add_argument("out_dir")
    e.add_argument("--rate", type=int, default=128, help="kb/s per channel (SBR below 128, as the reference)")
    e.add_argument("--decode", action="store_true", help="also decode each .pac to <stem>_<rate>.wav")
    m = sub.add_parser("manifest")
    m.add_argument("out_dir")
    a = ap.parse_args(argv)
    if a.cmd == "manifest":
        rows = merge_manifest(a.out_dir)
        print(f"{len(rows)} files")
        return 0
    import torch
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    os.makedirs(a.out_dir, exist_ok=True)
    files = plan(a.in_dir, a.out_dir, a.rate, world)[rank]
    n = encode_rank(files, a.out_dir, a.rate, a.decode, rank)
    print(f"rank {rank}/{world}: {n} files")
    return 0


if __name__ == "__main__":
    sys.exit(main())
