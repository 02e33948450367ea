"""Build libpac_b200.so in-tree with nvcc for sm_100a (B200).

    python -m audio_codec_b200.build [--force]
"""
import os
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / os.environ.get("PAC_LIB_NAME", "libpac_b200.so")   # experiments build variants beside the product library
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--use_fast_math=false"]
FLAGS = [f for f in FLAGS if f != "--use_fast_math=false"]  # never fast-math: fp64 parity path
FLAGS += os.environ.get("PAC_CFLAGS", "").split()          # e.g. -DPAC_PHASE_TIMING (debug builds only)


def sources():
    return sorted(CSRC.glob("*.cu"))


def stale():
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + \
        list((PKG.parent / "include").glob("*.h"))
    return any(p.stat().st_mtime > t for p in deps)


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    objs = []
    obj_dir = PKG / "build" / LIB.stem if LIB.name != "libpac_b200.so" else PKG / "build"
    obj_dir.mkdir(parents=True, exist_ok=True)
    procs = []
    for src in sources():
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    fail = False
    for src, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src.name}\n{out}\n")
        fail |= p.returncode != 0
    if fail:
        raise RuntimeError("nvcc failed")
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *map(str, objs)]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
