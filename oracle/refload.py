"""Import-time loader for the UNMODIFIED reference under /root/reference/coder.

TEST INFRASTRUCTURE ONLY.  Used by tests/golden/make_golden.py (and by opt-in
whole-file tests) in the build container, where /root/reference exists.  It
never runs on the GPU box and nothing in the product package imports it.

The reference targets NumPy < 1.24; no source file is copied or edited on
disk.  The loader patches the text in memory at import:
  * np.int / np.float aliases (removed in NumPy 1.24)
  * ndarray.tostring() (removed in NumPy 2)
  * bitpack.ReadBits: uint8 element no longer promotes to Python int (NEP 50)
  * matplotlib is stubbed (not installed; plotting only)
Reference module names (mdct, window, codec ...) are top-level, so call
`install()` only in a process that does not import this repo's shims under
the same bare names (the repo's shims live in the audio_codec_b200 package).
"""
import importlib.abc
import importlib.util
import os
import re
import sys
import types

REF_ROOT = os.environ.get("PAC_REFERENCE_ROOT", "/root/reference")
REF = os.path.join(REF_ROOT, "coder")


def available():
    return os.path.isdir(REF)


class _Loader(importlib.abc.SourceLoader):
    def __init__(self, path):
        self.path = path

    def get_filename(self, name):
        return self.path

    def get_data(self, path):
        src = open(path, "rb").read().decode("utf-8-sig")
        src = re.sub(r"\bnp\.int\b(?!\d)", "int", src)
        src = re.sub(r"\bnp\.float\b(?!\d)", "float", src)
        src = src.replace(".tostring()", ".tobytes()")
        if path.endswith("bitpack.py"):
            src = src.replace("&= self.data[self.iByte]",
                              "&= int(self.data[self.iByte])")
            src = src.replace("dataMask = self.data[self.iByte]",
                              "dataMask = int(self.data[self.iByte])")
        return src.encode()


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
        p = os.path.join(REF, name + ".py")
        if "." not in name and os.path.exists(p):
            return importlib.util.spec_from_loader(name, _Loader(p), origin=p)
        return None


_installed = False


def install(scratch=None):
    """Make `import mdct, codec, pacfile ...` resolve to the reference."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference tree not present at " + REF)
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.lines"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].lines = sys.modules["matplotlib.lines"]
    sys.meta_path.insert(0, _Finder())
    sys.dont_write_bytecode = True
    sys.setrecursionlimit(12000)
    # pacfile.py creates ../test_debug_long_out relative to cwd at import
    scratch = scratch or "/tmp/pac_ref_scratch/work"
    os.makedirs(scratch, exist_ok=True)
    os.chdir(scratch)
    _installed = True
