"""CPU oracle for the perceptual-coder hot path.  TEST INFRASTRUCTURE ONLY.

A NumPy/SciPy restatement of the reference algorithm (Abhipray/audio-codec,
`coder/*.py`), each function citing the reference file:line it follows.  It
exists to CHECK the CUDA path; it is never the product path.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference`
legs may import it.  Nothing under `audio_codec_b200/` imports it.

Parity status: PINNED.  `tests/test_oracle_golden.py` checks this restatement
against golden vectors produced by running the unmodified reference in the
build container (`tests/golden/make_golden.py`): per-function known answers,
`.pac` byte identity and decoded-PCM identity on excerpts of the four
test_signals at 96 and 128 kb/s (VQ, VQ+SBR) and the block-floating-point
variant, plus (opt-in, needs /root/reference) the reference's own shipped
fixtures `test_decoded_full/*`.

Third-party arithmetic the reference calls and this oracle calls identically:
numpy.fft (pocketfft), numpy ufuncs/reductions, scipy.ndimage.gaussian_filter1d,
scipy.interpolate.interp1d (NumPy 2.3.5 / SciPy 1.18.1 reproduce the shipped
fixtures byte for byte; the reference pins no versions).
"""
