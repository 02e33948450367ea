"""GPU: the no-allocation entry point pac_gain_shape_dequantize_static (
no counterpart in the reference) against the existing pac_gain_shape_dequantize, through the C ABI.

Collected last on purpose (test_zz_*).  Passes alone in a fresh process and after
the other tests (it builds everything it needs; the static scratch keeps the previous call's contents, which the
entry point must not depend on - two calls in a row with different shapes check that)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-12, 1e-15          # the tolerance of tests/test_ops_gpu.py::test_gain_shape_identical_to_oracle


def _streams(L, bits, n, seed):
    """n random bands of L lines quantised at `bits` bits per line -> (byte buffer [n, row], start bit per vector
    (absolute, rows back to back), budgets [n])."""
    import torch
    from audio_codec_b200 import bitpack, gain_shape_quantize as gs
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, L)) * 10.0 ** rng.uniform(-3, 0, (n, 1))
    x[1, L // 2:] *= 1e-3
    x[2, 1:] = 0.0                                      # a single pulse
    ba = np.full(n, bits * L, dtype=np.int32)
    idx, wid, nf = gs.quantize_gain_shape_batch(x, ba)
    row = (int(wid.sum(dim=1).max()) + 7) // 8 + 8
    rec = bitpack.pack_rows(idx, wid, nf, row)
    pos = torch.arange(n, dtype=torch.int64, device=rec.device) * (rec.shape[1] * 8)
    return rec.reshape(-1).contiguous(), pos, torch.from_numpy(ba).to(rec.device), wid.sum(dim=1).to(torch.int64)


def _call(fn, ctx, rec, pos, ba, L):
    import torch
    from audio_codec_b200 import _lib
    n = pos.numel()
    out = torch.full((n, L), float("nan"), dtype=torch.float64, device=rec.device)
    pos_out = torch.zeros(n, dtype=torch.int64, device=rec.device)
    err = C.c_int(0)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = fn(ctx.h, p(rec), rec.numel(), p(pos), p(ba), L, n, p(out), p(pos_out), C.byref(err), _lib.stream_ptr())
    return rc, err.value, out.cpu().numpy(), pos_out.cpu().numpy()


def test_no_alloc_dequantize_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    from audio_codec_b200 import _lib
    ctx = _lib.default_context(48000)
    lib = _lib.lib
    # two calls in a row with different band lengths and budgets: the second finds the first one's workspaces and
    # error word in the shared static scratch and must not be affected by them
    for L, bits, n, seed in ((149, 5, 96, 1), (13, 9, 200, 2), (363, 4, 64, 3)):
        rec, pos, ba, used = _streams(L, bits, n, seed)
        assert n <= lib.pac_gain_shape_dequantize_static_max_vectors(L)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc1, e1, y1, p1 = _call(lib.pac_gain_shape_dequantize_static, ctx, rec, pos, ba, L)
        rc0, e0, y0, p0 = _call(lib.pac_gain_shape_dequantize, ctx, rec, pos, ba, L)
        rc2, e2, y2, p2 = _call(lib.pac_gain_shape_dequantize_static, ctx, rec, pos, ba, L)
        assert (rc0, e0) == (0, 0) and (rc1, e1) == (0, 0) and (rc2, e2) == (0, 0)
        assert np.isfinite(y0).all() and np.abs(y0).max() > 0
        for y, p in ((y1, p1), (y2, p2)):
            np.testing.assert_allclose(y, y0, rtol=RTOL, atol=ATOL)
            assert np.array_equal(y, y0)                 # same operations on the same values: in fact identical
            assert np.array_equal(p, p0)
        assert np.array_equal(p0 - pos.cpu().numpy(), used.cpu().numpy())
        assert torch.cuda.memory_allocated() >= before   # (torch's allocator is not involved in the call at all)
    # a soft error raised by one call (truncated buffer) is reported by it and does not leak into the next
    rec, pos, ba, _ = _streams(41, 8, 8, 4)
    rc, e, _, _ = _call(lib.pac_gain_shape_dequantize_static, ctx, rec[:rec.numel() // 2].contiguous(), pos, ba, 41)
    rc0, e0, _, _ = _call(lib.pac_gain_shape_dequantize, ctx, rec[:rec.numel() // 2].contiguous(), pos, ba, 41)
    assert (rc, e) == (rc0, e0) and rc == -4 and e != 0
    rc, e, y, _ = _call(lib.pac_gain_shape_dequantize_static, ctx, rec, pos, ba, 41)
    rc0, e0, y0, _ = _call(lib.pac_gain_shape_dequantize, ctx, rec, pos, ba, 41)
    assert (rc, e) == (0, 0) and np.array_equal(y, y0)


def test_no_alloc_entry_point_rejects_requests_beyond_its_scratch():
    import torch
    from audio_codec_b200 import _lib
    ctx = _lib.default_context(48000)
    lib = _lib.lib
    assert lib.pac_gain_shape_dequantize_static_max_vectors(363) == (4 << 20) // (3 * 512 * 8) == 341
    assert lib.pac_gain_shape_dequantize_static_max_vectors(1) == (4 << 20) // 24
    assert lib.pac_gain_shape_dequantize_static_max_vectors(0) == 0
    L, n = 363, 342
    rec = torch.zeros(64, dtype=torch.uint8, device="cuda")
    pos = torch.zeros(n, dtype=torch.int64, device="cuda")
    ba = torch.zeros(n, dtype=torch.int32, device="cuda")
    rc, _, y, _ = _call(lib.pac_gain_shape_dequantize_static, ctx, rec, pos, ba, L)
    assert rc == -5 and np.isnan(y).all()                # PAC_ERR_TOO_LARGE, nothing written
    assert b"static scratch" in lib.pac_strerror(-5)
    rc, e, y, _ = _call(lib.pac_gain_shape_dequantize_static, ctx, rec, pos[:341].contiguous(), ba[:341].contiguous(), L)
    rc0, e0, y0, _ = _call(lib.pac_gain_shape_dequantize, ctx, rec, pos[:341].contiguous(), ba[:341].contiguous(), L)
    assert (rc, e) == (rc0, e0) == (0, 0) and np.array_equal(y, y0)      # the largest request that fits
