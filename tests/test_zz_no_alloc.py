"""srt_sparse_from_frame_static: the ImageData compaction for callers that must not allocate (include/srt.h).

No counterpart in the reference.  The entry point builds the reference's
ImageData lists (:28, :505-525) from a dense frame with its scratch in a static device array — no cudaMalloc, no staging.
Checked here against the existing entry point (srt_render_sparse + srt_fetch_sparse) on two calls in a row; the second
frame is smaller than the first, so whatever the first call left in the shared scratch must not leak into it.

Collected last on purpose: a failure here must not hide other tests.  Self-contained: builds its
own context and scene, so it passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from simple_raytracer_b200 import _abi, capi, host, scenes

pytestmark = [pytest.mark.gpu]


def _scene():
    cfg = scenes.CONFIGS["C3"]
    hs = host.HostScene()
    view = cfg.build(hs, 20.0)
    light = scenes.light_view_space(hs, view)
    arrays = hs.flatten()
    hs.close()
    return arrays, np.asarray(light, np.float32)


def _static_sparse(ctx, frame, rows, out_pts, out_col, out_n):
    H, W = frame.shape[:2]
    r0, r1 = rows if rows else (0, 0)
    capi._check(capi.lib().srt_sparse_from_frame_static(ctx.h, frame.ptr, W, H, r0, r1, out_pts.ptr, out_col.ptr, out_n.ptr))
    ctx.sync()
    n = int(out_n.array[0])
    return out_pts.array.reshape(-1, 2)[:n].copy(), out_col.array.reshape(-1, 3)[:n].copy()


def test_static_scratch_sparse_equals_render_sparse_twice_in_a_row():
    if capi.device_count() == 0:
        pytest.fail("gpu test selected but no CUDA device is visible — the CUDA path has no fallback")
    ctx = capi.Context(0)
    arrays, light = _scene()
    sc = ctx.upload(arrays)
    cap = 640 * 360
    # caller-owned buffers, allocated BEFORE the calls under test: pinned host memory the device reaches directly
    out_pts = capi.PinnedArray((cap * 2,), np.float32)
    out_col = capi.PinnedArray((cap * 3,), np.float32)
    out_n = capi.PinnedArray((1,), np.uint64)
    try:
        for W, H, rows in ((640, 360, None), (333, 201, None), (333, 201, (37, 160))):
            p = _abi.default_params(W, H, float(H), light, 4, True)
            if rows:
                p.row_begin, p.row_end = rows
            want_pts, want_col = ctx.render_sparse(sc, p)                 # the existing entry point
            q = _abi.default_params(W, H, float(H), light, 4, True)       # the dense frame a caller would hold: (0,0,0) = not drawn
            q.no_sky_fill = 1
            frame = capi.PinnedArray((H, W, 3), np.uint8)
            ctx.render_raw(sc, q, rgb=frame.ptr)
            ctx.sync()
            out_pts.array[:] = -7.0
            out_col.array[:] = -7.0
            out_n.array[0] = 12345
            got_pts, got_col = _static_sparse(ctx, frame, rows, out_pts, out_col, out_n)
            assert len(want_pts) > 1000
            assert len(got_pts) == len(want_pts)
            assert np.array_equal(got_pts, want_pts) and np.array_equal(got_col, want_col)      # integers as floats: exact
            assert np.all(out_pts.array[2 * len(got_pts):] == -7.0), "wrote past the last entry"
            frame.free()
        # limits and argument checks: a clear error, nothing launched
        big = capi.PinnedArray((8, 3), np.uint8)
        L = capi.lib()
        rc = L.srt_sparse_from_frame_static(ctx.h, big.ptr, 1 << 20, 4096, 0, 0, out_pts.ptr, out_col.ptr, out_n.ptr)
        assert rc == _abi.SRT_ERR_UNSUPPORTED and b"SRT_SPARSE_STATIC_MAX_SEGMENTS" in L.srt_last_error()
        pageable = np.zeros((8, 8, 3), np.uint8)
        rc = L.srt_sparse_from_frame_static(ctx.h, pageable.ctypes.data, 8, 8, 0, 0, out_pts.ptr, out_col.ptr, out_n.ptr)
        assert rc == _abi.SRT_ERR_INVALID_ARG and b"pageable" in L.srt_last_error()
        big.free()
    finally:
        out_pts.free(); out_col.free(); out_n.free()
        sc.destroy()
        ctx.destroy()
