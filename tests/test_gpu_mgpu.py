"""srt_mgpu_* (include/srt.h): several GPUs from one host thread through the C ABI.

The sharding logic is exact (pixels :511-523 and frames :534 are independent), so every result here is compared byte for
byte with a one-context render.  On a one-GPU box the device list repeats ordinal 0 (several shards on one GPU); where the
box has more GPUs the same tests run over all of them, and the shards of GPUs 1.. are written straight into GPU 0's frame
through peer-mapped memory.
"""
import ctypes as C

import numpy as np
import pytest

from simple_raytracer_b200 import _abi, capi, host, scenes

pytestmark = pytest.mark.gpu


def _scene(key, angle):
    cfg = scenes.CONFIGS[key]
    hs = host.HostScene()
    view = cfg.build(hs, angle)
    light = scenes.light_view_space(hs, view)
    arrays = hs.flatten()
    hs.close()
    return cfg, arrays, np.asarray(light, np.float32)


def _device_lists():
    n = capi.device_count()
    lists = [[0, 0, 0], [0]]
    if n >= 2:
        lists.append(list(range(n)))
        lists.append([1, 0])                 # the frame lives on GPU 1, GPU 0 writes into it
    return lists


@pytest.mark.parametrize("key,lights", [("C3", 8), ("C4", 2)])
def test_mgpu_render_frame_equals_one_context(gpu_ctx, key, lights):
    cfg, arrays, light = _scene(key, 200.0 if key == "C3" else 30.0)
    W, H = 482, 270                          # not a multiple of the tile width
    p = _abi.default_params(W, H, float(H), light, lights, True)
    sc = gpu_ctx.upload(arrays)
    want = gpu_ctx.render(sc, p)
    sc.destroy()
    for devices in _device_lists():
        m = capi.MultiGpu(devices)
        info = m.info()
        assert info["n_devices"] == len(devices) and 1 <= info["peer_writes"] <= len(devices)
        m.upload(arrays)
        for block in (0, 4, 32):
            got = m.render_frame(p, row_block=block)
            for k in ("rgb", "obj_id", "tri_id", "t"):
                assert np.array_equal(got[k], want[k]), (devices, block, k)
            assert got["stats"]["hit_pixels"] == want["stats"]["hit_pixels"]
            assert got["stats"]["primary_rays"] == W * H
        # a row range, rgb only, into a DEVICE frame on the first GPU (no download by the library)
        first = capi.Context(devices[0])
        dev_rgb = capi.lib().srt_device_alloc(first.h, W * H * 3)
        p2 = _abi.default_params(W, H, float(H), light, lights, True)
        p2.row_begin, p2.row_end = 12, 201
        m.render_frame_raw(p2, rgb=dev_rgb, row_block=8)
        m.sync()
        # read it back through a one-piece render's staging: compact it with the no-alloc entry point's sibling instead
        pts = capi.PinnedArray((W * H * 2,), np.float32); col = capi.PinnedArray((W * H * 3,), np.float32); cnt = capi.PinnedArray((1,), np.uint64)
        capi._check(capi.lib().srt_sparse_from_frame_static(first.h, dev_rgb, W, H, 12, 201, pts.ptr, col.ptr, cnt.ptr))
        first.sync()
        n = int(cnt.array[0])
        assert n == (201 - 12) * W                          # sky-filled frame: every pixel of the range is drawn
        xs = pts.array[:2 * n:2].astype(np.int64); ys = pts.array[1:2 * n:2].astype(np.int64)
        back = np.zeros((H, W, 3), np.uint8)
        back[ys, xs] = col.array[:3 * n].reshape(n, 3).astype(np.uint8)
        assert np.array_equal(back[12:201], want["rgb"][12:201]), devices
        capi.lib().srt_device_free(first.h, dev_rgb)
        pts.free(); col.free(); cnt.free()
        first.destroy()
        # the scene replaced in place (next frame of an animation), same handle
        _cfg, arrays2, light2 = _scene(key, 75.0)
        m.upload(arrays2)
        p3 = _abi.default_params(W, H, float(H), light2, lights, True)
        got = m.render_frame(p3, ids=False, want_t=False)
        sc2 = gpu_ctx.upload(arrays2)
        assert np.array_equal(got["rgb"], gpu_ctx.render(sc2, p3, ids=False, want_t=False)["rgb"]), devices
        sc2.destroy()
        m.destroy()


def test_mgpu_render_series_equals_frame_by_frame(gpu_ctx):
    cfg = scenes.CONFIGS["C3"]
    W, H, focal = scenes.scaled(cfg, 135)
    angles = [0.0, 36.0, 91.5, 135.0, 180.0, 200.0, 270.0, 300.0, 358.5]
    frames = [_scene("C3", a) for a in angles]
    params = [_abi.default_params(W, H, focal, light, 4, True) for (_c, _a, light) in frames]
    want = []
    for (_c, arrays, _l), p in zip(frames, params):
        sc = gpu_ctx.upload(arrays)
        want.append(gpu_ctx.render(sc, p, ids=False, want_t=False)["rgb"])
        sc.destroy()
    for devices in _device_lists():
        m = capi.MultiGpu(devices)
        outs = [capi.PinnedArray((H, W, 3), np.uint8) for _ in angles]
        m.render_series([a for (_c, a, _l) in frames], params, [o.ptr for o in outs])
        m.sync()
        for i, o in enumerate(outs):
            assert np.array_equal(o.array, want[i]), (devices, i)
        m.render_series([a for (_c, a, _l) in frames][::-1], params[::-1], [o.ptr for o in outs])     # lanes reused, scenes updated in place
        m.sync()
        for i, o in enumerate(outs):
            assert np.array_equal(o.array, want[len(angles) - 1 - i]), (devices, "second pass", i)
        for o in outs:
            o.free()
        m.destroy()


def test_mgpu_argument_errors():
    with pytest.raises(capi.SrtError):
        capi.MultiGpu([])
    with pytest.raises(capi.SrtError):
        capi.MultiGpu([capi.device_count() + 5])
    m = capi.MultiGpu([0])
    p = _abi.default_params(64, 64, 64.0, np.zeros(3, np.float32), 1, True)
    with pytest.raises(capi.SrtError):          # no scene yet
        m.render_frame(p)
    m.destroy()


_CHILD = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from simple_raytracer_b200 import _abi, capi, host, scenes
handle = bytes.fromhex(sys.argv[2]); W, H, shard, world = (int(x) for x in sys.argv[3:7])
cfg = scenes.CONFIGS["C3"]
hs = host.HostScene(); view = cfg.build(hs, 200.0); light = np.asarray(scenes.light_view_space(hs, view), np.float32); arrays = hs.flatten(); hs.close()
ctx = capi.Context(0)
ptr = ctx.ipc_open(handle)                       # the parent's frame, mapped into this process
sc = ctx.upload(arrays)
p = _abi.default_params(W, H, float(H), light, 4, True)
p.row_block, p.shard_index, p.shard_count = 16, shard, world
ctx.render_raw(sc, p, rgb=ptr)                   # this process's kernels store straight into the parent's memory
ctx.sync()
ctx.ipc_close(ptr); sc.destroy(); ctx.destroy()
print("CHILD OK")
"""


def test_ipc_shared_frame_between_two_processes(gpu_ctx, tmp_path):
    """srt_ipc_export / srt_ipc_open / srt_ipc_close: another PROCESS maps this process's frame and renders its shard of the
    rows straight into it (the one-process-per-GPU form of the peer stores; here both processes share GPU 0).  The union of
    the parent's shard and the child's equals the one-piece frame byte for byte."""
    import subprocess
    import sys
    from pathlib import Path
    torch = pytest.importorskip("torch")
    from simple_raytracer_b200 import multigpu
    repo = Path(__file__).resolve().parent.parent
    cfg, arrays, light = _scene("C3", 200.0)
    W, H = 480, 272
    p = _abi.default_params(W, H, float(H), light, 4, True)
    sc = gpu_ctx.upload(arrays)
    want = gpu_ctx.render(sc, p, ids=False, want_t=False)["rgb"]
    ptr = gpu_ctx.device_alloc(W * H * 3)
    try:
        frame = torch.as_tensor(multigpu._DeviceBuffer(ptr, W * H * 3), device=torch.device("cuda", 0))
        frame.zero_()
        torch.cuda.synchronize()
        handle = gpu_ctx.ipc_export(ptr)
        assert len(handle) == 64
        script = tmp_path / "child.py"
        script.write_text(_CHILD)
        r = subprocess.run([sys.executable, str(script), str(repo), handle.hex(), str(W), str(H), "1", "2"], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "CHILD OK" in r.stdout, r.stderr[-2000:]
        p.row_block, p.shard_index, p.shard_count = 16, 0, 2
        gpu_ctx.render_raw(sc, p, rgb=ptr)
        gpu_ctx.sync()
        got = frame.view(H, W, 3).cpu().numpy()
        assert np.array_equal(got, want)
    finally:
        gpu_ctx.device_free(ptr)
        sc.destroy()
