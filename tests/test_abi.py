"""CPU-side checks of the C ABI surface: the library loads, exports every symbol include/srt.h
declares, structure sizes agree, and compute entry points fail loudly without a device."""
import ctypes as C
import re
from pathlib import Path

import pytest

from simple_raytracer_b200 import _abi, capi

REPO = Path(__file__).resolve().parent.parent


def declared_functions():
    text = (REPO / "include" / "srt.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = capi.lib()
    names = declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/srt.h but not exported"
    assert sorted(capi.EXPORTS) == names
    assert lib.srt_abi_version() == 1


def test_struct_sizes_match_header(tmp_path):
    """sizeof() as the C compiler sees include/srt.h == the ctypes / numpy mirrors."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "srt.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(srt_render_params),sizeof(srt_frame),sizeof(srt_scene_desc),sizeof(srt_stats),'
                   'sizeof(srt_node),sizeof(srt_tri_geom),sizeof(srt_tri_attr),sizeof(srt_object),'
                   'sizeof(srt_texture_desc));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", str(REPO / "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert sizes == [C.sizeof(_abi.RenderParams), C.sizeof(_abi.Frame), C.sizeof(_abi.SceneDesc),
                     C.sizeof(_abi.Stats), _abi.NODE_DTYPE.itemsize, _abi.GEOM_DTYPE.itemsize,
                     _abi.ATTR_DTYPE.itemsize, _abi.OBJECT_DTYPE.itemsize, C.sizeof(_abi.TextureDesc)]
    p = _abi.RenderParams()
    capi.lib().srt_default_params(C.byref(p))
    assert (p.width, p.height, p.focal, p.light_count) == (600, 400, 400.0, 1)
    assert (p.light_step, p.shadow_div, p.tone_k) == (3.0, 5.0, 0.5)
    assert abs(p.gamma - 1.1) < 1e-6 and list(p.sky) == [173, 216, 230] and p.use_bvh == 1


def test_light_staircase():
    p = _abi.default_params(8, 8, 8.0, (1.0, 2.0, 3.0), 7)
    L = capi.light_positions(p)
    assert L.tolist() == [[1, 2, 3], [4, 2, 3], [4, 5, 3], [4, 5, 6], [7, 5, 6], [7, 8, 6], [7, 8, 9]]


def test_no_device_fails_loudly():
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.SrtError) as e:
        capi.Context(0)
    assert e.value.code == _abi.SRT_ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_light_sample_positions_staircase_and_jittered():
    """srt_light_positions (host arithmetic of the product library, no GPU needed) against the oracle: the reference's
    staircase (:363-383) and the counter-based jittered samples of light_mode (reproducible, inside the staircase's box)."""
    import numpy as np
    from oracle import pyoracle
    from simple_raytracer_b200 import _abi, capi
    olib = pyoracle.oracle_lib()
    p = _abi.default_params(64, 64, 400.0, (500.0, -300.0, -200.0), 16, True)
    for mode, seed in ((_abi.LIGHTS_STAIRCASE, 0), (_abi.LIGHTS_JITTERED, 0), (_abi.LIGHTS_JITTERED, 12345)):
        p.light_mode, p.light_seed = mode, seed
        got = capi.light_positions(p)
        want = np.zeros((16, 3), np.float32)
        olib.srt_oracle_light_positions_mode(p, want.ctypes.data_as(pyoracle._f32p))
        assert np.array_equal(got, want), (mode, seed)
        if mode == _abi.LIGHTS_JITTERED:
            lo = np.array([500.0, -300.0, -200.0], np.float32)
            assert (got >= lo).all() and (got < lo + 3.0 * 6).all()
            assert len({tuple(r) for r in got}) == 16
            assert np.array_equal(got, capi.light_positions(p))          # reproducible
    p.light_mode, p.light_seed = _abi.LIGHTS_JITTERED, 1
    a = capi.light_positions(p)
    p.light_seed = 2
    assert not np.array_equal(a, capi.light_positions(p))
    p.light_mode = 7
    try:
        capi.light_positions(p)
        assert False, "unknown light_mode accepted"
    except capi.SrtError as e:
        assert e.code == _abi.SRT_ERR_INVALID_ARG
