"""oracle/stage_assets.py — TEST / BUILD INFRASTRUCTURE.  Stages the reference's mesh and texture DATA
(no source) under oracle/_ref/assets/ for machines without /root/reference (the GPU box), repairs what is
broken in this checkout of the assets, and writes a pre-decoded `.rgb8` sidecar next to every texture
(decoded with the reference's own stb_image through oracle/_ref/libsrt_ref.so).

    python oracle/stage_assets.py /root/reference oracle/_ref

Repairs (documented in BASELINE.md / DESIGN.md):
  * obj/house/small_house_obj.mtl: material `Material.001` names obj/house/grass-texture-background-min.png,
    which is absent from this checkout (.MISSING_LARGE_BLOBS) -> remapped to ./obj/house/grass.png;
  * obj/bird/12248_Bird_v1_L2.mtl: `map_Kd 12248_Bird_v1_diff.jpg` is a bare file name, but the reference opens
    texture names relative to the process CWD (Object.cpp:55-57) -> ./obj/bird/12248_Bird_v1_diff.jpg.
"""
import ctypes as C
import re
import shutil
import struct
import sys
from pathlib import Path


def main():
    ref, out = Path(sys.argv[1]), Path(sys.argv[2])
    assets = out / "assets"
    assets.mkdir(parents=True, exist_ok=True)
    if (ref / "obj").is_dir():
        shutil.copytree(ref / "obj", assets / "obj", dirs_exist_ok=True, copy_function=shutil.copyfile)
        for n in ("cube.obj", "sphere.obj"):
            shutil.copyfile(ref / n, assets / n)
    fixes = [(assets / "obj/house/small_house_obj.mtl", "./obj/house/grass-texture-background-min.png", "./obj/house/grass.png"),
             (assets / "obj/bird/12248_Bird_v1_L2.mtl", "map_Kd 12248_Bird_v1_diff.jpg", "map_Kd ./obj/bird/12248_Bird_v1_diff.jpg")]
    for path, old, new in fixes:
        if path.exists():
            t = path.read_text(errors="replace")
            if old in t:
                path.write_text(t.replace(old, new))
    lib = C.CDLL(str(out / "libsrt_ref.so"))
    lib.ref_decode_texture.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    n_ok = 0
    for mtl in assets.rglob("*.mtl"):
        for m in re.finditer(r"^\s*map_Kd\s+(\S.*?)\s*$", mtl.read_text(errors="replace"), flags=re.M):
            tex = assets / m.group(1)
            if not tex.exists():
                continue
            w, h = C.c_int(0), C.c_int(0)
            if lib.ref_decode_texture(str(tex).encode(), None, 0, C.byref(w), C.byref(h)) != 0:
                continue
            buf = C.create_string_buffer(w.value * h.value * 3)
            if lib.ref_decode_texture(str(tex).encode(), buf, len(buf), C.byref(w), C.byref(h)) != 0:
                continue
            Path(str(tex) + ".rgb8").write_bytes(b"SRT8" + struct.pack("<ii", w.value, h.value) + buf.raw)
            n_ok += 1
    print(f"staged assets under {assets}; {n_ok} texture sidecars written")


if __name__ == "__main__":
    main()
