"""tools/stage_assets.py — BUILD INFRASTRUCTURE.  Copies the reference's mesh and texture DATA (no source) into
assets/ at the repository root for machines without /root/reference (the GPU box), and repairs the two texture
references that are broken in this checkout of the assets.  assets/ is git-ignored and travels with gpurun snapshots.

    python tools/stage_assets.py /root/reference assets

Textures are decoded by the product's own PNG / JPEG decoders (csrc/host/srt_png.cpp, srt_jpeg.cpp), bit-identical to the
reference's stb_image (tests/test_host_scene.py checks that against the live reference).  Nothing here, and nothing the
product reads, lives under oracle/.
"""
import re
import shutil
import sys
from pathlib import Path


def main():
    ref, assets = Path(sys.argv[1]), Path(sys.argv[2])
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
    print(f"staged assets under {assets}")


if __name__ == "__main__":
    main()
