"""Read-bandwidth sweep over buffer sizes (L2 -> SM for sizes below the L2 capacity, HBM above it): the measured
denominator of bench.py's `roofline_l2`.  Usage: python tools/l2_bandwidth.py > profiles/rNN_l2_bandwidth.txt"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_raytracer_b200 import capi  # noqa: E402


def main():
    ctx = capi.Context(0)
    info = ctx.info() if hasattr(ctx, "info") else {}
    print(f"# device info: {info}")
    print("# MiB   GB/s (best of 3; all SMs, 4 CTAs of 512 threads per SM, 16-byte ld.global.cg, 4 loads in flight per thread)")
    for mib in (4, 8, 16, 32, 48, 64, 96, 128, 192, 256, 1024, 4096):
        reps = max(3, min(60, 2048 // mib))
        best = max(ctx.measure_read_bandwidth(mib << 20, reps) for _ in range(3))
        print(f"{mib:6d}  {best:9.1f}")
    ctx.destroy()


if __name__ == "__main__":
    main()
