"""Build a variant of libvslam_b200.so with extra nvcc flags on chosen sources, for A/B runs of one kernel.

usage: python tools/build_variant.py NAME "-DVS_LEVEL_CTAS=5 ..." [source.cu ...]   (default source: orb_level.cu)
Writes visual-slam_b200/build/libvslam_NAME.so; run with VSLAM_LIB_PATH=visual-slam_b200/build/libvslam_NAME.so.
The other objects are the ones the normal build left in visual-slam_b200/build/ (run build.py first).
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "visual-slam_b200")
sys.path.insert(0, PKG)
import build as B  # noqa: E402


def main():
    name, extra = sys.argv[1], sys.argv[2].split()
    which = sys.argv[3:] or ["orb_level.cu"]
    B.build()
    objs = []
    for src in B.sources():
        base = os.path.basename(src)
        obj = os.path.join(B.BUILD, base[:-3] + ".o")
        if base in which:
            obj = os.path.join(B.BUILD, base[:-3] + "_" + name + ".o")
            subprocess.check_call([B.NVCC] + B.FLAGS + extra + ["-c", src, "-o", obj])
        objs.append(obj)
    out = os.path.join(B.BUILD, "libvslam_%s.so" % name)
    subprocess.check_call([B.NVCC, "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    print(out)


if __name__ == "__main__":
    main()
