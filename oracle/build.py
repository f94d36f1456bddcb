"""Build recipe for the parity oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

    python oracle/build.py            # builds what is stale
    python oracle/build.py --force

Outputs (both git-ignored, both travel to the GPU box with the gpurun snapshot):
  oracle/liboracle.so       gcc build of oracle/detection_oracle.c (the restatement)
  oracle/_ref/libtvref.so   the reference's own cpu/roi_align_kernel.cpp,
                            cpu/nms_kernel.cpp, cpu/roi_pool_kernel.cpp and
                            cpu/ps_roi_pool_kernel.cpp, cpu/ps_roi_align_kernel.cpp, compiled from /root/reference where
                            they lie via oracle/ref_wrap.cpp.  Only built when
                            /root/reference exists (it does not on the GPU box; the
                            prebuilt file is used there).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("B200DET_REFERENCE", "/root/reference")


def _stale(out, srcs):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in srcs)


def build_oracle(force=False):
    src = os.path.join(HERE, "detection_oracle.c")
    out = os.path.join(HERE, "liboracle.so")
    if force or _stale(out, [src]):
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fvisibility=hidden", "-shared",
               "-fPIC", "-Wall", "-o", out, src, "-lm"]
        subprocess.check_call(cmd)
    return out


def build_ref(force=False):
    """g++ directly on the reference's files (through ref_wrap.cpp's #includes)."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libtvref.so")
    ref_srcs = [os.path.join(REFERENCE, "torchvision/csrc/ops/cpu", f)
                for f in ("roi_align_kernel.cpp", "roi_align_common.h", "nms_kernel.cpp", "roi_pool_kernel.cpp",
                          "ps_roi_pool_kernel.cpp", "ps_roi_align_kernel.cpp")]
    if not all(os.path.exists(s) for s in ref_srcs):
        return out if os.path.exists(out) else None
    wrap = os.path.join(HERE, "ref_wrap.cpp")
    wrap_pool = os.path.join(HERE, "ref_wrap_roi_pool.cpp")
    wrap_ps = os.path.join(HERE, "ref_wrap_ps_roi_pool.cpp")
    wrap_psa = os.path.join(HERE, "ref_wrap_ps_roi_align.cpp")
    if not (force or _stale(out, [wrap, wrap_pool, wrap_ps, wrap_psa] + ref_srcs)):
        return out
    import torch
    from torch.utils import cpp_extension as ce

    os.makedirs(out_dir, exist_ok=True)
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O3", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI),
           "-I" + REFERENCE]
    cmd += ["-I" + p for p in ce.include_paths()]
    cmd += [wrap, wrap_pool, wrap_ps, wrap_psa, "-o", out, "-L" + torch_lib, "-ltorch", "-ltorch_cpu", "-lc10",
            "-Wl,-rpath," + torch_lib, "-pthread"]
    subprocess.check_call(cmd)
    return out


def main():
    force = "--force" in sys.argv
    print("oracle:", build_oracle(force))
    print("reference:", build_ref(force))


if __name__ == "__main__":
    main()
