"""Build recipe for baseline/_ref -- the UNMODIFIED reference package, TEST / BENCH INFRASTRUCTURE only.

    python baseline/build_ref.py [--force] [--cuda]

Produces (all git-ignored; they travel to the GPU box with the gpurun snapshot):

  baseline/_ref/torchvision/         the reference's Python package, copied as it is (no csrc/), plus
  baseline/_ref/torchvision/_C.so    its CPU-only extension built from the sources where they lie under
                                     /root/reference (the list setup.py:155-162 uses: csrc/*.cpp,
                                     csrc/ops/*.cpp, csrc/ops/{autocast,autograd,cpu,quantized/cpu}/*.cpp;
                                     one g++ call per file).  This is the configuration INTEGRATION.md
                                     describes: schemas, Autograd and Autocast kernels come from the
                                     reference, `cpu_vision_b200` adds only the CUDA kernels.
  baseline/_ref/test/                the reference's own test files for the path (test_ops.py, test_models.py, their
                                     helpers and the pickled expectations), copied as they are: tests/test_dropin_gpu.py
                                     runs them against the new CUDA kernels
  baseline/_ref/tvref_cuda.so        (--cuda) the reference's own cuda/roi_align_kernel.cu and
                                     cuda/nms_kernel.cu compiled unmodified for sm_100a: the stock GPU
                                     kernels as a reported comparison.  Loaded only by
                                     baseline/stock_cuda_bench.py in a process of its own, never together
                                     with the product library (both register torchvision:: CUDA kernels).

Nothing here is imported by cpu_vision_b200/.  /root/reference does not exist on the GPU box: the
prebuilt files are used there, and every consumer skips when they are absent.
"""
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200DET_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

VERSION_PY = ("__version__ = '0.20.0a0+ref'\ngit_version = 'reference'\n"
              "from torchvision.extension import _check_cuda_version\n"
              "if _check_cuda_version() > 0:\n    cuda = _check_cuda_version()\n")


def _flags():
    import torch
    from torch.utils import cpp_extension as ce

    csrc = os.path.join(REF, "torchvision", "csrc")
    inc = ["-I" + p for p in ce.include_paths()] + ["-I" + csrc, "-I" + sysconfig.get_paths()["include"]]
    abi = "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    return csrc, inc, abi, torch_lib


def available():
    return os.path.exists(os.path.join(OUT, "torchvision", "_C.so"))


def build_package(force=False):
    pkg = os.path.join(OUT, "torchvision")
    so = os.path.join(pkg, "_C.so")
    if os.path.exists(so) and not force:
        return OUT
    if not os.path.isdir(os.path.join(REF, "torchvision", "csrc")):
        return OUT if os.path.exists(so) else None
    csrc, inc, abi, torch_lib = _flags()
    obj_dir = os.path.join(OUT, "obj")
    if os.path.exists(pkg):
        _rmtree(pkg)
    os.makedirs(obj_dir, exist_ok=True)
    shutil.copytree(os.path.join(REF, "torchvision"), pkg,
                    ignore=shutil.ignore_patterns("csrc", "*.so", "__pycache__"))
    with open(os.path.join(pkg, "version.py"), "w") as f:
        f.write(VERSION_PY)
    srcs = glob.glob(os.path.join(csrc, "*.cpp")) + glob.glob(os.path.join(csrc, "ops", "*.cpp"))
    for sub in ("autocast", "autograd", "cpu", "quantized/cpu"):
        srcs += glob.glob(os.path.join(csrc, "ops", sub, "*.cpp"))

    def cc(src):
        obj = os.path.join(obj_dir, os.path.relpath(src, csrc).replace("/", "__") + ".o")
        subprocess.check_call(["g++", "-O3", "-std=c++17", "-fPIC", abi, "-DTORCH_EXTENSION_NAME=_C"] + inc +
                              ["-c", src, "-o", obj])
        return obj

    with cf.ThreadPoolExecutor(8) as ex:
        objs = list(ex.map(cc, srcs))
    subprocess.check_call(["g++", "-shared", "-o", so] + objs +
                          ["-L" + torch_lib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
                           "-Wl,-rpath," + torch_lib])
    shutil.rmtree(obj_dir)
    _writable(pkg)
    return OUT


def copy_tests(force=False):
    """The reference's tests that pin the boundary, unmodified."""
    dst = os.path.join(OUT, "test")
    src = os.path.join(REF, "test")
    if os.path.isdir(dst) and not force:
        return dst
    if not os.path.isdir(src):
        return dst if os.path.isdir(dst) else None
    if os.path.isdir(dst):
        _rmtree(dst)
    os.makedirs(dst)
    for f in ("test_ops.py", "test_models.py", "common_utils.py", "conftest.py", "_utils_internal.py",
              "optests_failures_dict.json", "common_extended_utils.py"):
        if os.path.exists(os.path.join(src, f)):
            shutil.copy(os.path.join(src, f), os.path.join(dst, f))
    shutil.copytree(os.path.join(src, "expect"), os.path.join(dst, "expect"))
    os.makedirs(os.path.join(dst, "assets"))
    for f in ("encode_jpeg",):  # (assets the model tests open)
        p = os.path.join(src, "assets", f)
        if os.path.isdir(p):
            shutil.copytree(p, os.path.join(dst, "assets", f))
    _writable(dst)
    return dst


def _writable(path):
    for root, dirs, files in os.walk(path):
        for n in dirs + files:
            os.chmod(os.path.join(root, n), 0o755 if n in dirs else 0o644)


def _rmtree(path):
    _writable(path)
    shutil.rmtree(path)


def build_stock_cuda(force=False):
    """The reference's CUDA kernels of the path, unmodified, for sm_100a."""
    so = os.path.join(OUT, "tvref_cuda.so")
    if os.path.exists(so) and not force:
        return so
    cuda_dir = os.path.join(REF, "torchvision", "csrc", "ops", "cuda")
    if not os.path.isdir(cuda_dir):
        return so if os.path.exists(so) else None
    csrc, inc, abi, torch_lib = _flags()
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(cuda_dir, f) for f in ("roi_align_kernel.cu", "nms_kernel.cu")]
    cmd = [NVCC, "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "--expt-relaxed-constexpr",
           "-Xcompiler", "-fPIC", abi, "-DTORCH_EXTENSION_NAME=tvref_cuda", "-shared"] + inc + \
          ["-I/usr/local/cuda/include"] + srcs + \
          ["-o", so, "-L" + torch_lib, "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda",
           "-Xlinker", "-rpath," + torch_lib]
    subprocess.check_call(cmd)
    return so


def main():
    force = "--force" in sys.argv
    print("package:", build_package(force))
    print("reference tests:", copy_tests(force))
    if "--cuda" in sys.argv:
        print("stock cuda kernels:", build_stock_cuda(force))


if __name__ == "__main__":
    main()
