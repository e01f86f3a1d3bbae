import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    need = [os.path.join(ROOT, "raytracer2_b200", "librt2host.so"), os.path.join(ROOT, "oracle", "librt2oracle.so"),
            os.path.join(ROOT, "raytracer2_b200", "librt2cuda.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.check_call(["make", "-C", ROOT, "all"], stdout=subprocess.DEVNULL)


@pytest.fixture(scope="session", autouse=True)
def built_libs():
    _ensure_built()


@pytest.fixture(scope="session")
def hostemu():
    """tests/hostemu: the product's host/device builder + traversal logic compiled for the CPU (test only)."""
    import ctypes as C
    d = os.path.join(ROOT, "tests", "hostemu")
    so = os.path.join(d, "libhostemu.so")
    src = os.path.join(d, "hostemu.cu")
    deps = [src] + [os.path.join(ROOT, "raytracer2_b200", "csrc", f) for f in ("rt2_build.cuh", "rt2_reinsert.cuh", "rt2_bvh.cuh", "rt2_math.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        nvcc = "/usr/local/cuda/bin/nvcc"
        if not os.path.exists(nvcc):
            pytest.skip("nvcc not available to build the host emulation")
        subprocess.check_call([nvcc, "-O2", "-std=c++17", "-fmad=false", "-Wno-deprecated-gpu-targets", "-Xcompiler",
                               "-fPIC,-ffp-contract=off,-mfma,-Wno-unknown-pragmas", "-shared", "-o", so, src])
    L = C.CDLL(so)
    L.hostemu_build.restype = C.c_void_p
    L.hostemu_build.argtypes = [C.c_void_p]
    L.hostemu_destroy.argtypes = [C.c_void_p]
    L.hostemu_trace_primary.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]
    L.hostemu_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_float, C.c_void_p, C.POINTER(C.c_uint64)]
    L.hostemu_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_uint32)] * 3
    L.hostemu_reinsertion_probe.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
    return L


def has_gpu():
    try:
        import ctypes as C
        lib = C.CDLL("libcudart.so.12")
        n = C.c_int(0)
        return lib.cudaGetDeviceCount(C.byref(n)) == 0 and n.value > 0
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False
