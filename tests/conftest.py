import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import pcnn_b200 as pc
        return pc.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # -m gpu on a box without a device must fail loudly, not skip silently; but a plain
    # `pytest` run here (no -m) should not try the GPU tests.
    if config.getoption("-m"):
        return
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_port():
    import pcoracle as po
    if not po.have("port"):
        po.build_oracle(ref=False)
    return po.Oracle("port")


@pytest.fixture(scope="session")
def oracle_ref():
    import pcoracle as po
    if not po.have("ref"):
        if os.path.isdir("/root/reference/Homework2/hw2/include"):
            po.build_oracle(ref=True)
        else:
            pytest.skip("oracle/_ref not built and the reference is not mounted")
    return po.Oracle("ref")


@pytest.fixture(scope="session")
def kitti():
    return np.load(os.path.join(GOLDEN, "kitti_100k.npy"))


@pytest.fixture(scope="session")
def golden_nn():
    return np.load(os.path.join(GOLDEN, "golden_nn.npz"))


@pytest.fixture(scope="session")
def golden_iss():
    return np.load(os.path.join(GOLDEN, "golden_iss.npz"))


@pytest.fixture(scope="session")
def golden_normals():
    return np.load(os.path.join(GOLDEN, "golden_normals.npz"))


@pytest.fixture(scope="session")
def golden_icp():
    return np.load(os.path.join(GOLDEN, "golden_icp.npz"))


@pytest.fixture(scope="session")
def hostsim():
    import ctypes as C
    d = os.path.join(ROOT, "tests", "hostsim")
    so = os.path.join(d, "libpcnn_hostsim.so")
    src = os.path.join(d, "hostsim.cpp")
    csrc = os.path.join(ROOT, "hands-on-point-cloud-processing_b200", "csrc")
    newest = max(os.path.getmtime(p) for p in [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")])
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off",
                               "-I/usr/local/cuda/include", "-I" + csrc, "-o", so, src])
    L = C.CDLL(so)
    L.hostsim_build.restype = C.c_void_p
    L.hostsim_build.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.hostsim_destroy.argtypes = [C.c_void_p]
    L.hostsim_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.hostsim_nn1.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.hostsim_order.argtypes = [C.c_void_p, C.c_void_p]
    return L


def _build_compat(tmp_path_factory, name):
    out = str(tmp_path_factory.mktemp(name) / name)
    pkg = os.path.join(ROOT, "hands-on-point-cloud-processing_b200")
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "compat", name + ".cpp"), "-o", out,
                           "-L" + pkg, "-lpcnn", "-Wl,-rpath," + pkg])
    return out


@pytest.fixture(scope="session")
def registration_binary(tmp_path_factory):
    """tests/compat/registration_main.cpp (a caller in the style of Homework9's main.cpp) built against
    include/pcnn_registration_compat.hpp with mock Eigen / PCL types, linked with libpcnn.so."""
    return _build_compat(tmp_path_factory, "registration_main")


@pytest.fixture(scope="session")
def compat_binary(tmp_path_factory):
    """tests/compat/compat_main.cpp (a caller in the reference's own style) built against
    include/pcnn_compat.hpp and linked with libpcnn.so."""
    out = str(tmp_path_factory.mktemp("compat") / "compat_main")
    pkg = os.path.join(ROOT, "hands-on-point-cloud-processing_b200")
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "compat", "compat_main.cpp"), "-o", out,
                           "-L" + pkg, "-lpcnn", "-Wl,-rpath," + pkg])
    return out
