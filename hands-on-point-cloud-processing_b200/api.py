"""ctypes binding of libpcnn.so (include/pcnn.h) -- the host-side mirror of the reference's
nearest-neighbour interface.

Two layers:

* ``Index`` / ``IcpState``: thin object wrappers over the C ABI.  Host arrays (numpy) go
  through the ``*_host`` entry points; CUDA tensors (torch) go through the device-pointer
  entry points on torch's current stream.  torch is used for device memory and streams
  only.
* The reference's own Python names (Homework2/hw2/script/kdtree.py:183-250,
  octree.py:109-194, result_set.py:16-89): ``kdtree_construction``, ``kdtree_knn_search``,
  ``kdtree_radius_search``, ``octree_construction``, ``octree_knn_search``,
  ``octree_radius_search``, ``KNNResultSet``, ``RadiusNNResultSet`` -- same arguments, same
  meaning, so a caller of the reference scripts can switch by changing the import.

There is no CPU fallback: if libpcnn.so is missing or no CUDA device is visible, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PCNN_LIB: developer override to load another BUILD of the same library (tools/tune_*.py)
_LIB_PATH = os.environ.get("PCNN_LIB") or os.path.join(_HERE, "libpcnn.so")
_lib = None

KDTREE, OCTREE = 0, 1
OK, ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED, ERR_CAPACITY = 0, -1, -2, -3, -4, -5
MAX_K = 128
ICP_NSUMS = 32
ICP_PEER_HANDLE_BYTES = 64  # PCNN_ICP_PEER_HANDLE_BYTES


class PcnnError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pcnn error %d: %s" % (code, msg))
        self.code = code


class IndexInfo(C.Structure):
    _fields_ = [("n_points", C.c_int64), ("kind", C.c_int32), ("device", C.c_int32),
                ("n_levels", C.c_int32), ("group_size", C.c_int32), ("device_bytes", C.c_int64),
                ("bbox_min", C.c_float * 3), ("bbox_max", C.c_float * 3), ("build_ms", C.c_float)]


class IcpParams(C.Structure):
    _fields_ = [("max_iter", C.c_int32), ("max_corres_dist", C.c_float),
                ("loss_epsilon", C.c_float), ("mode", C.c_int32)]


class IcpResult(C.Structure):
    _fields_ = [("T", C.c_float * 16), ("iters", C.c_int32), ("converged", C.c_int32),
                ("status", C.c_int32), ("last_loss", C.c_float), ("last_pairs", C.c_int64),
                ("loop_ms", C.c_float), ("last_searches", C.c_int32)]


class IssParams(C.Structure):
    _fields_ = [("local_radius", C.c_float), ("non_max_radius", C.c_float), ("gamma21", C.c_float),
                ("gamma32", C.c_float), ("min_neighbors", C.c_int32), ("weighted", C.c_int32)]


class RansacParams(C.Structure):
    _fields_ = [("max_iter", C.c_int32), ("dist_threshold", C.c_float), ("coplanar_min", C.c_float),
                ("reserved", C.c_int32), ("seed", C.c_uint64)]


class RansacResult(C.Structure):
    _fields_ = [("R", C.c_float * 9), ("t", C.c_float * 3), ("consensus", C.c_int64), ("best_iter", C.c_int64)]


EXPORTS = [
    "pcnn_last_error", "pcnn_version", "pcnn_device_count", "pcnn_index_build",
    "pcnn_index_build_host", "pcnn_index_destroy", "pcnn_index_get_info", "pcnn_knn",
    "pcnn_knn_host", "pcnn_set_option", "pcnn_get_profile", "pcnn_radius_count", "pcnn_radius_fill",
    "pcnn_radius_host", "pcnn_icp", "pcnn_icp_host", "pcnn_icp_create", "pcnn_icp_accumulate",
    "pcnn_icp_sums_dev", "pcnn_icp_update", "pcnn_icp_finish", "pcnn_icp_destroy", "pcnn_nn1_f32",
    "pcnn_release_cached", "pcnn_debug_warp_times", "pcnn_iss_detect", "pcnn_iss_detect_host",
    "pcnn_voxel_downsample", "pcnn_estimate_normals", "pcnn_estimate_normals_host",
    "pcnn_desc_nn1", "pcnn_match_descriptors", "pcnn_ransac",
    "pcnn_icp_peer_export", "pcnn_icp_peer_attach", "pcnn_icp_step_peer", "pcnn_icp_peer_detach",
    "pcnn_radius_count_noalloc",
]


def lib():
    """Load libpcnn.so; raises (loudly) if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ImportError("libpcnn.so not found at %s: build it with `python -c 'import "
                          "__graft_entry__ as g; g.build()'` (there is no CPU fallback)" % _LIB_PATH)
    L = C.CDLL(_LIB_PATH)
    vp, i32, i64, f32, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double
    L.pcnn_last_error.restype = C.c_char_p
    L.pcnn_last_error.argtypes = []
    L.pcnn_version.restype = i32
    L.pcnn_device_count.restype = i32
    L.pcnn_index_build.argtypes = [C.POINTER(vp), i32, vp, i64, i32, i32, f64, i32, vp]
    L.pcnn_index_build_host.argtypes = [C.POINTER(vp), i32, vp, i64, i32, i32, f64, i32]
    L.pcnn_index_destroy.argtypes = [vp]
    L.pcnn_index_get_info.argtypes = [vp, C.POINTER(IndexInfo)]
    L.pcnn_knn.argtypes = [vp, vp, i64, i32, i32, vp, vp, vp, vp]
    L.pcnn_knn_host.argtypes = [vp, vp, i64, i32, i32, vp, vp, vp]
    L.pcnn_set_option.argtypes = [vp, C.c_char_p, i32]
    L.pcnn_get_profile.argtypes = [vp, C.POINTER(C.c_float)]
    L.pcnn_radius_count.argtypes = [vp, vp, i64, i32, f64, vp, C.POINTER(i64), vp]
    L.pcnn_radius_fill.argtypes = [vp, vp, i64, i32, f64, vp, vp, vp, vp, vp]
    L.pcnn_radius_count_noalloc.argtypes = [vp, vp, i64, i32, f64, vp, vp]
    L.pcnn_radius_host.argtypes = [vp, vp, i64, i32, f64, vp, vp, vp, vp, i64]
    L.pcnn_icp.argtypes = [vp, vp, vp, i64, i32, vp, C.POINTER(IcpParams), C.POINTER(IcpResult), vp]
    L.pcnn_icp_host.argtypes = [vp, vp, vp, i64, i32, vp, C.POINTER(IcpParams), C.POINTER(IcpResult)]
    L.pcnn_icp_create.argtypes = [C.POINTER(vp), vp, vp, vp, i64, i32, vp, C.POINTER(IcpParams), i64, vp]
    L.pcnn_icp_accumulate.argtypes = [vp, vp]
    L.pcnn_icp_sums_dev.restype = vp
    L.pcnn_icp_sums_dev.argtypes = [vp]
    L.pcnn_icp_update.argtypes = [vp, vp]
    L.pcnn_icp_finish.argtypes = [vp, C.POINTER(IcpResult), vp]
    L.pcnn_icp_destroy.argtypes = [vp]
    L.pcnn_icp_peer_export.argtypes = [vp, vp, C.POINTER(vp)]
    L.pcnn_icp_peer_attach.argtypes = [vp, i32, i32, vp, C.POINTER(vp)]
    L.pcnn_icp_step_peer.argtypes = [vp, vp]
    L.pcnn_icp_peer_detach.argtypes = [vp]
    L.pcnn_nn1_f32.argtypes = [vp, vp, i64, i32, vp, vp, vp]
    L.pcnn_release_cached.argtypes = [i32]
    L.pcnn_debug_warp_times.argtypes = [vp, vp]
    L.pcnn_iss_detect.argtypes = [vp, C.POINTER(IssParams), vp, i64, C.POINTER(i64), vp, vp]
    L.pcnn_iss_detect_host.argtypes = [vp, C.POINTER(IssParams), vp, i64, C.POINTER(i64), vp]
    L.pcnn_estimate_normals.argtypes = [vp, i32, f64, vp, vp]
    L.pcnn_estimate_normals_host.argtypes = [vp, i32, f64, vp]
    L.pcnn_desc_nn1.argtypes = [vp, i64, vp, i64, i32, vp, vp, i32, vp]
    L.pcnn_match_descriptors.argtypes = [vp, i64, vp, i64, i32, i32, f32, vp, i64, C.POINTER(i64), i32, vp]
    L.pcnn_ransac.argtypes = [vp, vp, vp, i64, C.POINTER(RansacParams), C.POINTER(RansacResult), i32, vp]
    L.pcnn_voxel_downsample.argtypes = [vp, vp, i64, i32, i32, f32, f32, f32, vp, vp, i64, C.POINTER(i64), i32, vp]
    for name in EXPORTS:
        f = getattr(L, name)
        if name not in ("pcnn_last_error", "pcnn_icp_sums_dev"):
            f.restype = i32
    _lib = L
    return L


def _check(rc):
    if rc != OK:
        raise PcnnError(rc, lib().pcnn_last_error().decode("utf-8", "replace"))


def device_count():
    return int(lib().pcnn_device_count())


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _host_xyz(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim == 1 and a.size in (3, 4):
        a = a.reshape(1, -1)
    if a.ndim != 2 or a.shape[1] < 3:
        raise ValueError("expected an (N, >=3) float32 array, got shape %r" % (a.shape,))
    return a


def _dev_xyz(t):
    import torch
    if t.dtype != torch.float32 or not t.is_cuda:
        raise ValueError("expected a float32 CUDA tensor")
    if t.dim() != 2 or t.shape[1] < 3:
        raise ValueError("expected an (N, >=3) tensor, got %r" % (tuple(t.shape),))
    return t.contiguous()


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _t_ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())
