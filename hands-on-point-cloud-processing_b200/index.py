"""Object wrappers over the C ABI: Index (build / kNN / radius / 1-NN / ICP) and IcpState
(split-phase ICP for source clouds sharded over several GPUs)."""
import ctypes as C

import numpy as np

from . import api
from .api import (IcpParams, IcpResult, IndexInfo, IssParams, KDTREE, OCTREE, _check, _dev_xyz, _host_xyz,
                  _is_torch, _np_ptr, _stream, _t_ptr, lib)


class Index:
    """Spatial index over a reference cloud.

    Replaces ``KDTreeConstruction(db, leaf_size)`` (Homework2/hw2/include/kdtree.hpp:419-429)
    and ``OctreeConstruction(db, leaf_size, min_extent)`` (octree.hpp:242-271).
    ``xyz`` may be a numpy array (copied to ``device``) or a float32 CUDA tensor.
    """

    def __init__(self, xyz, kind=KDTREE, leaf_size=32, min_extent=1e-4, device=None):
        L = lib()
        self._h = C.c_void_p()
        if _is_torch(xyz):
            t = _dev_xyz(xyz)
            dev = t.device.index if device is None else device
            _check(L.pcnn_index_build(C.byref(self._h), kind, _t_ptr(t), t.shape[0], t.shape[1],
                                      leaf_size, float(min_extent), dev, _stream()))
            self._keep = t  # until the stream has consumed it
        else:
            a = _host_xyz(xyz) if len(xyz) else np.zeros((0, 3), np.float32)
            dev = 0 if device is None else device
            _check(L.pcnn_index_build_host(C.byref(self._h), kind, _np_ptr(a), a.shape[0],
                                           a.shape[1], leaf_size, float(min_extent), dev))
        self.kind, self.device = kind, dev

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().pcnn_index_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        out = IndexInfo()
        _check(lib().pcnn_index_get_info(self._h, C.byref(out)))
        return {"n_points": out.n_points, "kind": out.kind, "device": out.device,
                "n_levels": out.n_levels, "group_size": out.group_size,
                "device_bytes": out.device_bytes, "bbox_min": list(out.bbox_min),
                "bbox_max": list(out.bbox_max), "build_ms": out.build_ms}

    def set_option(self, name, value):
        _check(lib().pcnn_set_option(self._h, name.encode(), int(value)))

    def last_kernel_ms(self):
        """Device time of the last main search kernel (needs set_option("profile", 1))."""
        ms = C.c_float(0)
        _check(lib().pcnn_get_profile(self._h, C.byref(ms)))
        return float(ms.value)

    # ------------------------------------------------------------------ kNN
    def knn(self, q, k, dist="f64", out=None):
        """k nearest neighbours of every query: (idx[nq,k] int32, dist[nq,k]).

        Batched ``KNNResultSet rs(k); K?TreeKNNSearch(root, db, rs, q)``
        (benchmark.hpp:61-66).  ``dist``: "f64" (the reference's type), "f32" or None.
        """
        L = lib()
        if _is_torch(q):
            import torch
            t = _dev_xyz(q)
            nq = t.shape[0]
            if out is not None:
                idx, d = out
            else:
                idx = torch.empty((nq, k), dtype=torch.int32, device=t.device)
                d = None if dist is None else torch.empty(
                    (nq, k), dtype=torch.float64 if dist == "f64" else torch.float32, device=t.device)
            d32 = d if (d is not None and d.dtype == torch.float32) else None
            d64 = d if (d is not None and d.dtype == torch.float64) else None
            _check(L.pcnn_knn(self._h, _t_ptr(t), nq, t.shape[1], k, _t_ptr(idx), _t_ptr(d32),
                              _t_ptr(d64), _stream()))
            return idx, d
        a = _host_xyz(q)
        nq = a.shape[0]
        if out is not None:
            idx, d = out
        else:
            idx = np.empty((nq, k), np.int32)
            d = None if dist is None else np.empty((nq, k), np.float64 if dist == "f64" else np.float32)
        d32 = d if (d is not None and d.dtype == np.float32) else None
        d64 = d if (d is not None and d.dtype == np.float64) else None
        _check(L.pcnn_knn_host(self._h, _np_ptr(a), nq, a.shape[1], k, _np_ptr(idx), _np_ptr(d32),
                               _np_ptr(d64)))
        return idx, d

    # ------------------------------------------------------------------ radius
    def radius(self, q, radius, dist="f64"):
        """All neighbours within ``radius`` (inclusive): CSR (offsets[nq+1], idx, dist).

        Batched ``RadiusNNResultSet rs(r); K?TreeRadiusNNSearch(...)`` (benchmark.hpp:68-71).
        The order inside a row is deterministic but unspecified (the reference's is its own
        traversal order): compare rows as sets.
        """
        L = lib()
        if _is_torch(q):
            import torch
            t = _dev_xyz(q)
            nq = t.shape[0]
            off = torch.empty(nq + 1, dtype=torch.int64, device=t.device)
            total = C.c_int64(0)
            _check(L.pcnn_radius_count(self._h, _t_ptr(t), nq, t.shape[1], float(radius), _t_ptr(off),
                                       C.byref(total), _stream()))
            n = int(total.value)
            idx = torch.empty(n, dtype=torch.int32, device=t.device)
            d = None if dist is None else torch.empty(
                n, dtype=torch.float64 if dist == "f64" else torch.float32, device=t.device)
            d32 = d if (d is not None and d.dtype == torch.float32) else None
            d64 = d if (d is not None and d.dtype == torch.float64) else None
            _check(L.pcnn_radius_fill(self._h, _t_ptr(t), nq, t.shape[1], float(radius), _t_ptr(off),
                                      _t_ptr(idx), _t_ptr(d32), _t_ptr(d64), _stream()))
            return off, idx, d
        a = _host_xyz(q)
        nq = a.shape[0]
        off = np.zeros(nq + 1, np.int64)
        rc = L.pcnn_radius_host(self._h, _np_ptr(a), nq, a.shape[1], float(radius), _np_ptr(off),
                                None, None, None, 0)
        if rc not in (api.OK, api.ERR_CAPACITY):
            _check(rc)
        n = int(off[nq])
        idx = np.empty(n, np.int32)
        d = None if dist is None else np.empty(n, np.float64 if dist == "f64" else np.float32)
        d32 = d if (d is not None and d.dtype == np.float32) else None
        d64 = d if (d is not None and d.dtype == np.float64) else None
        if n > 0:
            _check(L.pcnn_radius_host(self._h, _np_ptr(a), nq, a.shape[1], float(radius), _np_ptr(off),
                                      _np_ptr(idx), _np_ptr(d32), _np_ptr(d64), n))
        return off, idx, d

    # ------------------------------------------------------------------ float32 1-NN
    def nn1(self, q):
        """Exact float32 1-NN as the ICP loop uses it: (idx[nq] int32, d2[nq] float32)."""
        import torch
        t = _dev_xyz(q) if _is_torch(q) else torch.from_numpy(_host_xyz(q)).cuda(self.device)
        idx = torch.empty(t.shape[0], dtype=torch.int32, device=t.device)
        d2 = torch.empty(t.shape[0], dtype=torch.float32, device=t.device)
        _check(lib().pcnn_nn1_f32(self._h, _t_ptr(t), t.shape[0], t.shape[1], _t_ptr(idx), _t_ptr(d2),
                                  _stream()))
        if _is_torch(q):
            return idx, d2
        return idx.cpu().numpy(), d2.cpu().numpy()

    # ------------------------------------------------------------------ normals by PCA over kNN
    def estimate_normals(self, max_nn=10, radius=5.0, device_out=False):
        """Normals of the indexed cloud: the per-point loop of pca_normal.py:85-101 (hybrid search
        radius=5, max_nn=10 there).  [n, 3] float32, original point order, sign arbitrary."""
        n = self.info()["n_points"]
        if device_out:
            import torch
            out = torch.empty((n, 3), dtype=torch.float32, device=torch.device("cuda", self.device))
            _check(lib().pcnn_estimate_normals(self._h, max_nn, float(radius), _t_ptr(out), _stream()))
            return out
        out = np.empty((n, 3), np.float32)
        _check(lib().pcnn_estimate_normals_host(self._h, max_nn, float(radius), _np_ptr(out)))
        return out

    # ------------------------------------------------------------------ ISS keypoints
    def iss_keypoints(self, local_radius, non_max_radius, gamma21=0.9, gamma32=0.9, min_neighbors=5,
                      weighted=True, device_out=False):
        """ISS keypoints of the indexed cloud: ``ISSKeypoint::compute`` with the setters'
        parameters (Homework7/hw7/src/iss_detector.cpp:38-111; defaults as main.cpp:95-101 uses
        them).  Returns (keypoint indices ascending, lambda3[n] with -1 for non-candidates)."""
        n = self.info()["n_points"]
        prm = IssParams(local_radius, non_max_radius, gamma21, gamma32, min_neighbors, 1 if weighted else 0)
        count = C.c_int64(0)
        if device_out:
            import torch
            dev = torch.device("cuda", self.device)
            idx = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
            lam = torch.empty(max(n, 1), dtype=torch.float32, device=dev)
            _check(lib().pcnn_iss_detect(self._h, C.byref(prm), _t_ptr(idx), n, C.byref(count), _t_ptr(lam),
                                         _stream()))
            return idx[:count.value], lam[:n]
        idx = np.empty(max(n, 1), np.int32)
        lam = np.empty(max(n, 1), np.float32)
        _check(lib().pcnn_iss_detect_host(self._h, C.byref(prm), _np_ptr(idx), n, C.byref(count), _np_ptr(lam)))
        return idx[:count.value].copy(), lam[:n]

    # ------------------------------------------------------------------ ICP
    def icp(self, src, tar_normals=None, init_T=None, max_iter=30, max_corres_dist=1.0,
            loss_epsilon=0.0, mode=0):
        """Registration::ICPpoint2point (mode 0) / ICPpoint2plane (mode 1) against this
        (target) cloud; registration.cpp:862-1011 / :710-859.  Returns a dict."""
        L = lib()
        prm = IcpParams(max_iter, max_corres_dist, loss_epsilon, mode)
        res = IcpResult()
        T0 = np.ascontiguousarray(np.eye(4) if init_T is None else init_T, dtype=np.float32)
        if _is_torch(src):
            import torch
            s = _dev_xyz(src)
            nrm = None
            if tar_normals is not None:
                nrm = tar_normals.contiguous()
                assert nrm.dtype == torch.float32 and nrm.shape[1] == 3
            _check(L.pcnn_icp(self._h, _t_ptr(nrm), _t_ptr(s), s.shape[0], s.shape[1], _np_ptr(T0),
                              C.byref(prm), C.byref(res), _stream()))
        else:
            s = _host_xyz(src)
            nrm = None if tar_normals is None else np.ascontiguousarray(tar_normals, np.float32).reshape(-1, 3)
            _check(L.pcnn_icp_host(self._h, _np_ptr(nrm), _np_ptr(s), s.shape[0], s.shape[1], _np_ptr(T0),
                                   C.byref(prm), C.byref(res)))
        return _icp_dict(res)


def _icp_dict(res):
    return {"T": np.array(list(res.T), np.float32).reshape(4, 4), "iters": res.iters,
            "converged": res.converged, "status": res.status, "loss": res.last_loss,
            "pairs": res.last_pairs, "loop_ms": res.loop_ms, "searches": res.last_searches}


class IcpState:
    """Split-phase ICP: ``accumulate()`` -> all-reduce ``sums`` -> ``update()`` per iteration.

    ``sums`` is a float64 CUDA tensor view of PCNN_ICP_NSUMS doubles: elements [0, 31) are
    summed across ranks; element 31, viewed as int64, takes the MAX across ranks.
    """

    def __init__(self, target, src, tar_normals=None, init_T=None, max_iter=30, max_corres_dist=1.0,
                 loss_epsilon=0.0, mode=0, src_index_base=0):
        import torch
        self.target = target
        self.prm = IcpParams(max_iter, max_corres_dist, loss_epsilon, mode)
        self._h = C.c_void_p()
        s = _dev_xyz(src)
        self._keep = (s, tar_normals)
        T0 = np.ascontiguousarray(np.eye(4) if init_T is None else init_T, dtype=np.float32)
        _check(lib().pcnn_icp_create(C.byref(self._h), target._h, _t_ptr(tar_normals), _t_ptr(s), s.shape[0],
                                     s.shape[1], _np_ptr(T0), C.byref(self.prm), src_index_base, _stream()))
        ptr = lib().pcnn_icp_sums_dev(self._h)
        self.sums = _tensor_from_ptr(ptr, api.ICP_NSUMS, torch.float64, s.device)

    def accumulate(self):
        _check(lib().pcnn_icp_accumulate(self._h, _stream()))

    def update(self):
        _check(lib().pcnn_icp_update(self._h, _stream()))

    def finish(self):
        res = IcpResult()
        _check(lib().pcnn_icp_finish(self._h, C.byref(res), _stream()))
        return _icp_dict(res)

    # ---- fused accumulate + all-reduce + update over peer memory (NVLink), one kernel per iteration
    def export_mailbox(self):
        """(64-byte CUDA IPC handle, device pointer) of this rank's mailbox (pcnn_icp_peer_export)."""
        handle = (C.c_ubyte * api.ICP_PEER_HANDLE_BYTES)()
        ptr = C.c_void_p()
        _check(lib().pcnn_icp_peer_export(self._h, handle, C.byref(ptr)))
        return bytes(handle), ptr.value

    def attach_peers(self, dist, group=None):
        """One process per GPU: exchange the mailbox handles through ``dist`` (host-side, once) and
        map every peer's mailbox.  After this, ``step_peer()`` needs no collective call."""
        import torch
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        handle, _ = self.export_mailbox()
        mine = torch.frombuffer(bytearray(handle), dtype=torch.uint8)
        backend = dist.get_backend(group)
        dev = self._keep[0].device if backend == "nccl" else torch.device("cpu")
        bufs = [torch.empty(api.ICP_PEER_HANDLE_BYTES, dtype=torch.uint8, device=dev) for _ in range(world)]
        dist.all_gather(bufs, mine.to(dev), group=group)
        blob = b"".join(bytes(b.cpu().numpy().tobytes()) for b in bufs)
        arr = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        _check(lib().pcnn_icp_peer_attach(self._h, rank, world, arr, None))
        self._peers = (dist, group)

    def attach_local_peers(self, rank, mailbox_ptrs):
        """Same process (several GPUs driven by one process, or tests): peers by device pointer."""
        world = len(mailbox_ptrs)
        arr = (C.c_void_p * world)(*[C.c_void_p(p) for p in mailbox_ptrs])
        _check(lib().pcnn_icp_peer_attach(self._h, rank, world, None, arr))

    def step_peer(self):
        _check(lib().pcnn_icp_step_peer(self._h, _stream()))

    def detach_peers(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            _check(lib().pcnn_icp_peer_detach(self._h))
        peers = getattr(self, "_peers", None)
        if peers is not None:
            peers[0].barrier(group=peers[1])  # every rank unmaps before any rank frees its mailbox
            self._peers = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            if getattr(self, "_peers", None) is not None:
                self.detach_peers()
            lib().pcnn_icp_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _tensor_from_ptr(ptr, n, dtype, device):
    """Zero-copy torch view of device memory owned by the C library."""
    import torch

    class _Holder:
        pass

    h = _Holder()
    itemsize = torch.empty((), dtype=dtype).element_size()
    typestr = {torch.float64: "<f8", torch.int64: "<i8", torch.float32: "<f4"}[dtype]
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False),
                                  "version": 2, "strides": (itemsize,)}
    return torch.as_tensor(h, device=device)
