"""The reference's own Python names for this path, backed by the CUDA library.

Mirrors Homework2/hw2/script/result_set.py:16-89, kdtree.py:183-250 and octree.py:109-194:
same function names, same argument order and meaning.  ``root`` is an ``Index`` instead of a
``Node`` / ``Octant``; ``db`` is accepted and ignored by the searches (the reference passes it
to every call, kdtree.py:183) because the index owns a copy of the cloud.

The single-query functions fill a result set exactly as the reference does (one launch per
query: correct, slow).  The ``*_batch`` functions hoist the loop of benchmark.py:58-72 into
one call and are the intended fast path.
"""
import bisect

import numpy as np

from .api import KDTREE, OCTREE
from .index import Index


class DistIndex:
    """result_set.py:4-13"""

    def __init__(self, distance, index):
        self.distance = distance
        self.index = index

    def __lt__(self, other):
        return self.distance < other.distance

    def __str__(self):
        return "Distance = %s, Index = %s" % (self.distance, self.index)


class KNNResultSet:
    """result_set.py:16-61: bounded ascending list, capacity k, slots start at (1e10, 0)."""

    def __init__(self, capacity):
        self.capacity = capacity
        self.count = 0
        self.worst_dist = 1e10
        self.dist_index_list = [DistIndex(self.worst_dist, 0) for _ in range(capacity)]
        self.comparision_count = 0

    def size(self):
        return self.count

    def full(self):
        return self.count == self.capacity

    def get_worst_dist(self):
        return self.worst_dist

    def list(self):
        for item in self.dist_index_list:
            print(item)

    def add_point(self, dist, index):
        """Same observable behaviour as result_set.py:41-61 (a candidate equal to the worst distance
        is accepted, lands behind entries of equal distance, and the entry pushed past the capacity
        is dropped), written as a binary-search insert over the live prefix."""
        self.comparision_count += 1
        if dist > self.worst_dist:
            return
        live = self.dist_index_list[:self.count]
        if self.count == self.capacity:
            live.pop()                      # full: the current last entry makes room
        else:
            self.count += 1
        at = bisect.bisect_right([e.distance for e in live], dist)
        live.insert(at, DistIndex(dist, index))
        self.dist_index_list[:self.count] = live
        self.worst_dist = self.dist_index_list[self.capacity - 1].distance

    def _fill(self, idx_row, dist_row):
        for j in range(self.capacity):
            if idx_row[j] >= 0:
                self.dist_index_list[j] = DistIndex(float(dist_row[j]), int(idx_row[j]))
                self.count = j + 1
        self.worst_dist = self.dist_index_list[self.capacity - 1].distance


class RadiusNNResultSet:
    """result_set.py:64-89: unbounded list of hits with dist <= radius."""

    def __init__(self, radius):
        self.radius = radius
        self.count = 0
        self.worst_dist = radius
        self.dist_index_list = []
        self.comparision_count = 0

    def get_worst_dist(self):
        return self.radius

    def size(self):
        return self.count

    def list(self):
        for item in self.dist_index_list:
            print(item)

    def add_point(self, distance, index):
        """result_set.py:83-89: every candidate is counted, those within the radius (inclusive) kept."""
        self.comparision_count += 1
        if distance <= self.radius:
            self.dist_index_list.append(DistIndex(distance, index))
            self.count = len(self.dist_index_list)


def kdtree_construction(db_np, leaf_size):
    """kdtree.py:240-250"""
    return Index(db_np, kind=KDTREE, leaf_size=leaf_size)


def octree_construction(db_np, leaf_size, min_extent):
    """octree.py:184-194"""
    return Index(db_np, kind=OCTREE, leaf_size=leaf_size, min_extent=min_extent)


def _knn_single(root, result_set, query):
    q = np.asarray(query, dtype=np.float32).reshape(1, -1)[:, :3]
    idx, dist = root.knn(q, result_set.capacity, dist="f64")
    result_set._fill(idx[0], dist[0])
    return False


def _radius_single(root, result_set, query):
    q = np.asarray(query, dtype=np.float32).reshape(1, -1)[:, :3]
    _, idx, dist = root.radius(q, result_set.radius, dist="f64")
    for d, i in zip(dist, idx):
        result_set.add_point(float(d), int(i))
    return False


def kdtree_knn_search(root, db, result_set, query):
    """kdtree.py:183-204"""
    return _knn_single(root, result_set, query)


def kdtree_radius_search(root, db, result_set, query):
    """kdtree.py:207-228"""
    return _radius_single(root, result_set, query)


def octree_knn_search(root, db, result_set, query):
    """octree.py:109-143"""
    return _knn_single(root, result_set, query)


def octree_radius_search(root, db, result_set, query):
    """octree.py:146-181"""
    return _radius_single(root, result_set, query)


def knn_search_batch(root, queries, k, dist="f64"):
    """The loop of benchmark.py:58-66 as one call: (idx[nq,k], dist[nq,k])."""
    return root.knn(queries, k, dist=dist)


def radius_search_batch(root, queries, radius, dist="f64"):
    """The loop of benchmark.py:68-71 as one call: CSR (offsets, idx, dist)."""
    return root.radius(queries, radius, dist=dist)


def read_velodyne_bin(path, duplicate_last=False):
    """KITTI .bin (x, y, z, intensity float32) -> (N, 3) float32; kdtree.py:253-264 and the C++
    ``readBinary`` (Homework2/hw2/include/test.hpp:11-33), which reads whole 16-byte records.
    ``duplicate_last=True`` reproduces what that C++ loop really returns: it tests ``eof`` only
    after a read has failed, so one more point is pushed from the stale record buffer -- the last
    point again (``124669 points read`` for the 124668-record frame), with its leading bytes
    replaced by whatever partial record trails the file.  An empty file yields no points (the
    reference would push one point of uninitialised stack)."""
    raw = np.fromfile(path, dtype=np.uint8)
    nrec, rem = divmod(len(raw), 16)
    rec = raw[: nrec * 16].view(np.float32).reshape(-1, 4)
    pts = rec[:, :3]
    if duplicate_last and nrec:
        stale = raw[(nrec - 1) * 16: nrec * 16].copy()
        stale[:rem] = raw[nrec * 16:]
        pts = np.concatenate([pts, stale.view(np.float32).reshape(1, 4)[:, :3]], axis=0)
    return np.ascontiguousarray(pts, dtype=np.float32)


# ------------------------------------------------------------------ cloud ingest + voxel grid
def read_xyz_normal_bin(path, trailing_record=True):
    """The 6-float (x, y, z, nx, ny, nz float32) record file of Homework9 -> (xyz [N,3], normals
    [N,3]); the reader of readBinaryAndVoxelDown (Homework9/hw9/src/registration.cpp:12-29).
    Its loop `for (; input.good() and !input.eof();)` runs once more after the last record, and
    the read that fails leaves a default-constructed pcl::PointNormal (all zeros) in the cloud:
    ``trailing_record=True`` reproduces that extra (0,0,0) point with a zero normal."""
    raw = np.fromfile(path, dtype=np.float32)
    raw = raw[: (len(raw) // 6) * 6].reshape(-1, 6)
    if trailing_record:
        raw = np.concatenate([raw, np.zeros((1, 6), np.float32)], axis=0)
    return np.ascontiguousarray(raw[:, :3]), np.ascontiguousarray(raw[:, 3:])


def voxel_downsample(xyz, normals=None, leaf=1.75, device=0):
    """pcl::VoxelGrid<pcl::PointNormal> as the reference configures it (registration.cpp:37-41,
    :688-691; leaf 1.75 is VoxelGridSampling's): one point per occupied voxel, ascending voxel
    index.  numpy in -> numpy out, CUDA tensors in -> CUDA tensors out.  Returns (xyz, normals)."""
    import ctypes as C
    from . import api
    from .api import _check, _is_torch, _np_ptr, _stream, _t_ptr, lib
    lx, ly, lz = (leaf, leaf, leaf) if np.isscalar(leaf) else leaf
    count = C.c_int64(0)
    if _is_torch(xyz):
        import torch
        x = xyz.contiguous()
        assert x.dtype == torch.float32 and x.dim() == 2 and x.shape[1] >= 3
        nr = None if normals is None else normals.contiguous()
        n = x.shape[0]
        ox = torch.empty((max(n, 1), 3), dtype=torch.float32, device=x.device)
        on = None if nr is None else torch.empty((max(n, 1), 3), dtype=torch.float32, device=x.device)
        _check(lib().pcnn_voxel_downsample(_t_ptr(x), _t_ptr(nr), n, x.shape[1], 3 if nr is None else nr.shape[1],
                                           lx, ly, lz, _t_ptr(ox), _t_ptr(on), n, C.byref(count), x.device.index,
                                           _stream()))
        return ox[:count.value], (None if on is None else on[:count.value])
    import torch
    dev = torch.device("cuda", device)
    tx = torch.from_numpy(np.ascontiguousarray(xyz, np.float32)).to(dev)
    tn = None if normals is None else torch.from_numpy(np.ascontiguousarray(normals, np.float32)).to(dev)
    ox, on = voxel_downsample(tx, tn, (lx, ly, lz))
    return ox.cpu().numpy(), (None if on is None else on.cpu().numpy())


def read_binary_and_voxel_down(path, voxel_size, device=0):
    """readBinaryAndVoxelDown(fileName, cloud, normals, voxel_size) (registration.cpp:8-68)."""
    xyz, nrm = read_xyz_normal_bin(path)
    return voxel_downsample(xyz, nrm, voxel_size, device)


# ------------------------------------------------------------------ global initialisation
def _cuda(a, dtype, device):
    import torch
    if isinstance(a, torch.Tensor):
        return a.to(dtype=dtype).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype=dtype, device=torch.device("cuda", device)).contiguous()


def desc_nn1(a, b, device=0):
    """Exact float32 1-NN of every row of ``a`` among the rows of ``b`` with nanoflann's arithmetic
    (registration.cpp:458-492).  Returns CUDA tensors (idx int32 [na], d2 float32 [na])."""
    import torch
    from .api import _check, _stream, _t_ptr, lib
    ta, tb = _cuda(a, torch.float32, device), _cuda(b, torch.float32, device)
    idx = torch.empty(max(len(ta), 1), dtype=torch.int32, device=ta.device)
    d2 = torch.empty(max(len(ta), 1), dtype=torch.float32, device=ta.device)
    _check(lib().pcnn_desc_nn1(_t_ptr(ta), len(ta), _t_ptr(tb), len(tb), ta.shape[1], _t_ptr(idx), _t_ptr(d2),
                               ta.device.index, _stream()))
    return idx[:len(ta)], d2[:len(ta)]


def match_descriptors(src_desc, tar_desc, mode="inter", rejection_rate=0.0, device=0):
    """findRANSACCorrespondencesInter / ...Union (registration.cpp:437-527 / :530-611):
    [(source index, target index)] int32 CUDA tensor, ascending descriptor distance."""
    import ctypes as C
    import torch
    from .api import _check, _stream, _t_ptr, lib
    ts, tt = _cuda(src_desc, torch.float32, device), _cuda(tar_desc, torch.float32, device)
    cap = len(ts) + len(tt)
    out = torch.empty((max(cap, 1), 2), dtype=torch.int32, device=ts.device)
    count = C.c_int64(0)
    _check(lib().pcnn_match_descriptors(_t_ptr(ts), len(ts), _t_ptr(tt), len(tt), ts.shape[1] if ts.dim() == 2 else 0,
                                        {"inter": 0, "union": 1}[mode], float(rejection_rate), _t_ptr(out), cap,
                                        C.byref(count), ts.device.index, _stream()))
    return out[:count.value]


def ransac(src_keypoints, tar_keypoints, pairs, max_iter=80000, dist_threshold=0.5, coplanar_min=0.15, seed=0, device=0):
    """Registration::RANSAC (registration.cpp:288-434).  Returns a dict (R [3,3], t [3], consensus, best_iter)."""
    import ctypes as C
    import torch
    from . import api
    from .api import _check, _stream, _t_ptr, lib
    s = _cuda(src_keypoints, torch.float32, device)[:, :3].contiguous()
    t = _cuda(tar_keypoints, torch.float32, device)[:, :3].contiguous()
    p = _cuda(pairs, torch.int32, device)
    prm = api.RansacParams(max_iter, dist_threshold, coplanar_min, 0, seed)
    res = api.RansacResult()
    _check(lib().pcnn_ransac(_t_ptr(s), _t_ptr(t), _t_ptr(p), len(p), C.byref(prm), C.byref(res), s.device.index, _stream()))
    return {"R": np.array(list(res.R), np.float32).reshape(3, 3), "t": np.array(list(res.t), np.float32),
            "consensus": int(res.consensus), "best_iter": int(res.best_iter)}
