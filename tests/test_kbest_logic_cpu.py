"""CPU model check of the k-best containers of the kNN kernels (pcnn_knn.cu: TopKf / topkf_offer, the seed
network of seed_first_group_f, HeapRef::offer / heapify / sort): the LOGIC restated step by step in Python
and run against the canonical answer -- the k smallest under (double key, original index), in that order
(Homework2/hw2/include/resultSet.hpp:65-91) -- on candidate streams full of float32 ties (equal double keys
with different indices, and double keys that differ by less than a float32 ulp).  The CUDA code itself is
checked on the GPU (tests/test_gpu_knn_paths.py); this pins the algorithm it implements."""
import numpy as np

F32 = np.float32
INF = F32(np.inf)


def ru32(d):
    f = F32(d)
    return f if np.float64(f) >= d else np.nextafter(f, INF)


class Cloud:
    """positions -> (double key, original index); what entry_less_exact recomputes from the points"""

    def __init__(self, keys, ids):
        self.keys, self.ids = keys, ids

    def less_exact(self, pa, pb):
        if pa < 0:
            return False
        if pb < 0:
            return True
        return (self.keys[pa], self.ids[pa]) < (self.keys[pb], self.ids[pb])


def topkf_offer(key, pos, f, ppos, cloud):
    K = len(key)
    if f > key[K - 1]:
        return False
    if not any(f == key[m] for m in range(K)):
        for j in range(K - 1, 0, -1):
            cj, cjm = f < key[j], f < key[j - 1]
            key[j], pos[j] = (key[j - 1], pos[j - 1]) if cjm else ((f, ppos) if cj else (key[j], pos[j]))
        if f < key[0]:
            key[0], pos[0] = f, ppos
        return True
    if not (f < key[K - 1] or cloud.less_exact(ppos, pos[K - 1])):
        return False
    key[K - 1], pos[K - 1] = f, ppos
    for j in range(K - 1, 0, -1):
        ka, kb, pa, pb = key[j - 1], key[j], pos[j - 1], pos[j]
        sw = kb < ka
        if kb == ka:
            sw = cloud.less_exact(pb, pa)
        if sw:
            key[j - 1], pos[j - 1], key[j], pos[j] = kb, pb, ka, pa
    return True


NET = [(0, 1), (2, 3), (4, 5), (6, 7), (0, 2), (1, 3), (4, 6), (5, 7), (1, 2), (5, 6), (0, 4), (3, 7), (1, 5), (2, 6),
       (1, 4), (3, 6), (2, 4), (3, 5), (3, 4)]


def seed_first_group(key, pos, fs, first, cloud):
    """eight candidates into an EMPTY list: network, general path on float ties"""
    f, ps = list(fs), list(range(first, first + 8))
    for a, b in NET:
        if f[b] < f[a]:
            f[a], f[b], ps[a], ps[b] = f[b], f[a], ps[b], ps[a]
    if any(f[j] == f[j + 1] for j in range(7)):
        for j in range(8):
            topkf_offer(key, pos, fs[j], first + j, cloud)
        return
    for j in range(min(len(key), 8)):
        key[j], pos[j] = f[j], ps[j]


class Heap:
    def __init__(self, k, cloud):
        self.k, self.cloud, self.K, self.I, self.filled = k, cloud, [INF] * k, [-1] * k, 0

    def less(self, fa, pa, fb, pb):
        return fa < fb or (fa == fb and self.cloud.less_exact(pa, pb))

    def sift(self, i, n, f, p):
        while True:
            c = 2 * i + 1
            if c >= n:
                break
            ck, ci = self.K[c], self.I[c]
            if c + 1 < n and self.less(ck, ci, self.K[c + 1], self.I[c + 1]):
                ck, ci, c = self.K[c + 1], self.I[c + 1], c + 1
            if not self.less(f, p, ck, ci):
                break
            self.K[i], self.I[i] = ck, ci
            i = c
        self.K[i], self.I[i] = f, p

    def heapify(self):
        for i in range(self.k // 2 - 1, -1, -1):
            self.sift(i, self.k, self.K[i], self.I[i])

    def offer(self, f, p):
        if self.filled < self.k:
            self.K[self.filled], self.I[self.filled] = f, p
            self.filled += 1
            if self.filled == self.k:
                self.heapify()
            return
        if self.less(f, p, self.K[0], self.I[0]):
            self.sift(0, self.k, f, p)

    def sorted_positions(self):
        if self.filled < self.k:
            self.heapify()
        for n in range(self.k - 1, 0, -1):
            lk, li = self.K[n], self.I[n]
            self.K[n], self.I[n] = self.K[0], self.I[0]
            self.sift(0, n, lk, li)
        return [p for p in self.I if p >= 0]


def _stream(rng, n):
    """double keys with exact ties and sub-float32-ulp neighbours, shuffled original indices"""
    base = np.abs(rng.normal(0, 1, n // 3 + 1)) * 10.0 ** rng.integers(-3, 4)
    keys = np.concatenate([base, base, np.nextafter(base, np.inf)])[:n]
    keys[rng.integers(0, n, n // 10)] = base[0]
    rng.shuffle(keys)
    return keys.astype(np.float64), rng.permutation(n).astype(np.int64)


def _canonical(keys, ids, k):
    order = sorted(range(len(keys)), key=lambda i: (keys[i], ids[i]))[:k]
    return order


def test_sorted_list_with_seed_network():
    rng = np.random.default_rng(1)
    for it in range(120):
        n = int(rng.integers(8, 120))
        keys, ids = _stream(rng, n)
        cloud = Cloud(keys, ids)
        fs = [ru32(d) for d in keys]
        for K in (1, 2, 4, 8, 12, 16):
            key, pos = [INF] * K, [-1] * K
            seed_first_group(key, pos, fs[:8], 0, cloud)
            for p in range(8, n):
                topkf_offer(key, pos, fs[p], p, cloud)
            assert [p for p in pos if p >= 0] == _canonical(keys, ids, K)


def test_heap_with_fill_phase():
    rng = np.random.default_rng(2)
    for it in range(120):
        n = int(rng.integers(1, 150))
        keys, ids = _stream(rng, n)
        cloud = Cloud(keys, ids)
        for k in (17, 20, 32):
            h = Heap(k, cloud)
            for p in range(n):
                h.offer(ru32(keys[p]), p)
            assert h.sorted_positions() == _canonical(keys, ids, k)
