"""B200-native nearest-neighbour search and ICP (drop-in for the Homework2 KD-tree/octree
search and the Homework9 ICP loop of yf26/Hands-On-Point-Cloud-Processing).

The directory name is not an importable identifier; import it through the repo-root shim
``pcnn_b200`` (``import pcnn_b200 as pc``).
"""
from .api import (KDTREE, OCTREE, MAX_K, ICP_NSUMS, PcnnError, device_count, lib, EXPORTS)  # noqa: F401
from .index import Index, IcpState  # noqa: F401
from .compat import (DistIndex, KNNResultSet, RadiusNNResultSet, kdtree_construction,  # noqa: F401
                     kdtree_knn_search, kdtree_radius_search, octree_construction,
                     octree_knn_search, octree_radius_search, knn_search_batch,
                     radius_search_batch, read_velodyne_bin, read_xyz_normal_bin, voxel_downsample,
                     read_binary_and_voxel_down, desc_nn1, match_descriptors, ransac)
from .sharding import shard_range, allreduce_icp_sums  # noqa: F401
from . import synth  # noqa: F401
