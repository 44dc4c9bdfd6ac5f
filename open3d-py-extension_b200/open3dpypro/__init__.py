"""open3dpypro — B200-native drop-in for the point-cloud hot path of qinhy/Open3D-py-extension.

Same import name and the same PointCloud / processors surface as the reference package; the
Open3D CPU calls underneath (transform, voxel_down_sample, estimate_normals, segment_plane) are
replaced by sm_100a CUDA kernels behind the C ABI in include/o3dpp.h.  There is no CPU fallback.
"""
__version__ = "0.1.0"
