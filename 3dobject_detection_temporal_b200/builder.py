"""``build(voxel_config)`` -- the one-import replacement for
second/builder/voxel_builder.py (which does ``from spconv.utils import
VoxelGenerator``).  Accepts the reference's ``voxel_generator_pb2.VoxelGenerator``
message -- or any object with the same three fields, which is what a message
revived by ``compat.protos`` on a modern protobuf is."""
from .voxel_generator import VoxelGenerator

_FIELDS = ("voxel_size", "point_cloud_range", "max_number_of_points_per_voxel")


def build(voxel_config, **kwargs):
    if not all(hasattr(voxel_config, f) for f in _FIELDS):
        # same exception type and wording as second/builder/voxel_builder.py:20-22
        raise ValueError('input_reader_config not of type '
                         'input_reader_pb2.InputReader.')
    return VoxelGenerator(
        voxel_size=list(voxel_config.voxel_size),
        point_cloud_range=list(voxel_config.point_cloud_range),
        max_num_points=voxel_config.max_number_of_points_per_voxel,
        max_voxels=20000,
        **kwargs)
