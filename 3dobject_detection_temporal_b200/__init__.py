"""B200-native voxelization hot path for jhkohpav/3dobject_detection_temporal.

multi-sweep merge -> point-to-voxel -> mean VFE -> BEV scatter, as hand-written
sm_100a CUDA behind the C ABI of ``include/bvx.h``; this package is the Python
mirror of the reference interfaces for that path and nothing else.

The directory name starts with a digit, so import it with
``importlib.import_module("3dobject_detection_temporal_b200")``.
"""
from . import _lib, synth  # noqa: F401
from .builder import build as build_voxel_generator  # noqa: F401
from .voxel_generator import VoxelGenerator, collate_points, pinned_empty  # noqa: F401
from .shard import shard_frames  # noqa: F401


def __getattr__(name):
    # torch-dependent pieces are imported on first use
    if name in ("VoxelFeatureExtractorV3", "SimpleVoxel", "PointPillarsScatter", "mean_vfe",
                "scatter_bev", "VFE_CLASS_DICT", "MIDDLE_CLASS_DICT", "modules", "VFELayer",
                "VoxelFeatureExtractor", "VoxelFeatureExtractorV2", "get_paddings_indicator", "vfe_augment",
                "masked_max", "scatter_index_bev"):
        import importlib
        mod = importlib.import_module(__name__ + ".modules")
        return mod if name == "modules" else getattr(mod, name)
    if name in ("anchors_area", "anchors_mask", "sparse_sum_for_anchors_mask", "fused_get_anchors_area",
                "anchors"):
        import importlib
        mod = importlib.import_module(__name__ + ".anchors")
        return mod if name == "anchors" else getattr(mod, name)
    raise AttributeError(name)
