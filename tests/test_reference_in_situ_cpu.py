"""The drop-in boundary tested IN SITU: the reference's own second/builder/voxel_builder.py, with the
one-import patch of INTEGRATION.md section 2.1 applied to its source text, builds a voxel generator from a
reference .config file -- and what comes out is this package's VoxelGenerator with the attributes the
reference's other builders read.  Needs /root/reference (this container; skipped on the GPU box, where the
reference does not exist) and no GPU (construction only; generate() itself is covered by the GPU parity tests)."""
import glob
import importlib
import os
import sys
import types

import numpy as np
import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "second", "builder")),
                                reason="the reference tree is not on this machine")

PATCH_OLD = "from spconv.utils import VoxelGenerator"
PATCH_NEW = ('import importlib\n'
             'VoxelGenerator = importlib.import_module("3dobject_detection_temporal_b200").VoxelGenerator')


def _patched_voxel_builder(pkg):
    src = open(os.path.join(REF, "second", "builder", "voxel_builder.py")).read()
    assert src.count(PATCH_OLD) == 1                      # the line INTEGRATION.md 2.1 names is there, once
    src = src.replace(PATCH_OLD, PATCH_NEW)
    # `from second.protos import voxel_generator_pb2`: the stale generated module does not import on a modern
    # protobuf; compat.protos revives its message classes from the serialized descriptor (nothing is executed)
    protos = importlib.import_module("3dobject_detection_temporal_b200.compat.protos")
    pool = protos.load_pool(REF)
    vg_pb2 = types.ModuleType("second.protos.voxel_generator_pb2")
    vg_pb2.VoxelGenerator = protos.message_class(pool, "second.protos.VoxelGenerator")
    saved = {k: sys.modules.get(k) for k in ("second", "second.protos", "second.protos.voxel_generator_pb2")}
    second = types.ModuleType("second"); second.__path__ = []
    sp = types.ModuleType("second.protos"); sp.__path__ = []
    sp.voxel_generator_pb2 = vg_pb2
    second.protos = sp
    sys.modules.update({"second": second, "second.protos": sp, "second.protos.voxel_generator_pb2": vg_pb2})
    try:
        mod = types.ModuleType("reference_voxel_builder_patched")
        exec(compile(src, "second/builder/voxel_builder.py (patched)", "exec"), mod.__dict__)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod, protos, pool


def test_reference_voxel_builder_builds_our_generator(pkg):
    from google.protobuf import text_format
    mod, protos, pool = _patched_voxel_builder(pkg)
    pipeline = protos.message_class(pool, "second.protos.TrainEvalPipelineConfig")
    configs = sorted(glob.glob(os.path.join(REF, "second", "configs", "**", "*.config"), recursive=True))
    assert configs
    built = 0
    for path in configs[:12]:
        cfg = pipeline()
        text_format.Merge(open(path).read(), cfg)               # one descriptor pool: the reference's isinstance check holds
        vcfg = cfg.model.second.voxel_generator
        gen = mod.build(vcfg)                                    # the reference's own build(), patched import
        assert type(gen) is pkg.VoxelGenerator
        vs = np.asarray(list(vcfg.voxel_size), np.float32)
        rg = np.asarray(list(vcfg.point_cloud_range), np.float32)
        assert np.array_equal(gen.voxel_size, vs) and np.array_equal(gen.point_cloud_range, rg)
        assert gen.max_num_points_per_voxel == vcfg.max_number_of_points_per_voxel
        # what second/pytorch/builder/second_builder.py:17-18 and dataset_builder.py:46-48 read
        assert gen.grid_size.dtype == np.int64
        assert np.array_equal(gen.grid_size, np.round((rg[3:] - rg[:3]) / vs).astype(np.int64))
        built += 1
    assert built > 0
    # the reference's own type check still guards the entry (voxel_builder.py:20-22)
    with pytest.raises(ValueError):
        mod.build(object())
