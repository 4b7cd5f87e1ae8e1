"""ctypes binding of libbvx.so (C ABI: include/bvx.h).

There is no CPU fallback and no silent degradation: if the shared library has
not been built, or a call fails, an exception is raised.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BVX_LIB: an experiment build of the same sources (tools/variants.sh); the product is libbvx.so
LIB_PATH = os.environ.get("BVX_LIB") or os.path.join(_HERE, "libbvx.so")
ABI_VERSION = 5

BVX_OK = 0
BVX_ERR_BAD_ARGUMENT = 1
BVX_ERR_CUDA = 2
BVX_ERR_WORKSPACE = 3
BVX_ERR_NO_DEVICE = 4

PROFILE_SLOTS = 8
KERNEL_NAMES = ["k_insert", "k_mid", "", "", "k_overflow", "k_gather", "", ""]
OVERFLOW = {"break": 0, "continue": 1}
GRID = {"hashed": 0, "dense": 1}


class BvxConfig(ctypes.Structure):
    _fields_ = [
        ("voxel_size", ctypes.c_float * 3),
        ("point_cloud_range", ctypes.c_float * 6),
        ("max_num_points", ctypes.c_int32),
        ("max_voxels", ctypes.c_int32),
        ("overflow_mode", ctypes.c_int32),
        ("grid_kind", ctypes.c_int32),
        ("max_points", ctypes.c_int32),
        ("max_frames", ctypes.c_int32),
        ("max_point_features", ctypes.c_int32),
        ("device", ctypes.c_int32),
    ]


class BvxSweep(ctypes.Structure):
    _fields_ = [
        ("points", ctypes.c_void_p),
        ("n", ctypes.c_int32),
        ("has_transform", ctypes.c_int32),
        ("R", ctypes.c_float * 9),
        ("t", ctypes.c_float * 3),
        ("dt", ctypes.c_float),
        ("perm", ctypes.c_void_p),
    ]


class BvxCrop(ctypes.Structure):
    _fields_ = [
        ("n_planes", ctypes.c_int32),
        ("reserved", ctypes.c_int32),
        ("planes", (ctypes.c_double * 4) * 8),
    ]


_vp = ctypes.c_void_p
_i32 = ctypes.c_int32

# name -> (restype, argtypes); every symbol include/bvx.h declares
SYMBOLS = {
    "bvx_create": (ctypes.c_int, [ctypes.POINTER(BvxConfig), ctypes.POINTER(_vp)]),
    "bvx_destroy": (ctypes.c_int, [_vp]),
    "bvx_grid_size": (ctypes.c_int, [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float),
                                     ctypes.POINTER(ctypes.c_int64)]),
    "bvx_workspace_bytes": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_size_t)]),
    "bvx_generate": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "bvx_generate_sweeps": (ctypes.c_int, [_vp, ctypes.POINTER(BvxSweep), _i32, _i32, _i32, _vp, _vp, _vp,
                                           _vp, _vp]),
    "bvx_generate_batched": (ctypes.c_int, [_vp, _vp, ctypes.POINTER(_i32), _i32, _i32, _i32, _vp, _vp, _vp,
                                            _vp, _vp, _i32, _vp]),
    "bvx_generate_batched_sweeps": (ctypes.c_int, [_vp, ctypes.POINTER(BvxSweep), ctypes.POINTER(_i32), _i32, _i32,
                                                   _i32, ctypes.POINTER(BvxCrop), _i32, _vp, _vp, _vp, _vp, _vp,
                                                   _i32, _vp]),
    "bvx_generate_host": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, ctypes.POINTER(_i32),
                                         _vp]),
    "bvx_generate_batched_host": (ctypes.c_int, [_vp, _vp, ctypes.POINTER(_i32), _i32, _i32, _i32, _vp, _vp,
                                                 _vp, _vp, _vp]),
    "bvx_host_alloc": (ctypes.c_int, [ctypes.POINTER(_vp), ctypes.c_size_t]),
    "bvx_host_free": (ctypes.c_int, [_vp]),
    "bvx_set_prefix_upload": (ctypes.c_int, [_vp, _i32]),
    "bvx_set_host_assemble": (ctypes.c_int, [_vp, _i32]),
    "bvx_set_host_threads": (ctypes.c_int, [_vp, _i32]),
    "bvx_host_assemble_rows": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _i32]),
    "bvx_host_assemble_compact": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _i32]),
    "bvx_last_transfer": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64),
                                         ctypes.POINTER(_i32)]),
    "bvx_frame_stats": (ctypes.c_int, [_vp, _i32, ctypes.POINTER(_i32), ctypes.POINTER(_i32), ctypes.POINTER(_i32),
                                       ctypes.POINTER(_i32), _vp]),
    "bvx_mean_vfe": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "bvx_vfe_augment": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp]),
    "bvx_masked_max": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "bvx_scatter_index_bev": (ctypes.c_int, [_vp, _i32, _i32, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "bvx_scatter_bev": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "bvx_scatter_bev_static": (ctypes.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "bvx_scatter_bev_dev": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "bvx_anchors_area_workspace_bytes": (ctypes.c_size_t, [_i32, _i32, _i32, _i32, _i32]),
    "bvx_anchors_area": (ctypes.c_int, [_vp, _i32, _i32, _i32, ctypes.POINTER(ctypes.c_float),
                                        ctypes.POINTER(ctypes.c_float), ctypes.POINTER(_i32), _i32, _vp, _i32,
                                        _vp, ctypes.c_size_t, _vp, _vp]),
    "bvx_sparse_sum_for_anchors_mask": (ctypes.c_int, [_vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "bvx_fused_get_anchors_area": (ctypes.c_int, [_vp, _i32, _i32, _vp, _i32, ctypes.POINTER(ctypes.c_float),
                                                  ctypes.POINTER(ctypes.c_float), ctypes.POINTER(_i32), _vp,
                                                  _vp]),
    "bvx_profile_enable": (ctypes.c_int, [_vp, _i32]),
    "bvx_profile_read": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64),
                                        _i32]),
    "bvx_last_error": (ctypes.c_char_p, [_vp]),
    "bvx_launch_count": (ctypes.c_int64, []),
    "bvx_abi_version": (ctypes.c_int, []),
}

_LIB = None


def load():
    """Load libbvx.so; raises if it has not been built (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()' or make -C "
            f"{os.path.join(_HERE, 'csrc')}). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    if lib.bvx_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libbvx.so ABI {lib.bvx_abi_version()} != expected {ABI_VERSION}; rebuild")
    _LIB = lib
    return lib


def check(rc, handle=None):
    """Map a bvx_status to the reference's error convention (Python exceptions:
    ValueError for bad input, second/builder/voxel_builder.py:20-22)."""
    if rc == BVX_OK:
        return
    msg = load().bvx_last_error(handle)
    msg = msg.decode() if msg else f"bvx status {rc}"
    if rc == BVX_ERR_BAD_ARGUMENT:
        raise ValueError(msg)
    if rc == BVX_ERR_WORKSPACE:
        raise WorkspaceError(msg)
    raise RuntimeError(msg)


class WorkspaceError(RuntimeError):
    """Handle capacity exceeded (the Python wrapper grows the handle and retries)."""


def launch_count():
    return int(load().bvx_launch_count())


def build(verbose=False):
    """Compile libbvx.so in-tree for sm_100a (csrc/Makefile)."""
    import subprocess
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc")], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libbvx.so failed")
    return LIB_PATH
