"""``VoxelGenerator`` -- drop-in for ``spconv.utils.VoxelGenerator`` as the reference
uses it (constructed at second/builder/voxel_builder.py:23-27; ``generate`` called
inside the reference's ``prep_pointcloud``, wired at
second/builder/dataset_builder.py:51-58 and second/core/inference.py:43-48).

Same constructor keywords, same attributes (``voxel_size`` f32[3],
``point_cloud_range`` f32[6], ``grid_size`` int64[3] xyz,
``max_num_points_per_voxel``), same return tuple ``(voxels[M,T,F] f32,
coordinates[M,3] i32 zyx, num_points_per_voxel[M] i32)``.  The work runs in
hand-written sm_100a kernels behind the C ABI of include/bvx.h.

* numpy in  -> numpy out (host form: H2D, kernels, D2H of the M result rows)
* CUDA torch tensor in -> CUDA torch tensors out (no copies, caller's stream)

There is no CPU fallback: without a GPU ``generate`` raises.
"""
import ctypes

import numpy as np

from . import _lib


def _as_f32x(v, n, name):
    a = np.asarray(v, dtype=np.float32).reshape(-1)
    if a.shape[0] != n:
        raise ValueError(f"{name} must have {n} elements, got {a.shape[0]}")
    return a


def _is_torch(x):
    return type(x).__module__.split(".")[0] == "torch"


class VoxelGenerator:
    def __init__(self, voxel_size, point_cloud_range, max_num_points, max_voxels=20000,
                 overflow="break", grid="hashed", device=None, prefix_upload=True, host_assemble=True,
                 host_threads=None):
        if overflow not in _lib.OVERFLOW:
            raise ValueError(f"overflow must be 'break' or 'continue', got {overflow!r}")
        if grid not in _lib.GRID:
            raise ValueError(f"grid must be 'hashed' or 'dense', got {grid!r}")
        self._voxel_size = _as_f32x(voxel_size, 3, "voxel_size")
        self._point_cloud_range = _as_f32x(point_cloud_range, 6, "point_cloud_range")
        self._max_num_points = int(max_num_points)
        self._max_voxels = int(max_voxels)
        if self._max_num_points < 1 or self._max_voxels < 1:
            raise ValueError("max_num_points and max_voxels must be >= 1")
        self._overflow = overflow
        self._grid = grid
        self._device = device
        self._prefix_upload = bool(prefix_upload)
        self._host_assemble = bool(host_assemble)
        if host_threads is None:
            # several ranks on one host (torchrun): each takes its share of the cores
            import os
            ranks = int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1)
            host_threads = max(1, (os.cpu_count() or 1) // ranks) if ranks > 1 else 0
        self._host_threads = int(host_threads)
        lib = _lib.load()
        gs = (ctypes.c_int64 * 3)()
        _lib.check(lib.bvx_grid_size(
            self._voxel_size.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
            self._point_cloud_range.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), gs))
        self._grid_size = np.array(list(gs), dtype=np.int64)
        self._handle = None
        self._captured = []          # (graph, private generator) pairs made by capture()
        self._cap = dict(max_points=0, max_frames=0, max_voxels=0, max_point_features=0)

    # ---- attributes the reference reads (second_builder.py:17-18, train.py:135)
    @property
    def voxel_size(self):
        return self._voxel_size

    @property
    def point_cloud_range(self):
        return self._point_cloud_range

    @property
    def grid_size(self):
        return self._grid_size

    @property
    def max_num_points_per_voxel(self):
        return self._max_num_points

    @property
    def overflow(self):
        return self._overflow

    # ---- handle management
    def _device_index(self):
        if self._device is None:
            return 0
        if isinstance(self._device, int):
            return self._device
        idx = getattr(self._device, "index", None)
        return 0 if idx is None else int(idx)

    def _ensure(self, n_points, n_frames, max_voxels, n_feat, device_index=None):
        c = self._cap
        dev = self._device_index() if device_index is None else device_index
        if (self._handle is not None and n_points <= c["max_points"] and n_frames <= c["max_frames"]
                and max_voxels <= c["max_voxels"] and n_feat <= c["max_point_features"]
                and dev == c.get("device")):
            return self._handle
        self.close()
        need = dict(max_points=max(n_points, 1), max_frames=max(n_frames, 1),
                    max_voxels=max(max_voxels, self._max_voxels),
                    max_point_features=max(n_feat, 4))
        cap = {}
        for k, v in need.items():
            old = c[k]
            # first allocation is exact; later growth is geometric so that frames of
            # slowly varying size do not re-create the handle every call
            cap[k] = v if old == 0 else (old if v <= old else int(v * 1.25) + 1)
        cap["max_voxels"] = max(need["max_voxels"], c["max_voxels"])
        cap["max_point_features"] = max(need["max_point_features"], c["max_point_features"])
        cap["device"] = dev
        cfg = _lib.BvxConfig()
        cfg.voxel_size[:] = self._voxel_size.tolist()
        cfg.point_cloud_range[:] = self._point_cloud_range.tolist()
        cfg.max_num_points = self._max_num_points
        cfg.max_voxels = cap["max_voxels"]
        cfg.overflow_mode = _lib.OVERFLOW[self._overflow]
        cfg.grid_kind = _lib.GRID[self._grid]
        cfg.max_points = cap["max_points"]
        cfg.max_frames = cap["max_frames"]
        cfg.max_point_features = cap["max_point_features"]
        cfg.device = dev
        h = ctypes.c_void_p()
        _lib.check(_lib.load().bvx_create(ctypes.byref(cfg), ctypes.byref(h)))
        if not self._prefix_upload:
            _lib.check(_lib.load().bvx_set_prefix_upload(h, 0), h)
        if not self._host_assemble:
            _lib.check(_lib.load().bvx_set_host_assemble(h, 0), h)
        if self._host_threads:
            _lib.check(_lib.load().bvx_set_host_threads(h, self._host_threads), h)
        self._handle = h
        self._cap = cap
        return h

    def close(self):
        if self._handle is not None:
            _lib.load().bvx_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def workspace_bytes(self):
        if self._handle is None:
            return 0
        out = ctypes.c_size_t()
        _lib.check(_lib.load().bvx_workspace_bytes(self._handle, ctypes.byref(out)), self._handle)
        return int(out.value)

    def last_transfer(self):
        """(h2d_bytes, d2h_bytes, sub_batches_redone) of the last numpy (host form) call."""
        if self._handle is None:
            return 0, 0, 0
        a, b, r = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_int32()
        _lib.check(_lib.load().bvx_last_transfer(self._handle, ctypes.byref(a), ctypes.byref(b), ctypes.byref(r)),
                   self._handle)
        return int(a.value), int(b.value), int(r.value)

    def frame_stats(self, b):
        """Per-frame facts of the last call (frames 0 .. b-1): dict of int32 arrays ``cells`` (distinct in-range
        cells among the points looked at), ``voxels`` (M), ``cutoff`` (break mode: the point the reference's loop
        stops at) and ``processed`` (points that went through the insert kernel: a prefix of the frame in break
        mode).  Synchronises."""
        if self._handle is None:
            raise RuntimeError("frame_stats() needs a live handle: call generate once first")
        arrs = [np.zeros(b, np.int32) for _ in range(4)]
        ptrs = [a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) for a in arrs]
        _lib.check(_lib.load().bvx_frame_stats(self._handle, b, *ptrs, None), self._handle)
        return dict(zip(("cells", "voxels", "cutoff", "processed"), arrs))

    def profile(self, on=True):
        """Per-kernel device timing (CUDA events around every launch) on/off."""
        if self._handle is None:
            raise RuntimeError("profile() needs a live handle: call generate once first")
        _lib.check(_lib.load().bvx_profile_enable(self._handle, 1 if on else 0), self._handle)

    def profile_read(self, reset=True):
        """{kernel: (total_ms, launches)} accumulated since the last reset."""
        ms = (ctypes.c_double * _lib.PROFILE_SLOTS)()
        n = (ctypes.c_int64 * _lib.PROFILE_SLOTS)()
        _lib.check(_lib.load().bvx_profile_read(self._handle, ms, n, 1 if reset else 0), self._handle)
        return {_lib.KERNEL_NAMES[i]: (float(ms[i]), int(n[i])) for i in range(_lib.PROFILE_SLOTS)
                if _lib.KERNEL_NAMES[i] and n[i] > 0}

    # ---- VoxelGenerator.generate
    def generate(self, points, max_voxels=None):
        """(voxels, coordinates, num_points_per_voxel) of one frame."""
        max_voxels = int(max_voxels or self._max_voxels)
        if _is_torch(points):
            return self._generate_device(points, max_voxels)
        pts = np.ascontiguousarray(points, dtype=np.float32)
        if pts.ndim != 2 or pts.shape[1] < 3:
            raise ValueError(f"points must be [N, F>=3], got {pts.shape}")
        n, f = pts.shape
        h = self._ensure(n, 1, max_voxels, f)
        t = self._max_num_points
        voxels = np.empty((max_voxels, t, f), dtype=np.float32)
        coors = np.empty((max_voxels, 3), dtype=np.int32)
        num = np.empty((max_voxels,), dtype=np.int32)
        m = ctypes.c_int32(0)
        _lib.check(_lib.load().bvx_generate_host(
            h, pts.ctypes.data, n, f, max_voxels, voxels.ctypes.data, coors.ctypes.data,
            num.ctypes.data, ctypes.byref(m), None), h)
        return voxels[:m.value], coors[:m.value], num[:m.value]

    def _generate_device(self, points, max_voxels, padded=False):
        import torch
        if not points.is_cuda:
            raise ValueError("torch points must live on a CUDA device (numpy arrays take the host form)")
        if points.dtype != torch.float32 or points.dim() != 2 or points.shape[1] < 3:
            raise ValueError("points must be a float32 [N, F>=3] tensor")
        points = points.contiguous()
        n, f = points.shape
        dev = points.device
        h = self._ensure(n, 1, max_voxels, f, dev.index)
        t = self._max_num_points
        with torch.cuda.device(dev):
            voxels = torch.empty((max_voxels, t, f), dtype=torch.float32, device=dev)
            coors = torch.empty((max_voxels, 3), dtype=torch.int32, device=dev)
            num = torch.empty((max_voxels,), dtype=torch.int32, device=dev)
            cnt = torch.empty((1,), dtype=torch.int32, device=dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(_lib.load().bvx_generate(
                h, points.data_ptr(), n, f, max_voxels, voxels.data_ptr(), coors.data_ptr(),
                num.data_ptr(), cnt.data_ptr(), st), h)
        if padded:
            return voxels, coors, num, cnt
        m = int(cnt.item())
        return voxels[:m], coors[:m], num[:m]

    def capture(self, points, max_voxels=None):
        """CUDA-graph the padded device form on a static ``points`` buffer: returns ``(graph, outputs)``
        with ``outputs = (voxels, coors, num_points, voxel_num[1])`` at capacity.  Copy a new frame of the
        same size into ``points`` and ``graph.replay()``: the three kernels (with their programmatic
        dependent-launch edges) run without any Python or launch overhead.

        The graph gets a generator (handle, workspace, descriptors) of its own, kept alive by this one:
        later calls on this generator -- other frame sizes, handle growth -- cannot disturb a replay."""
        import torch
        own = VoxelGenerator(self._voxel_size, self._point_cloud_range, self._max_num_points, self._max_voxels,
                             overflow=self._overflow, grid=self._grid, device=self._device,
                             prefix_upload=self._prefix_upload, host_assemble=self._host_assemble,
                             host_threads=self._host_threads)
        own.generate_padded(points, max_voxels)          # outside capture: allocates the handle, warms the kernels
        torch.cuda.synchronize(points.device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = own.generate_padded(points, max_voxels)
        self._captured.append((graph, own))
        return graph, out

    def generate_padded(self, points, max_voxels=None):
        """Device form without the host sync: padded outputs + device count [1]."""
        return self._generate_device(points, int(max_voxels or self._max_voxels), padded=True)

    # ---- temporal input (north star)
    def generate_sweeps(self, sweeps, max_voxels=None):
        """``sweeps``: list of (points[N_s, F_in], R[3,3] or None, t[3] or None, dt).
        Key frame first.  The merged cloud (xyz transformed, dt appended) is never
        materialised; F_out = F_in + 1.  numpy or CUDA tensors (all the same kind)."""
        max_voxels = int(max_voxels or self._max_voxels)
        if len(sweeps) < 1:
            raise ValueError("need at least one sweep")
        use_torch = _is_torch(sweeps[0][0])
        import torch
        if use_torch:
            dev = sweeps[0][0].device
        else:
            dev = torch.device("cuda", self._device_index())
        keep, arr = [], (_lib.BvxSweep * len(sweeps))()
        f_in, total = self._sweep_structs(sweeps, arr, 0, dev, use_torch, keep, None)
        f_out = f_in + 1
        h = self._ensure(total, 1, max_voxels, f_out, dev.index)
        tt = self._max_num_points
        with torch.cuda.device(dev):
            voxels = torch.empty((max_voxels, tt, f_out), dtype=torch.float32, device=dev)
            coors = torch.empty((max_voxels, 3), dtype=torch.int32, device=dev)
            num = torch.empty((max_voxels,), dtype=torch.int32, device=dev)
            cnt = torch.empty((1,), dtype=torch.int32, device=dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(_lib.load().bvx_generate_sweeps(
                h, arr, len(sweeps), f_in, max_voxels, voxels.data_ptr(), coors.data_ptr(),
                num.data_ptr(), cnt.data_ptr(), st), h)
            m = int(cnt.item())
        out = voxels[:m], coors[:m], num[:m]
        if not use_torch:
            out = tuple(o.cpu().numpy() for o in out)
        return out

    def _sweep_structs(self, sweeps, arr, base, dev, use_torch, keep, f_in):
        """Fill arr[base:base+len(sweeps)]; returns (f_in, points in these sweeps)."""
        import torch
        total = 0
        for i, sw in enumerate(sweeps):
            pts, R, t, dt = sw[:4]
            perm = sw[4] if len(sw) > 4 else None
            if use_torch:
                p = pts.contiguous()
                if p.dtype != torch.float32 or not p.is_cuda:
                    raise ValueError("sweep points must be float32 CUDA tensors")
            else:
                p = torch.from_numpy(np.ascontiguousarray(pts, dtype=np.float32)).to(dev)
            if p.dim() != 2 or p.shape[1] < 3:
                raise ValueError("sweep points must be [N, F>=3]")
            if f_in is None:
                f_in = int(p.shape[1])
            elif f_in != p.shape[1]:
                raise ValueError("all sweeps need the same feature count")
            keep.append(p)
            a = arr[base + i]
            a.points = p.data_ptr()
            a.n = int(p.shape[0])
            a.has_transform = 0 if (R is None and t is None) else 1
            Rm = np.eye(3, dtype=np.float32) if R is None else np.asarray(R, dtype=np.float32).reshape(3, 3)
            tv = np.zeros(3, dtype=np.float32) if t is None else np.asarray(t, dtype=np.float32).reshape(3)
            a.R[:] = Rm.reshape(-1).tolist()
            a.t[:] = tv.tolist()
            a.dt = float(np.float32(dt))
            a.perm = None
            if perm is not None:
                # the reference's shuffle_points as a gather on the load: row j = points[perm[j]]
                q = perm if _is_torch(perm) else torch.from_numpy(np.ascontiguousarray(perm))
                q = q.to(device=p.device, dtype=torch.int32).contiguous()
                if q.dim() != 1 or q.shape[0] != p.shape[0]:
                    raise ValueError("perm must be a permutation of the sweep's rows")
                keep.append(q)
                a.perm = q.data_ptr()
            total += int(p.shape[0])
        return f_in, total

    def generate_batched_sweeps(self, frames, max_voxels=None, append_dt=True, mean_channels=0, crops=None,
                                padded=False, out=None):
        """B frames, each a list of sweeps ``(points[N_s, F_in], R or None, t or None, dt[, perm])``, key frame
        first: the batched form of ``generate_sweeps``.  ``append_dt=False`` keeps F_out = F_in; with
        one transformed sweep per frame that is the reference's global rotation / scaling / flip
        (second/core/preprocess.py:756-804) applied in registers while the points are loaded.
        ``crops``: optional list, one entry per frame, ``None`` or ``(normal_vec [K, 3], d [K])`` with
        K <= 8 -- the surfaces of a convex polygon as the reference's ``surface_equ_3d`` returns them
        (second/core/geometry.py:84-93); points with ``x*a + y*b + z*c + d >= 0`` for any surface are
        dropped before the transform, which is ``remove_outside_points`` (box_np_ops.py:645-657).
        Returns ``(voxels [sum M, T, F_out], coors [sum M, 4], num_points, frame_voxel_offsets
        [B + 1][, mean])`` -- CUDA tensors for CUDA input, numpy arrays for numpy input.  ``padded=True`` (CUDA
        input) skips the host sync and returns the capacity-sized buffers; ``out=`` reuses a previous such return."""
        import torch
        max_voxels = int(max_voxels or self._max_voxels)
        b = len(frames)
        if b < 1 or any(len(fr) < 1 for fr in frames):
            raise ValueError("need at least one frame and one sweep per frame")
        use_torch = _is_torch(frames[0][0][0])
        dev = frames[0][0][0].device if use_torch else torch.device("cuda", self._device_index())
        nsw = sum(len(fr) for fr in frames)
        arr = (_lib.BvxSweep * nsw)()
        so = np.zeros(b + 1, dtype=np.int32)
        keep, f_in, total, base = [], None, 0, 0
        for i, fr in enumerate(frames):
            f_in, n = self._sweep_structs(fr, arr, base, dev, use_torch, keep, f_in)
            total += n
            base += len(fr)
            so[i + 1] = base
        carr = None
        if crops is not None:
            if len(crops) != b:
                raise ValueError("crops needs one entry per frame")
            carr = (_lib.BvxCrop * b)()
            for i, cr in enumerate(crops):
                if cr is None:
                    continue
                nv = np.asarray(cr[0], dtype=np.float64).reshape(-1, 3)
                dd = np.asarray(cr[1], dtype=np.float64).reshape(-1)
                if nv.shape[0] != dd.shape[0] or nv.shape[0] > 8:
                    raise ValueError("a crop is (normal_vec [K, 3], d [K]) with K <= 8")
                carr[i].n_planes = nv.shape[0]
                for k in range(nv.shape[0]):
                    carr[i].planes[k][:] = [nv[k, 0], nv[k, 1], nv[k, 2], dd[k]]
        f_out = f_in + (1 if append_dt else 0)
        h = self._ensure(total, max(b, nsw), max_voxels, f_out, dev.index)
        tt, cap = self._max_num_points, b * max_voxels
        with torch.cuda.device(dev):
            if out is not None:
                voxels, coors, num, fvo = out[:4]
                mean = out[4] if mean_channels > 0 else None
                if (tuple(voxels.shape) != (cap, tt, f_out) or tuple(coors.shape) != (cap, 4) or num.shape[0] != cap
                        or fvo.shape[0] != b + 1):
                    raise ValueError("out= buffers do not match this call's capacity shapes")
            else:
                voxels = torch.empty((cap, tt, f_out), dtype=torch.float32, device=dev)
                coors = torch.empty((cap, 4), dtype=torch.int32, device=dev)
                num = torch.empty((cap,), dtype=torch.int32, device=dev)
                fvo = torch.empty((b + 1,), dtype=torch.int32, device=dev)
                mean = (torch.empty((cap, mean_channels), dtype=torch.float32, device=dev)
                        if mean_channels > 0 else None)
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(_lib.load().bvx_generate_batched_sweeps(
                h, arr, so.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), b, f_in, 1 if append_dt else 0,
                carr, max_voxels, voxels.data_ptr(), coors.data_ptr(), num.data_ptr(), fvo.data_ptr(),
                mean.data_ptr() if mean is not None else None, mean_channels, st), h)
            if padded and use_torch:
                self._keep = keep                     # the sweep tensors must outlive the enqueued kernels
                return (voxels, coors, num, fvo) + ((mean,) if mean is not None else ())
            m = int(fvo[-1].item())
        out = (voxels[:m], coors[:m], num[:m], fvo) + ((mean[:m],) if mean is not None else ())
        if not use_torch:
            out = tuple(o.cpu().numpy() for o in out)
        return out

    # ---- B frames per call (+ the voxel part of merge_second_batch)
    def generate_bev(self, points, frame_offsets, max_voxels=None, channels=4, out=None):
        """Fused pillar path, points -> BEV canvas, device tensors only and no host sync: voxelization
        with the mean VFE fused into the gather (the padded ``[M, T, F]`` rows are never written) and the
        PointPillars scatter of ``middle.py:513-526`` reading the voxel count from the device.  Equals
        ``PointPillarsScatter(...)(VoxelFeatureExtractorV3()(voxels, num), coors, B)`` on the outputs of
        ``generate_batched``.  Returns ``(canvas [B, C*D, H, W], voxel_features [cap, C], coors [cap, 4],
        num_points [cap], frame_voxel_offsets [B+1])`` at capacity; ``out=`` reuses a previous return."""
        import torch
        max_voxels = int(max_voxels or self._max_voxels)
        if not _is_torch(points) or not points.is_cuda or points.dtype != torch.float32 or points.dim() != 2:
            raise ValueError("points must be a float32 CUDA [N, F] tensor")
        points = points.contiguous()
        n, f = points.shape
        offs = np.ascontiguousarray(frame_offsets, dtype=np.int32)
        b = offs.shape[0] - 1
        if b < 1 or offs[0] != 0 or int(offs[-1]) != n:
            raise ValueError("frame_offsets must be [B+1], start at 0 and end at the number of points")
        c = int(channels)
        if not 1 <= c <= f:
            raise ValueError(f"channels must be in [1, {f}]")
        dev = points.device
        h = self._ensure(n, b, max_voxels, f, dev.index)
        cap = b * max_voxels
        gx, gy, gz = (int(x) for x in self._grid_size)
        lib = _lib.load()
        with torch.cuda.device(dev):
            if out is not None:
                canvas, feats, coors, num, fvo = out
                if (tuple(canvas.shape) != (b, c * gz, gy, gx) or tuple(feats.shape) != (cap, c)
                        or tuple(coors.shape) != (cap, 4) or num.shape[0] != cap or fvo.shape[0] != b + 1):
                    raise ValueError("out= buffers do not match this call's shapes")
            else:
                canvas = torch.empty((b, c * gz, gy, gx), dtype=torch.float32, device=dev)
                feats = torch.empty((cap, c), dtype=torch.float32, device=dev)
                coors = torch.empty((cap, 4), dtype=torch.int32, device=dev)
                num = torch.empty((cap,), dtype=torch.int32, device=dev)
                fvo = torch.empty((b + 1,), dtype=torch.int32, device=dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.bvx_generate_batched(
                h, points.data_ptr(), offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), b, f, max_voxels,
                None, coors.data_ptr(), num.data_ptr(), fvo.data_ptr(), feats.data_ptr(), c, st), h)
            _lib.check(lib.bvx_scatter_bev_dev(
                feats.data_ptr(), coors.data_ptr(), fvo.data_ptr() + 4 * b, cap, c, b, gz, gy, gx,
                canvas.data_ptr(), st))
        return canvas, feats, coors, num, fvo

    def generate_batched(self, points, frame_offsets=None, max_voxels=None, mean_channels=0,
                         padded=False, out=None):
        """Voxelize B frames in one call.

        ``points``: list of per-frame arrays, or one [sum N, F] array/tensor with
        ``frame_offsets`` [B+1].  Returns ``(voxels[sumM,T,F], coordinates[sumM,4]
        (b,z,y,x), num_points[sumM], frame_voxel_offsets[B+1])`` (+ mean features
        [sumM, mean_channels] when ``mean_channels > 0``: the mean VFE fused into the
        gather kernel).  ``padded=True`` (device form) skips the host sync and returns
        capacity-sized buffers.  ``out`` = previously returned capacity-sized buffers
        ``(voxels, coors, num, offsets[, mean])`` to write into instead of allocating
        (CUDA tensors for the device form, ideally pinned numpy arrays for the host form).
        """
        max_voxels = int(max_voxels or self._max_voxels)
        if isinstance(points, (list, tuple)):
            if len(points) == 0:
                raise ValueError("need at least one frame")
            sizes = [int(p.shape[0]) for p in points]
            frame_offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
            if _is_torch(points[0]):
                import torch
                points = torch.cat([p.contiguous() for p in points], dim=0)
            else:
                points = np.concatenate([np.asarray(p, dtype=np.float32) for p in points], axis=0)
        if frame_offsets is None:
            raise ValueError("frame_offsets required with a concatenated points array")
        offs = np.ascontiguousarray(frame_offsets, dtype=np.int32)
        b = offs.shape[0] - 1
        if b < 1 or offs[0] != 0:
            raise ValueError("frame_offsets must be [B+1] starting at 0")
        offs_p = offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        t = self._max_num_points
        lib = _lib.load()
        if _is_torch(points):
            import torch
            if not points.is_cuda or points.dtype != torch.float32 or points.dim() != 2:
                raise ValueError("points must be a float32 CUDA [N, F] tensor")
            points = points.contiguous()
            n, f = points.shape
            if int(offs[-1]) != n:
                raise ValueError("frame_offsets[-1] must equal the number of points")
            dev = points.device
            h = self._ensure(n, b, max_voxels, f, dev.index)
            cap = b * max_voxels
            with torch.cuda.device(dev):
                if out is not None:
                    voxels, coors, num, fvo = out[:4]
                    mean = out[4] if mean_channels > 0 else None
                    if (tuple(voxels.shape) != (cap, t, f) or tuple(coors.shape) != (cap, 4)
                            or num.shape[0] != cap or fvo.shape[0] != b + 1
                            or (mean is not None and tuple(mean.shape) != (cap, mean_channels))):
                        raise ValueError("out= buffers do not match this call's capacity shapes")
                else:
                    voxels = torch.empty((cap, t, f), dtype=torch.float32, device=dev)
                    coors = torch.empty((cap, 4), dtype=torch.int32, device=dev)
                    num = torch.empty((cap,), dtype=torch.int32, device=dev)
                    fvo = torch.empty((b + 1,), dtype=torch.int32, device=dev)
                    mean = (torch.empty((cap, mean_channels), dtype=torch.float32, device=dev)
                            if mean_channels > 0 else None)
                st = torch.cuda.current_stream(dev).cuda_stream
                _lib.check(lib.bvx_generate_batched(
                    h, points.data_ptr(), offs_p, b, f, max_voxels, voxels.data_ptr(), coors.data_ptr(),
                    num.data_ptr(), fvo.data_ptr(), mean.data_ptr() if mean is not None else None,
                    mean_channels, st), h)
            if padded:
                return (voxels, coors, num, fvo) + ((mean,) if mean is not None else ())
            m = int(fvo[-1].item())
            return (voxels[:m], coors[:m], num[:m], fvo) + ((mean[:m],) if mean is not None else ())
        if mean_channels:
            raise ValueError("mean_channels is only available in the device form")
        pts = points if (isinstance(points, np.ndarray) and points.dtype == np.float32
                         and points.flags.c_contiguous) else np.ascontiguousarray(points, dtype=np.float32)
        n, f = pts.shape
        if int(offs[-1]) != n:
            raise ValueError("frame_offsets[-1] must equal the number of points")
        h = self._ensure(n, b, max_voxels, f)
        cap = b * max_voxels
        if out is not None:
            voxels, coors, num, fvo = out[:4]
            if (voxels.shape != (cap, t, f) or coors.shape != (cap, 4) or num.shape != (cap,)
                    or fvo.shape != (b + 1,) or voxels.dtype != np.float32 or coors.dtype != np.int32
                    or num.dtype != np.int32 or fvo.dtype != np.int32):
                raise ValueError("out= buffers do not match this call's capacity shapes/dtypes")
        else:
            voxels = np.empty((cap, t, f), dtype=np.float32)
            coors = np.empty((cap, 4), dtype=np.int32)
            num = np.empty((cap,), dtype=np.int32)
            fvo = np.empty((b + 1,), dtype=np.int32)
        _lib.check(lib.bvx_generate_batched_host(
            h, pts.ctypes.data, offs_p, b, f, max_voxels, voxels.ctypes.data, coors.ctypes.data,
            num.ctypes.data, fvo.ctypes.data, None), h)
        m = int(fvo[-1])
        return voxels[:m], coors[:m], num[:m], fvo


class _PinnedBuffer:
    """Owner of one bvx_host_alloc block; numpy arrays made from it keep it alive (``arr.base`` chain) and
    the block goes back to the driver when the last of them is collected."""

    def __init__(self, nbytes):
        p = ctypes.c_void_p()
        _lib.check(_lib.load().bvx_host_alloc(ctypes.byref(p), max(nbytes, 16)))
        self.ptr = p
        self.nbytes = max(nbytes, 16)
        self.__array_interface__ = {"shape": (self.nbytes,), "typestr": "|u1", "data": (p.value, False), "version": 3}

    def __del__(self):
        try:
            if self.ptr is not None:
                _lib.load().bvx_host_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


def pinned_empty(shape, dtype=np.float32):
    """numpy array backed by pinned host memory (bvx_host_alloc); the host forms DMA straight from/to it.
    The memory is released (bvx_host_free) when the array and every view of it are gone."""
    dtype = np.dtype(dtype)
    count = int(np.prod(shape))
    owner = _PinnedBuffer(count * dtype.itemsize)
    raw = np.asarray(owner)                               # uint8 view; raw.base is the owner
    return raw[:count * dtype.itemsize].view(dtype).reshape(shape)


def collate_points(frames, pinned=True):
    """Raw-points collate for the batched forms: per-sample ``[n_i, F]`` arrays ->
    ``(points [sum n_i, F] float32, frame_offsets [B + 1] int32)``.  This is the point-cloud part of
    what the reference's ``merge_second_batch`` does for voxels (concatenate in batch order;
    second/pytorch/train.py:17,265) moved in front of the voxelizer; with ``pinned=True`` the batch
    lands in pinned host memory, which ``generate_batched`` DMAs from in place (pipelined host form)."""
    if len(frames) < 1:
        raise ValueError("need at least one frame")
    frames = [np.asarray(f, dtype=np.float32) for f in frames]
    f = frames[0].shape[1]
    if any(x.ndim != 2 or x.shape[1] != f for x in frames):
        raise ValueError("frames must be [n_i, F] arrays with one feature count")
    offs = np.concatenate([[0], np.cumsum([x.shape[0] for x in frames])]).astype(np.int32)
    out = pinned_empty((int(offs[-1]), f), np.float32) if pinned else np.empty((int(offs[-1]), f), np.float32)
    for x, a, b in zip(frames, offs[:-1], offs[1:]):
        out[a:b] = x
    return out, offs
