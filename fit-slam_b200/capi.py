"""ctypes view of include/fitslam_cuda.h (the C ABI of libfitslam_cuda.so).

This is test/bench plumbing: the product is the shared library and the C++ classes in
fit-slam_b200/cpp/.  There is no fallback: if the library is missing, import fails.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FITSLAM_LIB") or os.path.join(_HERE, "libfitslam_cuda.so")


class OccupancyParams(C.Structure):
    """fitslam_occupancy_params (include/fitslam_cuda.h)."""
    _fields_ = [("lethal_obstacle_threshold", C.c_int32), ("lethal_cost_threshold", C.c_int32), ("unknown_cost_value", C.c_int32),
                ("trinary_costmap", C.c_int32), ("track_unknown_space", C.c_int32)]


class LethalArea(C.Structure):
    """fitslam_lethal_area (include/fitslam_cuda.h): kind 0 = addNewMarkedArea circle, 1 = addNewMarkedAreaFOV triangle."""
    _fields_ = [("wx", C.c_double), ("wy", C.c_double), ("extent", C.c_double), ("yaw", C.c_double), ("kind", C.c_int32), ("reserved", C.c_int32)]


class SensorParams(C.Structure):
    """fitslam_sensor_params (include/fitslam_cuda.h)."""
    _fields_ = [("sensor_min_angle", C.c_double), ("sensor_max_angle", C.c_double), ("angular_resolution", C.c_double),
                ("sensor_max_range", C.c_double)]


class Params(C.Structure):
    _fields_ = [
        ("min_frontier_cluster_size", C.c_int32), ("max_frontier_cluster_size", C.c_int32),
        ("max_frontier_distance", C.c_double), ("lethal_threshold", C.c_int32), ("_pad0", C.c_int32),
        ("max_camera_depth", C.c_double), ("delta_theta", C.c_double), ("camera_fov", C.c_double),
        ("robot_radius", C.c_double), ("max_arrival_factor", C.c_double), ("min_arrival_factor", C.c_double),
        ("alpha", C.c_double), ("beta", C.c_double), ("vx_max", C.c_double), ("wz_max", C.c_double),
        ("num_frontiers_in_local_area", C.c_int32), ("_pad1", C.c_int32),
        ("local_frontier_search_radius", C.c_double),
        ("voxel_step", C.c_float), ("crowd_exponent", C.c_float), ("fisher_information_threshold", C.c_double),
        ("lut_min_x", C.c_float), ("lut_max_x", C.c_float), ("lut_min_y", C.c_float), ("lut_max_y", C.c_float),
        ("lut_min_z", C.c_float), ("lut_max_z", C.c_float),
        ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
        ("img_max_u", C.c_float), ("img_max_v", C.c_float), ("candidate_pose_z", C.c_double),
    ]


FRONTIER_DTYPE = np.dtype([("goal_x", "<f8"), ("goal_y", "<f8"), ("goal_mx", "<u4"), ("goal_my", "<u4"),
                           ("size", "<i4"), ("component", "<u4"), ("chunk", "<u4"), ("first_cell", "<u4")])
ARRIVAL_DTYPE = np.dtype([("information", "<i4"), ("max_index", "<i4"), ("theta_s_star", "<f8"),
                          ("achievable", "u1"), ("footprint_in_lethal", "u1"), ("ray_failed", "u1"),
                          ("_pad", "u1", (5,))])
POSE_DTYPE = np.dtype([("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("qx", "<f8"), ("qy", "<f8"), ("qz", "<f8"),
                       ("qw", "<f8")])
RECORD_DTYPE = np.dtype([("information", "<i4"), ("max_index", "<i4"), ("fisher", "<f4"), ("flags", "<u4")])
KEYFRAME_DTYPE = np.dtype([("Rwc", "<f4", (9,)), ("twc", "<f4", (3,)), ("first", "<u4"), ("count", "<u4")])  # fitslam_keyframe
assert FRONTIER_DTYPE.itemsize == 40 and ARRIVAL_DTYPE.itemsize == 24 and POSE_DTYPE.itemsize == 56 and KEYFRAME_DTYPE.itemsize == 56


class Selection(C.Structure):
    _fields_ = [("n_sorted", C.c_uint32), ("n_local", C.c_uint32), ("global_begin", C.c_uint32),
                ("n_global", C.c_uint32), ("closest_global", C.c_int32), ("_pad", C.c_int32)]


class CycleResult(C.Structure):
    _fields_ = [("n_frontiers", C.c_uint32), ("n_frontier_cells", C.c_uint32), ("n_components", C.c_uint32),
                ("n_candidates", C.c_uint32), ("n_fisher_poses", C.c_uint32), ("bfs_levels", C.c_uint32),
                ("ambiguous_sorts", C.c_uint32), ("_pad", C.c_uint32), ("selection", Selection),
                ("stage_ms", C.c_float * 8), ("widest_level", C.c_uint32), ("wavefront_ctas", C.c_uint32)]


OK, E_INVALID, E_CUDA, E_CAPACITY, E_NO_COSTMAP, E_ROBOT_OUT_OF_BOUNDS, E_NO_FRONTIERS, E_NO_POLYGON, \
    E_ARRIVAL_UTILITY, E_COST_OUT_OF_BOUNDS, E_NO_LOOKUP_TABLE, E_NO_MAP_POINTS, E_IO = range(13)
DIST_GIVEN, DIST_EUCLIDEAN = 0, 1
LIV_STATIC_MAX_POINTS = 4194304  # FITSLAM_LIV_STATIC_MAX_POINTS
FISHER_OFF, FISHER_BEST_YAW, FISHER_ALL_YAWS = 0, 1, 2

_vp, _u32p, _dp = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/fitslam_cuda.h declares
PROTOTYPES = {
    "fitslam_abi_version": (C.c_int, []),
    "fitslam_default_params": (None, [C.POINTER(Params)]),
    "fitslam_create": (C.c_int, [C.POINTER(Params), C.c_int, C.POINTER(_vp)]),
    "fitslam_destroy": (None, [_vp]),
    "fitslam_last_error": (C.c_char_p, [_vp]),
    "fitslam_status_string": (C.c_char_p, [C.c_int]),
    "fitslam_set_shard": (C.c_int, [_vp, C.c_int, C.c_int]),
    "fitslam_stream": (_vp, [_vp]),
    "fitslam_launch_count": (C.c_uint64, [_vp]),
    "fitslam_set_costmap": (C.c_int, [_vp, _vp, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double, _vp]),
    "fitslam_get_costmap": (C.c_int, [_vp, _vp, C.c_uint64]),
    "fitslam_set_occupancy_grid": (C.c_int, [_vp, _vp, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double, _vp, _vp]),
    "fitslam_default_occupancy_params": (None, [_vp]),
    "fitslam_mark_lethal": (C.c_int, [_vp, _vp, C.c_uint32, _u32p]),
    "fitslam_default_sensor_params": (None, [_vp]),
    "fitslam_sensor_mark_rays": (C.c_int, [_vp, _vp, _vp, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                           C.c_double, _vp, _u32p]),
    "fitslam_set_map_points": (C.c_int, [_vp, _vp, C.c_uint64]),
    "fitslam_search_frontiers": (C.c_int, [_vp, C.c_double, C.c_double, _vp, _u32p, _vp, _u32p]),
    "fitslam_set_search_distance": (C.c_int, [_vp, C.c_double]),
    "fitslam_roadmap_connectable": (C.c_int, [_vp, _vp, C.c_uint32, C.c_double, _vp, _vp]),
    "fitslam_arrival_limits": (C.c_int, [_vp, _dp, _dp, C.POINTER(C.c_int32)]),
    "fitslam_score_arrival": (C.c_int, [_vp, _vp, C.c_uint32, _vp, _vp, _vp, _vp]),
    "fitslam_ray_shape": (C.c_int, [_vp, _u32p, _u32p, _u32p, _u32p]),
    "fitslam_fisher_generate_lookup": (C.c_int, [_vp, C.c_char_p]),
    "fitslam_fisher_load_lookup": (C.c_int, [_vp, C.c_char_p]),
    "fitslam_fisher_lookup_dims": (C.c_int, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "fitslam_fisher_lookup_copy": (C.c_int, [_vp, _vp, C.c_uint64]),
    "fitslam_score_fisher": (C.c_int, [_vp, _vp, C.c_uint32, _vp, _vp, _vp]),
    "fitslam_landmarks_in_view": (C.c_int, [_vp, _vp, _vp, C.c_uint32, _u32p]),
    "fitslam_landmarks_in_view_static": (C.c_int, [_vp, _vp, C.c_uint64, _vp, _vp, C.c_uint32, _vp]),
    "fitslam_set_keyframes": (C.c_int, [_vp, _vp, C.c_uint32, _vp, C.c_uint64]),
    "fitslam_landmarks_in_view_keyframes": (C.c_int, [_vp, _vp, C.c_int32, C.c_float, C.c_float, C.c_int, _vp, C.c_uint32, _u32p, _u32p]),
    "fitslam_score_fisher_keyframes": (C.c_int, [_vp, _vp, C.c_int32, C.c_float, C.c_int, _vp, _vp, _u32p, _u32p]),
    "fitslam_normalise_select": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint32, _vp, _vp, _vp, _vp, _vp,
                                           C.POINTER(Selection)]),
    "fitslam_cycle": (C.c_int, [_vp, C.c_double, C.c_double, C.c_double, _vp, C.c_int, C.POINTER(CycleResult)]),
    "fitslam_cycle_begin": (C.c_int, [_vp, C.c_double, C.c_double, C.c_double, _vp, C.c_int, _u32p]),
    "fitslam_cycle_end": (C.c_int, [_vp, C.POINTER(CycleResult)]),
    "fitslam_cycle_fetch": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint32]),
    "fitslam_cycle_fetch_candidates": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint32, _vp]),
    "fitslam_comm_unique_id": (C.c_int, [_vp, C.c_uint64]),
    "fitslam_comm_init": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "fitslam_comm_destroy": (C.c_int, [_vp]),
    "fitslam_cycle_merge": (C.c_int, [_vp]),
    "fitslam_score_records": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(C.c_uint64)]),
}


def load_library(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback for the exploration hot path)")
    lib = C.CDLL(path)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


def default_occupancy_params():
    p = OccupancyParams()
    load_library().fitslam_default_occupancy_params(C.byref(p))
    return p


def default_sensor_params():
    p = SensorParams()
    load_library().fitslam_default_sensor_params(C.byref(p))
    return p


def lethal_areas(areas):
    """[(wx, wy, extent, yaw, kind), ...] -> a fitslam_lethal_area array."""
    arr = (LethalArea * max(len(areas), 1))()
    for i, (wx, wy, extent, yaw, kind) in enumerate(areas):
        arr[i] = LethalArea(wx, wy, extent, yaw, kind, 0)
    return arr


def default_params(lib=None):
    lib = lib or load_library()
    p = Params()
    lib.fitslam_default_params(C.byref(p))
    return p


class FitslamError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"fitslam status {status}: {msg}")
        self.status = status


def _prefer_bundled_nccl():
    """A Python process that also imports torch must not end up with two NCCL versions under one soname: point the library at
    the NCCL PyTorch bundles (nvidia-nccl wheel) unless the caller named one or torch has loaded it already."""
    if os.environ.get("FITSLAM_NCCL_LIB"):
        return
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except (ImportError, ValueError):
        spec = None
    for d in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
        cand = os.path.join(d, "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["FITSLAM_NCCL_LIB"] = cand
            return


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return int(a)  # raw (device) address


class Context:
    """One fitslam_ctx.  Thin: every method is one C-ABI call plus numpy buffer allocation."""

    def __init__(self, params=None, device=0, lib=None):
        self.lib = lib or load_library()
        self.params = params or default_params(self.lib)
        h = _vp()
        st = self.lib.fitslam_create(C.byref(self.params), device, C.byref(h))
        if st != OK:
            raise FitslamError(st, "fitslam_create: " + self.lib.fitslam_status_string(st).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.fitslam_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, st, allow=()):
        if st != OK and st not in allow:
            raise FitslamError(st, self.lib.fitslam_last_error(self.h).decode())
        return st

    @property
    def stream(self):
        return self.lib.fitslam_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.fitslam_launch_count(self.h))

    def set_shard(self, rank, nranks):
        self._check(self.lib.fitslam_set_shard(self.h, rank, nranks))

    # ---- one NCCL communicator per context (csrc/comm.cu)
    @staticmethod
    def comm_unique_id():
        """128 bytes (an ncclUniqueId): made on rank 0, carried to the other ranks by the host."""
        _prefer_bundled_nccl()
        buf = C.create_string_buffer(128)
        st = load_library().fitslam_comm_unique_id(buf, 128)
        if st != OK:
            raise FitslamError(st, "fitslam_comm_unique_id: libnccl.so.2 not loadable")
        return buf.raw

    def comm_init(self, unique_id, rank, nranks):
        assert len(unique_id) == 128
        _prefer_bundled_nccl()
        self._check(self.lib.fitslam_comm_init(self.h, C.c_char_p(bytes(unique_id)), rank, nranks))

    def comm_destroy(self):
        self._check(self.lib.fitslam_comm_destroy(self.h))

    def cycle_merge(self):
        self._check(self.lib.fitslam_cycle_merge(self.h))

    def set_costmap(self, data, res, ox, oy, dirty=None, shape=None):
        """data: uint8 numpy array [H, W] (host) or a raw device address with shape=(H, W); None with shape=(H, W) on the
        receiving ranks of a communicator (comm_init)."""
        if isinstance(data, np.ndarray):
            assert data.dtype == np.uint8 and data.ndim == 2
            data = np.ascontiguousarray(data)
            H, W = data.shape
        else:
            H, W = shape
        d = None if dirty is None else np.asarray(dirty, dtype=np.uint32)
        self._check(self.lib.fitslam_set_costmap(self.h, _ptr(data), W, H, res, ox, oy, _ptr(d)))
        self.W, self.H = W, H

    def set_occupancy_grid(self, data, res, ox, oy, params=None, dirty=None, shape=None):
        """data: int8 numpy array [H, W] (nav_msgs/OccupancyGrid cells, host) or a raw device address with shape=(H, W)."""
        if isinstance(data, np.ndarray):
            assert data.dtype == np.int8 and data.ndim == 2
            data = np.ascontiguousarray(data)
            H, W = data.shape
        else:
            H, W = shape
        d = None if dirty is None else np.asarray(dirty, dtype=np.uint32)
        self._check(self.lib.fitslam_set_occupancy_grid(self.h, _ptr(data), W, H, res, ox, oy,
                                                        C.byref(params) if params is not None else None, _ptr(d)))
        self.W, self.H = W, H

    def costmap_copy(self):
        """The resident costmap as a [H, W] uint8 array (Costmap2D::getCharMap)."""
        out = np.zeros((self.H, self.W), np.uint8)
        self._check(self.lib.fitslam_get_costmap(self.h, _ptr(out), out.size))
        return out

    def mark_lethal(self, areas):
        """LethalMarker: stamp 253 over the ray fans of [(wx, wy, extent, yaw, kind), ...] in the resident costmap;
        returns the number of areas whose centre was on the map."""
        arr, n = lethal_areas(areas), C.c_uint32(0)
        self._check(self.lib.fitslam_mark_lethal(self.h, C.cast(arr, _vp), len(areas), C.byref(n)))
        return n.value

    def sensor_mark_rays(self, latest, explored, res, ox, oy, px, py, yaw, params=None, shape=None):
        """SensorSimulator ray marking.  latest / explored: int8 numpy arrays [H, W] (explored is updated in place) or two
        raw device addresses with shape=(H, W).  Returns the number of rays cast."""
        if isinstance(latest, np.ndarray):
            assert latest.dtype == np.int8 and explored.dtype == np.int8 and latest.shape == explored.shape
            assert latest.flags.c_contiguous and explored.flags.c_contiguous
            H, W = latest.shape
        else:
            H, W = shape
        n = C.c_uint32(0)
        self._check(self.lib.fitslam_sensor_mark_rays(self.h, _ptr(latest), _ptr(explored), W, H, res, ox, oy, px, py, yaw,
                                                      C.byref(params) if params is not None else None, C.byref(n)))
        return n.value

    def set_map_points(self, xyz, n=None):
        if isinstance(xyz, np.ndarray):
            xyz = np.ascontiguousarray(xyz, dtype=np.float32)
            n = xyz.shape[0]
        self._check(self.lib.fitslam_set_map_points(self.h, _ptr(xyz), n))

    def ray_shape(self):
        v = [C.c_uint32() for _ in range(4)]
        self._check(self.lib.fitslam_ray_shape(self.h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)  # n_rays, k, n_windows, max_length

    def roadmap_connectable(self, segments, radius, want_unknown=True):
        """FrontierRoadMap::isConnectable for n node pairs (n x 4 doubles, world frame) on the resident costmap."""
        seg = np.ascontiguousarray(segments, np.float64).reshape(-1, 4)
        n = len(seg)
        ok = np.zeros(n, np.uint8)
        unk = np.zeros(n, np.int32) if want_unknown else None
        self._check(self.lib.fitslam_roadmap_connectable(self.h, _ptr(seg), n, radius, _ptr(ok), _ptr(unk) if want_unknown else None))
        return ok, unk

    def arrival_limits(self):
        a, b, m = C.c_double(), C.c_double(), C.c_int32()
        self._check(self.lib.fitslam_arrival_limits(self.h, C.byref(a), C.byref(b), C.byref(m)))
        return a.value, b.value, m.value

    def set_search_distance(self, d):
        self._check(self.lib.fitslam_set_search_distance(self.h, d))

    def search_frontiers(self, rx, ry, capacity=1 << 16, cell_capacity=1 << 21, want_cells=True):
        out = np.zeros(capacity, dtype=FRONTIER_DTYPE)
        n = C.c_uint32(capacity)
        cells = np.zeros(cell_capacity, dtype=np.uint32) if want_cells else None
        nc = C.c_uint32(cell_capacity if want_cells else 0)
        st = self._check(self.lib.fitslam_search_frontiers(self.h, rx, ry, _ptr(out), C.byref(n), _ptr(cells),
                                                           C.byref(nc)), allow=(E_ROBOT_OUT_OF_BOUNDS,))
        if st == E_ROBOT_OUT_OF_BOUNDS:
            return st, out[:0], (cells[:0] if want_cells else None)
        return st, out[:n.value], (cells[:nc.value] if want_cells else None)

    def score_arrival(self, frontiers, poly, want_rays=False, want_windows=False, n=None):
        """frontiers: FRONTIER_DTYPE array, or None to score the list resident from the last search."""
        if frontiers is not None:
            frontiers = np.ascontiguousarray(frontiers)
            n = len(frontiers)
        n_rays, k, n_win, _ = self.ray_shape()
        out = np.zeros(n, dtype=ARRIVAL_DTYPE)
        rays = np.zeros((n, n_rays), dtype=np.uint16) if want_rays else None
        wins = np.zeros((n, n_win), dtype=np.int32) if want_windows else None
        poly = None if poly is None else np.asarray(poly, dtype=np.float64)
        self._check(self.lib.fitslam_score_arrival(self.h, _ptr(frontiers), n, _ptr(poly), _ptr(out), _ptr(rays),
                                                   _ptr(wins)))
        return out, rays, wins

    def fisher_generate_lookup(self, path=None):
        self._check(self.lib.fitslam_fisher_generate_lookup(self.h, path.encode() if path else None))

    def fisher_load_lookup(self, path):
        self._check(self.lib.fitslam_fisher_load_lookup(self.h, path.encode()))

    def fisher_lookup(self):
        dims, org = (C.c_int32 * 3)(), (C.c_int32 * 3)()
        self._check(self.lib.fitslam_fisher_lookup_dims(self.h, dims, org))
        out = np.zeros(tuple(dims), dtype=np.float32)
        self._check(self.lib.fitslam_fisher_lookup_copy(self.h, _ptr(out), out.size))
        return out, tuple(org)

    def score_fisher(self, poses):
        poses = np.ascontiguousarray(poses)
        n = len(poses)
        info = np.zeros(n, dtype=np.float32)
        safe = np.zeros(n, dtype=np.uint8)
        nvis = np.zeros(n, dtype=np.uint32)
        self._check(self.lib.fitslam_score_fisher(self.h, _ptr(poses), n, _ptr(info), _ptr(safe), _ptr(nvis)))
        return info, safe, nvis

    def landmarks_in_view(self, pose, capacity=None):
        """GetLandmarksInView (exhaustive) for one pose (a POSE_DTYPE record): (xyz [count, 3] float32 in the pose frame, count)."""
        pose = np.ascontiguousarray(pose, dtype=POSE_DTYPE).reshape(1)
        cap = int(capacity if capacity is not None else 1 << 22)
        out = np.zeros((cap, 3), np.float32)
        n = C.c_uint32(0)
        st = self._check(self.lib.fitslam_landmarks_in_view(self.h, _ptr(pose), _ptr(out), cap, C.byref(n)), allow=(E_CAPACITY,))
        return st, out[:min(n.value, cap)], n.value

    def set_keyframes(self, keyframes, point_index):
        """The atlas of the keyframe-structured landmark query: KEYFRAME_DTYPE records in service order + the concatenated
        map-point lists (indices into the points of set_map_points)."""
        keyframes = np.ascontiguousarray(keyframes, dtype=KEYFRAME_DTYPE)
        point_index = np.ascontiguousarray(point_index, dtype=np.uint32)
        self._check(self.lib.fitslam_set_keyframes(self.h, _ptr(keyframes), len(keyframes), _ptr(point_index), len(point_index)))

    def landmarks_in_view_keyframes(self, pose, max_landmarks=1000, max_distance=5.0, exhaustive=False, capacity=None):
        """mapPointsVisibleFromPose over the keyframe atlas: (status, xyz [count, 3] in the pose frame, count, keyframes checked)."""
        pose = np.ascontiguousarray(pose, dtype=POSE_DTYPE).reshape(1)
        cap = int(capacity if capacity is not None else 1 << 22)
        out = np.zeros((cap, 3), np.float32)
        n, chk = C.c_uint32(0), C.c_uint32(0)
        st = self._check(self.lib.fitslam_landmarks_in_view_keyframes(self.h, _ptr(pose), max_landmarks, max_distance, 0.0, int(exhaustive),
                                                                       _ptr(out), cap, C.byref(n), C.byref(chk)), allow=(E_CAPACITY,))
        return st, out[:min(n.value, cap)], n.value, chk.value

    def score_fisher_keyframes(self, pose, max_landmarks=1000, max_distance=5.0, exhaustive=False):
        """isPoseSafe behind the keyframe-structured service: (information, safe, landmarks returned, landmarks used)."""
        pose = np.ascontiguousarray(pose, dtype=POSE_DTYPE).reshape(1)
        info, safe = np.zeros(1, np.float32), np.zeros(1, np.uint8)
        n, used = C.c_uint32(0), C.c_uint32(0)
        self._check(self.lib.fitslam_score_fisher_keyframes(self.h, _ptr(pose), max_landmarks, max_distance, int(exhaustive), _ptr(info),
                                                            _ptr(safe), C.byref(n), C.byref(used)))
        return float(info[0]), int(safe[0]), n.value, used.value

    def landmarks_in_view_static(self, d_xyz, n_points, d_pose, d_out, capacity, d_count):
        """fitslam_landmarks_in_view_static: raw DEVICE addresses in and out, nothing allocated, nothing synchronised (the
        kernels run on self.stream).  Returns the status; LIV_STATIC_MAX_POINTS is the largest n_points."""
        return self.lib.fitslam_landmarks_in_view_static(self.h, _ptr(d_xyz), n_points, _ptr(d_pose), _ptr(d_out), capacity,
                                                         _ptr(d_count))

    def normalise_select(self, path_length_m, path_heading, blacklisted=None):
        n = len(path_length_m)
        pl = np.ascontiguousarray(path_length_m, dtype=np.float64)
        ph = np.ascontiguousarray(path_heading, dtype=np.float64)
        bl = None if blacklisted is None else np.ascontiguousarray(blacklisted, dtype=np.uint8)
        wc, ua, ud = (np.zeros(n, dtype=np.float64) for _ in range(3))
        ach = np.zeros(n, dtype=np.uint8)
        srt = np.zeros(n, dtype=np.uint32)
        sel = Selection()
        st = self._check(self.lib.fitslam_normalise_select(self.h, _ptr(pl), _ptr(ph), _ptr(bl), n, _ptr(wc), _ptr(ua),
                                                           _ptr(ud), _ptr(ach), _ptr(srt), C.byref(sel)),
                         allow=(E_ARRIVAL_UTILITY, E_COST_OUT_OF_BOUNDS))
        return st, wc, ua, ud, ach, srt[:sel.n_sorted], sel

    def cycle(self, rx, ry, ryaw, poly, fisher_mode=FISHER_OFF):
        res = CycleResult()
        poly = np.asarray(poly, dtype=np.float64)
        st = self._check(self.lib.fitslam_cycle(self.h, rx, ry, ryaw, _ptr(poly), fisher_mode, C.byref(res)),
                         allow=(E_ROBOT_OUT_OF_BOUNDS, E_NO_FRONTIERS))
        return st, res

    def cycle_begin(self, rx, ry, ryaw, poly, fisher_mode=FISHER_OFF):
        n = C.c_uint32()
        poly = np.asarray(poly, dtype=np.float64)
        st = self._check(self.lib.fitslam_cycle_begin(self.h, rx, ry, ryaw, _ptr(poly), fisher_mode, C.byref(n)),
                         allow=(E_ROBOT_OUT_OF_BOUNDS, E_NO_FRONTIERS))
        return st, n.value

    def cycle_end(self):
        res = CycleResult()
        self._check(self.lib.fitslam_cycle_end(self.h, C.byref(res)))
        return res

    @staticmethod
    def fetch_buffers(capacity, n_poses=0, alloc=np.zeros):
        """Host buffers for cycle_fetch / cycle_fetch_candidates that a caller keeps between cycles (`alloc(n, dtype=...)`:
        np.zeros, or an allocator of pinned memory)."""
        b = dict(frontiers=alloc(capacity, dtype=FRONTIER_DTYPE), arrival=alloc(capacity, dtype=ARRIVAL_DTYPE),
                 path_length_m=alloc(capacity, dtype=np.float64), path_heading=alloc(capacity, dtype=np.float64),
                 weighted_cost=alloc(capacity, dtype=np.float64), fisher=alloc(capacity, dtype=np.float32),
                 sorted=alloc(capacity, dtype=np.uint32))
        if n_poses:
            b.update(candidate_fisher=alloc(n_poses, dtype=np.float32), candidate_safe=alloc(n_poses, dtype=np.uint8),
                     candidate_landmarks=alloc(n_poses, dtype=np.uint32))
        return b

    def cycle_fetch(self, n, out=None):
        """Results of the last cycle for its n frontiers.  `out`: buffers from fetch_buffers() to fill (views of the first n
        entries are returned); default: fresh arrays."""
        if out is not None and len(out["frontiers"]) < n:
            raise FitslamError(E_CAPACITY, "cycle_fetch: the caller's buffers hold fewer than n frontiers")
        b = out if out is not None else self.fetch_buffers(n)
        fr, ar, pl, ph, wc, fi, srt = (b[k] for k in ("frontiers", "arrival", "path_length_m", "path_heading", "weighted_cost",
                                                       "fisher", "sorted"))
        self._check(self.lib.fitslam_cycle_fetch(self.h, _ptr(fr), _ptr(ar), _ptr(pl), _ptr(ph), _ptr(wc), _ptr(fi),
                                                 _ptr(srt), len(fr)))
        return dict(frontiers=fr[:n], arrival=ar[:n], path_length_m=pl[:n], path_heading=ph[:n], weighted_cost=wc[:n], fisher=fi[:n],
                    sorted=srt[:n])

    def cycle_fetch_candidates(self, n_poses, out=None):
        """Per-candidate Fisher results of the last cycle: (information, safe, n_visible), pose = frontier * ppf + window.
        `out`: buffers from fetch_buffers(..., n_poses) to fill."""
        if out is not None and len(out["candidate_fisher"]) >= n_poses:
            info, safe, nvis = out["candidate_fisher"], out["candidate_safe"], out["candidate_landmarks"]
        else:
            info = np.zeros(n_poses, dtype=np.float32)
            safe = np.zeros(n_poses, dtype=np.uint8)
            nvis = np.zeros(n_poses, dtype=np.uint32)
        n = C.c_uint32()
        self._check(self.lib.fitslam_cycle_fetch_candidates(self.h, _ptr(info), _ptr(safe), _ptr(nvis), len(info), C.byref(n)))
        return info[:n.value], safe[:n.value], nvis[:n.value]

    def records_tensor(self):
        """Zero-copy torch view (int32, [nranks * B, 4]) of the device score records, for the NCCL all-gather."""
        import torch
        ptr, nbytes = self.score_records()

        class _Dev:
            __cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<i4", "data": (ptr, False), "version": 2}
        return torch.as_tensor(_Dev(), device="cuda").view(-1, 4)

    def score_records(self):
        p, b = _vp(), C.c_uint64()
        self._check(self.lib.fitslam_score_records(self.h, C.byref(p), C.byref(b)))
        return p.value, b.value
