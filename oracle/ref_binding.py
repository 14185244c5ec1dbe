"""ctypes view of oracle/_ref/libfitslam_ref.so: the REFERENCE's own FrontierSearch / Frontier / ray helpers, compiled
from /root/reference where it lies (oracle/Makefile).  TEST INFRASTRUCTURE ONLY: it pins the restatement in oracle.cpp
(tests/test_ref_pin.py) and gives the GPU tests a second, independent expectation.  The .so is git-ignored and travels
to the GPU box with the snapshot; nothing here reads /root/reference at run time."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libfitslam_ref.so")
LIB_PATH_RE = os.path.join(_HERE, "_ref", "libfitslam_ref_re.so")  # the roadmap-explorer package's own copies (rows 8f-2, 8f-4)
LIB_PATH_ORBW = os.path.join(_HERE, "_ref", "libfitslam_ref_orbw.so")  # the ORB-SLAM3 wrapper's landmark query (rows F1, F2, 8f-1)


def available():
    return os.path.exists(LIB_PATH)


def available_re():
    return os.path.exists(LIB_PATH_RE)


def available_orbw():
    return os.path.exists(LIB_PATH_ORBW)


class ReferenceORBW:
    """oracle/_ref/libfitslam_ref_orbw.so: ORBSLAM3Interface::mapPointsVisibleFromPose, the landmark service handler and the
    type conversions under them, compiled from the reference against stand-in ORB_SLAM3 / Sophus / Eigen declarations
    (oracle/ref_harness_orbw.cpp)."""

    def __init__(self):
        self.lib = C.CDLL(LIB_PATH_ORBW)

    def landmarks_in_view(self, xyz, keyframes, point_index, pose, params, max_landmarks=1000, max_distance=5.0, exhaustive=False):
        """(landmarks in the query-pose frame as the response's float64 positions, keyframes checked)."""
        xyz = np.ascontiguousarray(xyz, np.float32)
        kf_pose = np.ascontiguousarray(np.concatenate([keyframes["Rwc"].reshape(-1, 9), keyframes["twc"].reshape(-1, 3)], axis=1), np.float32)
        first = np.ascontiguousarray(keyframes["first"], np.uint32)
        count = np.ascontiguousarray(keyframes["count"], np.uint32)
        idx = np.ascontiguousarray(point_index, np.uint32)
        p = np.asarray(pose).reshape(-1)[0]
        pose7 = np.array([p["x"], p["y"], p["z"], p["qx"], p["qy"], p["qz"], p["qw"]], np.float64)
        cam = np.array([params.fx, params.fy, params.cx, params.cy], np.float32)
        cap = len(xyz) + 1
        out = np.zeros((cap, 3), np.float64)
        chk = C.c_int(-1)
        n = self.lib.ref_orbw_landmarks_in_view(C.c_void_p(pose7.ctypes.data), C.c_void_p(xyz.ctypes.data), C.c_uint64(len(xyz)),
                                                C.c_void_p(kf_pose.ctypes.data), C.c_void_p(first.ctypes.data), C.c_void_p(count.ctypes.data),
                                                C.c_uint32(len(first)), C.c_void_p(idx.ctypes.data), C.c_int(max_landmarks),
                                                C.c_float(max_distance), C.c_int(int(exhaustive)), C.c_void_p(cam.ctypes.data),
                                                C.c_void_p(out.ctypes.data), C.c_uint32(cap), C.byref(chk))
        return out[:n], chk.value

    def euler_ros(self, R):
        R = np.ascontiguousarray(R, np.float32)
        out = np.zeros(3, np.float32)
        self.lib.ref_orbw_euler_ros(C.c_void_p(R.ctypes.data), C.c_void_p(out.ctypes.data))
        return out


class ReferenceRE:
    """oracle/_ref/libfitslam_ref_re.so: FrontierRoadMap::isConnectable and SensorSimulator::markRay of the roadmap-explorer
    package, compiled from the reference (oracle/ref_harness_re.cpp)."""

    def __init__(self):
        self.lib = C.CDLL(LIB_PATH_RE)

    def connectable(self, a, meta, segments, radius):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        seg = np.ascontiguousarray(segments, np.float64).reshape(-1, 4)
        out = np.zeros(len(seg), np.uint8)
        self.lib.ref_re_connectable(C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_double(meta["res"]), C.c_double(meta["ox"]),
                                    C.c_double(meta["oy"]), C.c_double(radius), C.c_void_p(seg.ctypes.data), C.c_uint32(len(seg)),
                                    C.c_void_p(out.ctypes.data))
        return out

    def mark_rays(self, latest, explored, res, ox, oy, px, py, yaw, params):
        lat = np.ascontiguousarray(latest, np.int8)
        exp = np.array(explored, np.int8, copy=True, order="C")
        H, W = lat.shape
        self.lib.ref_re_mark_rays.restype = C.c_uint32
        n = self.lib.ref_re_mark_rays(C.c_void_p(lat.ctypes.data), C.c_void_p(exp.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_double(res),
                                      C.c_double(ox), C.c_double(oy), C.c_double(px), C.c_double(py), C.c_double(yaw),
                                      C.c_double(params.sensor_min_angle), C.c_double(params.sensor_max_angle),
                                      C.c_double(params.angular_resolution), C.c_double(params.sensor_max_range))
        return exp, n


class Reference:
    def __init__(self):
        self.lib = C.CDLL(LIB_PATH)

    @staticmethod
    def _map(a, meta):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        return a, (C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_double(meta["res"]), C.c_double(meta["ox"]),
                   C.c_double(meta["oy"]))

    def search_frontiers(self, a, meta, params, rx, ry, capacity=1 << 16, cell_capacity=1 << 21):
        """FrontierSearch::searchFrom + getAllFrontiers.  Returns (n or -1, frontiers[n, 3] = goal x, goal y, size;
        every_frontier[m, 2] = x, y)."""
        a, m = self._map(a, meta)
        fr = np.zeros((capacity, 3), dtype=np.float64)
        cells = np.zeros((cell_capacity, 2), dtype=np.float64)
        nc = C.c_uint32()
        n = self.lib.ref_search_frontiers(*m, C.c_double(params.min_frontier_cluster_size), C.c_double(params.max_frontier_cluster_size),
                                          C.c_double(params.max_frontier_distance), C.c_int(params.lethal_threshold), C.c_double(rx),
                                          C.c_double(ry), C.c_void_p(fr.ctypes.data), C.c_uint32(capacity), C.c_void_p(cells.ctypes.data),
                                          C.c_uint32(cell_capacity), C.byref(nc))
        if n < 0:
            return n, fr[:0], cells[:0]
        assert n <= capacity and nc.value <= cell_capacity
        return n, fr[:n], cells[:nc.value]

    def traced_cells(self, a, meta, sx, sy, wx, wy, max_length, obstacle=(240, 254), trace=(255, 255)):
        """getTracedCells with a fresh RayTracedCells.  Returns (ok, cells pushed, hit_obstacle, unknown cells, all cells)."""
        a, m = self._map(a, meta)
        out = (C.c_int32 * 4)()
        ok = self.lib.ref_traced_cells(*m, C.c_double(sx), C.c_double(sy), C.c_double(wx), C.c_double(wy), C.c_double(max_length),
                                       C.c_int(obstacle[0]), C.c_int(obstacle[1]), C.c_int(trace[0]), C.c_int(trace[1]), out)
        return bool(ok), out[0], bool(out[1]), out[2], out[3]

    def footprint_in_lethal(self, a, meta, cx, cy, radius_cells):
        a, m = self._map(a, meta)
        return bool(self.lib.ref_footprint_in_lethal(*m, C.c_uint32(cx), C.c_uint32(cy), C.c_double(radius_cells)))

    def nearest_free_cell(self, a, start, val):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        r = C.c_uint32()
        ok = self.lib.ref_nearest_free_cell(C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_uint32(start), C.c_int(val), C.byref(r))
        return bool(ok), r.value

    def nhood(self, W, H, idx, eight):
        out = (C.c_uint32 * 8)()
        n = self.lib.ref_nhood(C.c_uint32(W), C.c_uint32(H), C.c_uint32(idx), C.c_int(1 if eight else 0), out)
        return [out[i] for i in range(n)]

    def score_arrival(self, a, meta, params, goals, sizes, poly):
        """FrontierCostCalculator::setMaxArrivalInformation + setArrivalInformationForFrontier per frontier
        (DEP/src/CostCalculator.cpp:23-191, compiled from the reference).  Returns (rows [n, 6] = arrival information, achievable,
        goal orientation x y z w; limits = (return value, max_arrival_info_gt, min_arrival_info_gt))."""
        a, m = self._map(a, meta)
        g = np.ascontiguousarray(np.column_stack([np.asarray(goals, np.float64), np.asarray(sizes, np.float64)]))
        n = len(g)
        poly = np.ascontiguousarray(poly, dtype=np.float64)
        out = np.zeros((n, 6), dtype=np.float64)
        lim = np.zeros(3, dtype=np.float64)
        self.lib.ref_score_arrival(*m, C.c_double(params.max_camera_depth), C.c_double(params.delta_theta), C.c_double(params.camera_fov),
                                   C.c_double(params.robot_radius), C.c_void_p(g.ctypes.data), C.c_uint32(n), C.c_void_p(poly.ctypes.data),
                                   C.c_void_p(out.ctypes.data), C.c_void_p(lim.ctypes.data))
        return out, tuple(lim)

    def plan_euclidean(self, rx, ry, quat_xyzw, goals, achievable, arrival):
        """setPlanForFrontierEuclidean + recomputeNormalizationFactors (DEP/src/CostCalculator.cpp:445-485, 512-520).  Returns
        (rows [n, 3] = achievable, path length, path heading; (min, max plan distance, min arrival information); robot yaw as
        quatToEuler computes it)."""
        robot = np.array([rx, ry, *quat_xyzw], dtype=np.float64)
        g = np.ascontiguousarray(goals, dtype=np.float64)
        ach = np.ascontiguousarray(achievable, dtype=np.uint8)
        arr = np.ascontiguousarray(arrival, dtype=np.float64)
        n = len(g)
        out = np.zeros((n, 3), dtype=np.float64)
        norm = np.zeros(3, dtype=np.float64)
        yaw = C.c_double()
        self.lib.ref_plan_euclidean(C.c_void_p(robot.ctypes.data), C.c_void_p(g.ctypes.data), C.c_void_p(ach.ctypes.data),
                                    C.c_void_p(arr.ctypes.data), C.c_uint32(n), C.c_void_p(out.ctypes.data), C.c_void_p(norm.ctypes.data),
                                    C.byref(yaw))
        return out, tuple(norm), yaw.value

    def assign_costs(self, a, meta, params, goals, sizes, poly, rx, ry, quat_xyzw, blacklisted=None):
        """FrontierCostsManager::assignCosts({"ArrivalInformation", "EuclideanDistance"}) then
        FullPathOptimizer::getFilteredFrontiersN(n = num_frontiers_in_local_area), both compiled from the reference
        (DEP/src/FrontierCostsManager.cpp:47-213, DEP/src/FullPathOptimizer.cpp:157-227).  Returns (status, rows [n, 7] = weighted
        cost, arrival utility, distance utility, achievable, arrival information, path length, path heading; local indices, global
        indices, closest global index)."""
        a, m = self._map(a, meta)
        g = np.ascontiguousarray(np.column_stack([np.asarray(goals, np.float64), np.asarray(sizes, np.float64)]))
        n = len(g)
        poly = np.ascontiguousarray(poly, dtype=np.float64)
        robot = np.array([rx, ry, *quat_xyzw], dtype=np.float64)
        bl = None if blacklisted is None else np.ascontiguousarray(blacklisted, dtype=np.uint8)
        out = np.zeros((n, 7), dtype=np.float64)
        sel = np.zeros(3, dtype=np.int32)
        loc, glo = np.full(n + 1, -1, np.int32), np.full(n + 1, -1, np.int32)
        st = self.lib.ref_assign_costs(*m, C.c_double(params.max_camera_depth), C.c_double(params.delta_theta), C.c_double(params.camera_fov),
                                       C.c_double(params.robot_radius), C.c_double(params.alpha), C.c_double(params.beta),
                                       C.c_double(params.vx_max), C.c_double(params.wz_max), C.c_void_p(g.ctypes.data), C.c_uint32(n),
                                       C.c_void_p(poly.ctypes.data), C.c_void_p(robot.ctypes.data),
                                       C.c_void_p(bl.ctypes.data) if bl is not None else None, C.c_uint32(params.num_frontiers_in_local_area),
                                       C.c_void_p(out.ctypes.data), C.c_void_p(sel.ctypes.data), C.c_void_p(loc.ctypes.data),
                                       C.c_void_p(glo.ctypes.data))
        return st, out, loc[:sel[0]], glo[:sel[1]], int(sel[2])

    # ---- stage F: the reference's FisherInformationManager, compiled whole
    def fisher_generate(self, path, params):
        """generateLookupTable(minX .. maxZ) into `path` (DEP/src/fisher_information/FisherInfoManager.cpp:117-229)."""
        self.lib.ref_fisher_generate(C.c_char_p(path.encode()), C.c_float(params.lut_min_x), C.c_float(params.lut_max_x),
                                     C.c_float(params.lut_min_y), C.c_float(params.lut_max_y), C.c_float(params.lut_min_z),
                                     C.c_float(params.lut_max_z))

    def fisher_manager(self, path):
        """A FisherInformationManager whose constructor loaded the table at `path` (None if the file is missing)."""
        self.lib.ref_fisher_create.restype = C.c_void_p
        h = self.lib.ref_fisher_create(C.c_char_p(path.encode()))
        return C.c_void_p(h) if h else None

    def fisher_destroy(self, h):
        self.lib.ref_fisher_destroy(h)

    def fisher_pose(self, h, landmarks, threshold):
        """isPoseSafe with the landmark service answering `landmarks` ([n, 3], pose frame): (safe, information)."""
        lm = np.ascontiguousarray(landmarks, dtype=np.float64)
        info = C.c_float()
        safe = self.lib.ref_fisher_pose(h, C.c_void_p(lm.ctypes.data), C.c_uint32(len(lm)), C.c_double(threshold), C.byref(info))
        return bool(safe), info.value

    def fisher_voxel(self, x, y, z):
        out = (C.c_float * 3)()
        ok = self.lib.ref_fisher_voxel(C.c_float(x), C.c_float(y), C.c_float(z), out)
        return bool(ok), (out[0], out[1], out[2])

    def crowd_factor(self, num, s):
        self.lib.ref_crowd_factor.restype = C.c_float
        return self.lib.ref_crowd_factor(C.c_int(num), C.c_float(s))

    def information_of_point(self, x, y, z):
        self.lib.ref_information_of_point.restype = C.c_float
        return self.lib.ref_information_of_point(C.c_float(x), C.c_float(y), C.c_float(z))

    def mark_lethal(self, a, meta, areas):
        """LethalMarker (the reference's own lethal_marker.cpp) on a copy of the costmap; areas: (wx, wy, extent, yaw, kind)."""
        a, m = self._map(np.array(a, np.uint8, copy=True), meta)
        ar = np.ascontiguousarray(np.asarray(areas, np.float64).reshape(-1, 5))
        n = self.lib.ref_mark_lethal(*m, C.c_void_p(ar.ctypes.data), C.c_uint32(len(ar)))
        return a, n

    def threshold_occupancy(self, grid, lethal_threshold):
        """TraversabilityThresholdNode::callback (the reference's own text) on an int8 grid: the published data."""
        g = np.ascontiguousarray(grid, np.int8)
        out = np.zeros(g.shape, np.int8)
        self.lib.ref_threshold_occupancy(C.c_void_p(g.ctypes.data), C.c_uint64(g.size), C.c_int(lethal_threshold), C.c_void_p(out.ctypes.data))
        return out
