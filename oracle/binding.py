"""ctypes view of oracle/liboracle.so.  TEST INFRASTRUCTURE ONLY (parity: see the oracle.cpp header and oracle/ref_binding.py):
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the product."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B" if force else "-s"])
    return LIB_PATH


class Oracle:
    def __init__(self, capi, params=None):
        """capi: the fitslam_b200.capi module (struct layouts are shared through include/fitslam_cuda.h)."""
        self.capi = capi
        self.lib = C.CDLL(build())
        self.lib.orc_fisher_create.restype = C.c_void_p
        self.lib.orc_generate_uid.restype = C.c_uint64
        self.lib.orc_fisher_num_records.restype = C.c_uint64
        self.lib.orc_fisher_num_entries.restype = C.c_uint64
        for f in ("orc_fisher_voxel_information", "orc_information_of_point_local", "orc_crowd_factor"):
            getattr(self.lib, f).restype = C.c_float
        if params is None:
            params = capi.Params()
            self.lib.orc_default_params(C.byref(params))
        self.params = params
        self._fisher = None

    def _map(self, a, meta):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        return a, (C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_double(meta["res"]),
                   C.c_double(meta["ox"]), C.c_double(meta["oy"]))

    def search_frontiers(self, a, meta, rx, ry, capacity=1 << 16, cell_capacity=1 << 21):
        a, m = self._map(a, meta)
        out = np.zeros(capacity, dtype=self.capi.FRONTIER_DTYPE)
        cells = np.zeros(cell_capacity, dtype=np.uint32)
        seeds = np.zeros(capacity, dtype=np.uint32)
        cent = np.zeros((capacity, 2), dtype=np.float64)
        n, nc, ns = C.c_uint32(capacity), C.c_uint32(cell_capacity), C.c_uint32(capacity)
        st = self.lib.orc_search_frontiers(*m, C.byref(self.params), C.c_double(rx), C.c_double(ry),
                                           C.c_void_p(out.ctypes.data), C.byref(n), C.c_void_p(cells.ctypes.data),
                                           C.byref(nc), C.c_void_p(seeds.ctypes.data), C.byref(ns),
                                           C.c_void_p(cent.ctypes.data))
        return st, out[:n.value], cells[:nc.value], seeds[:ns.value], cent[:n.value]

    def frontier_mask(self, a):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        mask = np.zeros((H, W), dtype=np.uint8)
        self.lib.orc_frontier_mask(C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.byref(self.params),
                                   C.c_void_p(mask.ctypes.data))
        return mask

    def nearest_free_cell(self, a, start, val):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        H, W = a.shape
        r = C.c_uint32()
        ok = self.lib.orc_nearest_free_cell(C.c_void_p(a.ctypes.data), C.c_uint32(W), C.c_uint32(H), C.c_uint32(start),
                                            C.c_int(val), C.byref(r))
        return bool(ok), r.value

    def ray_shape(self, res):
        v = [C.c_uint32() for _ in range(4)]
        self.lib.orc_ray_shape(C.byref(self.params), C.c_double(res), *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def arrival_limits(self, a, meta):
        a, m = self._map(a, meta)
        mx, mn, mv = C.c_double(), C.c_double(), C.c_int32()
        self.lib.orc_arrival_limits(*m, C.byref(self.params), C.byref(mx), C.byref(mn), C.byref(mv))
        return mx.value, mn.value, mv.value

    def score_arrival(self, a, meta, frontiers, poly):
        a, m = self._map(a, meta)
        fr = np.ascontiguousarray(frontiers)
        n = len(fr)
        n_rays, k, n_win, _ = self.ray_shape(meta["res"])
        out = np.zeros(n, dtype=self.capi.ARRIVAL_DTYPE)
        rays = np.zeros((n, n_rays), dtype=np.uint16)
        wins = np.zeros((n, n_win), dtype=np.int32)
        poly = np.asarray(poly, dtype=np.float64)
        self.lib.orc_score_arrival(*m, C.byref(self.params), C.c_void_p(fr.ctypes.data), C.c_uint32(n),
                                   C.c_void_p(poly.ctypes.data), C.c_void_p(out.ctypes.data),
                                   C.c_void_p(rays.ctypes.data), C.c_void_p(wins.ctypes.data))
        return out, rays, wins

    def occupancy_to_cost(self, grid, occ_params):
        """threshold node + StaticLayer::interpretValue on an int8 occupancy grid -> uint8 costs."""
        g = np.ascontiguousarray(grid, np.int8)
        out = np.zeros(g.shape, np.uint8)
        self.lib.orc_occupancy_to_cost(C.c_void_p(g.ctypes.data), C.c_uint64(g.size), C.byref(occ_params), C.c_void_p(out.ctypes.data))
        return out

    def mark_lethal(self, a, meta, areas):
        """LethalMarker on a copy of the costmap: (stamped map, areas that were on the map)."""
        a, m = self._map(np.array(a, np.uint8, copy=True), meta)
        arr, n = self.capi.lethal_areas(areas), C.c_uint32(0)
        self.lib.orc_mark_lethal(*m, C.cast(arr, C.c_void_p), C.c_uint32(len(areas)), C.byref(n))
        return a, n.value

    def sensor_mark_rays(self, latest, explored, res, ox, oy, px, py, yaw, params):
        """SensorSimulator ray marking: (updated copy of explored, rays cast)."""
        lat = np.ascontiguousarray(latest, np.int8)
        exp = np.array(explored, np.int8, copy=True, order="C")
        H, W = lat.shape
        n = C.c_uint32(0)
        self.lib.orc_sensor_mark_rays(C.c_void_p(lat.ctypes.data), C.c_void_p(exp.ctypes.data), C.c_uint32(W), C.c_uint32(H),
                                      C.c_double(res), C.c_double(ox), C.c_double(oy), C.c_double(px), C.c_double(py),
                                      C.c_double(yaw), C.byref(params), C.byref(n))
        return exp, n.value

    def roadmap_connectable(self, a, meta, segments, radius):
        """FrontierRoadMap::isConnectable for n node pairs (n x 4 doubles): (flags uint8, unknown counts int32)."""
        a, m = self._map(a, meta)
        seg = np.ascontiguousarray(segments, np.float64).reshape(-1, 4)
        n = len(seg)
        ok = np.zeros(n, np.uint8)
        unk = np.zeros(n, np.int32)
        self.lib.orc_roadmap_connectable(*m, C.c_void_p(seg.ctypes.data), C.c_uint32(n), C.c_double(radius),
                                         C.c_void_p(ok.ctypes.data), C.c_void_p(unk.ctypes.data))
        return ok, unk

    def plan_euclidean(self, rx, ry, ryaw, frontiers, achievable):
        fr = np.ascontiguousarray(frontiers)
        n = len(fr)
        ach = np.ascontiguousarray(achievable, dtype=np.uint8).copy()
        pl, ph = np.zeros(n), np.zeros(n)
        self.lib.orc_plan_euclidean(C.c_double(rx), C.c_double(ry), C.c_double(ryaw), C.c_void_p(fr.ctypes.data),
                                    C.c_uint32(n), C.c_void_p(ach.ctypes.data), C.c_void_p(pl.ctypes.data),
                                    C.c_void_p(ph.ctypes.data))
        return ach, pl, ph

    def normalise(self, arrival, achievable, path_length, path_heading, blacklisted, max_arrival_info_gt):
        n = len(arrival)
        ar = np.ascontiguousarray(arrival)
        ach = np.ascontiguousarray(achievable, dtype=np.uint8)
        pl = np.ascontiguousarray(path_length, dtype=np.float64)
        ph = np.ascontiguousarray(path_heading, dtype=np.float64)
        bl = None if blacklisted is None else np.ascontiguousarray(blacklisted, dtype=np.uint8)
        wc, ua, ud = np.zeros(n), np.zeros(n), np.zeros(n)
        err = C.c_int32(-1)
        st = self.lib.orc_normalise(C.byref(self.params), C.c_void_p(ar.ctypes.data), C.c_void_p(ach.ctypes.data),
                                    C.c_void_p(pl.ctypes.data), C.c_void_p(ph.ctypes.data),
                                    C.c_void_p(bl.ctypes.data) if bl is not None else None, C.c_uint32(n),
                                    C.c_double(max_arrival_info_gt), C.c_void_p(wc.ctypes.data),
                                    C.c_void_p(ua.ctypes.data), C.c_void_p(ud.ctypes.data), C.byref(err))
        return st, wc, ua, ud, err.value

    def select(self, eligible, path_length_m):
        n = len(eligible)
        el = np.ascontiguousarray(eligible, dtype=np.uint8)
        pl = np.ascontiguousarray(path_length_m, dtype=np.float64)
        srt = np.zeros(max(n, 1), dtype=np.uint32)
        sel = self.capi.Selection()
        self.lib.orc_select(C.byref(self.params), C.c_void_p(el.ctypes.data), C.c_void_p(pl.ctypes.data), C.c_uint32(n),
                            C.c_void_p(srt.ctypes.data), C.byref(sel))
        return srt[:sel.n_sorted], sel

    def generate_uid(self, x, y):
        return int(self.lib.orc_generate_uid(C.c_double(x), C.c_double(y)))

    # ---- Fisher
    @property
    def fisher(self):
        if self._fisher is None:
            self._fisher = C.c_void_p(self.lib.orc_fisher_create(C.byref(self.params)))
        return self._fisher

    def fisher_generate(self, path=None):
        return self.lib.orc_fisher_generate(self.fisher, path.encode() if path else None)

    def fisher_load(self, path):
        return self.lib.orc_fisher_load(self.fisher, path.encode())

    def fisher_records(self):
        n = int(self.lib.orc_fisher_num_records(self.fisher))
        out = np.zeros((n, 4), dtype=np.float32)
        self.lib.orc_fisher_records(self.fisher, C.c_void_p(out.ctypes.data), C.c_uint64(n))
        return out

    def fisher_num_entries(self):
        return int(self.lib.orc_fisher_num_entries(self.fisher))

    def voxel_information(self, x, y, z):
        return float(self.lib.orc_fisher_voxel_information(self.fisher, C.c_float(x), C.c_float(y), C.c_float(z)))

    def voxel_coordinate(self, x, y, z):
        out = (C.c_float * 3)()
        self.lib.orc_fisher_voxel_coordinate(self.fisher, C.c_float(x), C.c_float(y), C.c_float(z), out)
        return tuple(out)

    def information_of_point_local(self, x, y, z):
        return float(self.lib.orc_information_of_point_local(C.c_float(x), C.c_float(y), C.c_float(z)))

    def crowd_factor(self, num):
        return float(self.lib.orc_crowd_factor(self.fisher, C.c_int(num)))

    def fisher_poses(self, xyz, poses):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        poses = np.ascontiguousarray(poses)
        n = len(poses)
        info = np.zeros(n, dtype=np.float32)
        safe = np.zeros(n, dtype=np.uint8)
        nvis = np.zeros(n, dtype=np.uint32)
        nused = np.zeros(n, dtype=np.uint32)
        self.last_information_f64 = np.zeros(n, dtype=np.float64)
        st = self.lib.orc_fisher_poses(self.fisher, C.c_void_p(xyz.ctypes.data), C.c_uint64(len(xyz)),
                                       C.c_void_p(poses.ctypes.data), C.c_uint32(n), C.c_void_p(info.ctypes.data),
                                       C.c_void_p(safe.ctypes.data), C.c_void_p(nvis.ctypes.data),
                                       C.c_void_p(nused.ctypes.data), C.c_void_p(self.last_information_f64.ctypes.data))
        return st, info, safe, nvis, nused

    def landmarks_in_view(self, xyz, pose, cap=1 << 20):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        pose = np.ascontiguousarray(pose)
        out = np.zeros((cap, 3), dtype=np.float32)
        cnt = C.c_uint32()
        self.lib.orc_landmarks_in_view(self.fisher, C.c_void_p(xyz.ctypes.data), C.c_uint64(len(xyz)),
                                       C.c_void_p(pose.ctypes.data), C.c_void_p(out.ctypes.data), C.c_uint32(cap),
                                       C.byref(cnt))
        return out[:cnt.value]

    def landmarks_in_view_keyframes(self, xyz, keyframes, point_index, pose, max_landmarks=1000, max_distance=5.0,
                                    exhaustive=False, cap=1 << 20):
        """mapPointsVisibleFromPose over a keyframe atlas (ORBW/src/orb_slam3_interface.cpp:188-317): (xyz in the pose frame,
        keyframes checked)."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        keyframes, point_index, pose = np.ascontiguousarray(keyframes), np.ascontiguousarray(point_index, dtype=np.uint32), np.ascontiguousarray(pose)
        out = np.zeros((cap, 3), dtype=np.float32)
        cnt, chk = C.c_uint32(), C.c_uint32()
        self.lib.orc_landmarks_in_view_keyframes(self.fisher, C.c_void_p(xyz.ctypes.data), C.c_uint64(len(xyz)),
                                                 C.c_void_p(keyframes.ctypes.data), C.c_uint32(len(keyframes)),
                                                 C.c_void_p(point_index.ctypes.data), C.c_void_p(pose.ctypes.data),
                                                 C.c_int32(max_landmarks), C.c_float(max_distance), C.c_int(int(exhaustive)),
                                                 C.c_void_p(out.ctypes.data), C.c_uint32(cap), C.byref(cnt), C.byref(chk))
        return out[:cnt.value], chk.value

    def fisher_pose_keyframes(self, xyz, keyframes, point_index, pose, max_landmarks=1000, max_distance=5.0, exhaustive=False):
        """isPoseSafe behind that service: (status, information, information accumulated in double, safe, landmarks, used)."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        keyframes, point_index, pose = np.ascontiguousarray(keyframes), np.ascontiguousarray(point_index, dtype=np.uint32), np.ascontiguousarray(pose)
        info, info64, safe = C.c_float(), C.c_double(), C.c_uint8()
        n, used = C.c_uint32(), C.c_uint32()
        st = self.lib.orc_fisher_pose_keyframes(self.fisher, C.c_void_p(xyz.ctypes.data), C.c_uint64(len(xyz)),
                                                C.c_void_p(keyframes.ctypes.data), C.c_uint32(len(keyframes)),
                                                C.c_void_p(point_index.ctypes.data), C.c_void_p(pose.ctypes.data),
                                                C.c_int32(max_landmarks), C.c_float(max_distance), C.c_int(int(exhaustive)),
                                                C.byref(info), C.byref(info64), C.byref(safe), C.byref(n), C.byref(used))
        return st, info.value, info64.value, safe.value, n.value, used.value

    def se3_roundtrip(self, R, t):
        R = np.ascontiguousarray(R, dtype=np.float32)
        t = np.ascontiguousarray(t, dtype=np.float32)
        Ro, to = np.zeros((3, 3), dtype=np.float32), np.zeros(3, dtype=np.float32)
        self.lib.orc_se3_roundtrip(C.c_void_p(R.ctypes.data), C.c_void_p(t.ctypes.data), C.c_void_p(Ro.ctypes.data),
                                   C.c_void_p(to.ctypes.data))
        return Ro, to

    def orientation_around_z(self, angle):
        q = (C.c_double * 4)()
        self.lib.orc_orientation_around_z(C.c_double(angle), q)
        return tuple(q)

    def cycle(self, a, meta, rx, ry, ryaw, poly):
        a, m = self._map(a, meta)
        poly = np.asarray(poly, dtype=np.float64)
        nf, ne = C.c_uint32(), C.c_uint32()
        st = self.lib.orc_cycle(*m, C.byref(self.params), C.c_double(rx), C.c_double(ry), C.c_double(ryaw),
                                C.c_void_p(poly.ctypes.data), C.byref(nf), C.byref(ne))
        return st, nf.value, ne.value
