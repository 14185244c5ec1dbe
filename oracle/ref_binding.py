"""ctypes view of oracle/_ref/libfitslam_ref.so: the REFERENCE's own FrontierSearch / Frontier / ray helpers, compiled
from /root/reference where it lies (oracle/Makefile).  TEST INFRASTRUCTURE ONLY: it pins the restatement in oracle.cpp
(tests/test_ref_pin.py) and gives the GPU tests a second, independent expectation.  The .so is git-ignored and travels
to the GPU box with the snapshot; nothing here reads /root/reference at run time."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libfitslam_ref.so")


def available():
    return os.path.exists(LIB_PATH)


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
