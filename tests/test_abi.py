"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol the header declares,
and refuses to run without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "fitslam_cuda.h")).read()
    return sorted(set(re.findall(r"FITSLAM_API[^;(]*?\b(fitslam_\w+)\s*\(", txt)))


def test_header_and_binding_agree(capi):
    assert header_symbols() == sorted(capi.PROTOTYPES)


def test_library_exports_every_symbol(capi):
    lib = C.CDLL(capi.LIB_PATH)
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_abi_version_and_defaults(capi):
    lib = capi.load_library()
    assert lib.fitslam_abi_version() == 2
    p = capi.default_params(lib)
    assert (p.min_frontier_cluster_size, p.max_frontier_cluster_size, p.lethal_threshold) == (1, 20, 160)
    assert (p.max_camera_depth, p.delta_theta, p.camera_fov, p.robot_radius) == (2.0, 0.10, 1.04, 0.60)
    assert (p.alpha, p.beta, p.vx_max, p.wz_max) == (0.25, 1.0, 0.5, 0.5)
    assert p.fisher_information_threshold == 550.0 and abs(p.voxel_step - 0.3) < 1e-7


def test_struct_layouts_match_header(capi):
    # sizes the C side static_asserts (csrc/api.cu); keeps ctypes/numpy views honest
    assert C.sizeof(capi.Params) == 192
    assert capi.FRONTIER_DTYPE.itemsize == 40 and capi.ARRIVAL_DTYPE.itemsize == 24
    assert C.sizeof(capi.Selection) == 24 and C.sizeof(capi.CycleResult) == 32 + 24 + 32 + 8


def test_params_shared_with_oracle(capi, oracle):
    a, b = capi.default_params(), oracle.params
    assert bytes(a) == bytes(b)


def test_no_cpu_fallback(capi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present")
    with pytest.raises(capi.FitslamError) as e:
        capi.Context()
    assert e.value.status == capi.E_CUDA
