import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_package():
    """The package directory is `fit-slam_b200` (hyphen): import it under the name fitslam_b200."""
    if "fitslam_b200" in sys.modules:
        return sys.modules["fitslam_b200"]
    pkg_dir = os.path.join(ROOT, "fit-slam_b200")
    spec = importlib.util.spec_from_file_location("fitslam_b200", os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["fitslam_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    return load_package()


@pytest.fixture(scope="session")
def capi(pkg):
    return pkg.capi


@pytest.fixture(scope="session")
def synth(pkg):
    import importlib
    return importlib.import_module("fitslam_b200.synth")


@pytest.fixture()
def oracle(capi):
    from oracle.binding import Oracle
    return Oracle(capi)


@pytest.fixture()
def ctx(capi):
    c = capi.Context()
    yield c
    c.close()
