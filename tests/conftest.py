import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# several ranks x lanes of the tile renderer share one GPU in the tests: enough hardware queues that their streams never alias
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import cudapathtracer_b200 as nb
        return nb.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Builds the product library and the oracle once per session (both are in-tree, prebuilt on the GPU box)."""
    import __graft_entry__ as ge
    ge.build()
    return True


@pytest.fixture(scope="session")
def scene_dirs(built, tmp_path_factory):
    """Generated stand-in scenes (small tessellations for tests)."""
    import cudapathtracer_b200 as nb
    root = tmp_path_factory.mktemp("scenes")
    out = {}
    for name, detail in (("cornell", 100), ("dragon", 8), ("water", 12)):
        d = str(root / name)
        nb.make_scene(name, d, detail)
        out[name] = d
    return out


@pytest.fixture(scope="session")
def py_scenes(scene_dirs):
    from scene_utils import load_scene_py
    return {k: load_scene_py(v) for k, v in scene_dirs.items()}


@pytest.fixture(scope="session")
def oracle_scenes(py_scenes):
    from oracle.oracle_binding import OracleScene
    out = {}
    for k, (cfg, pos, nrm, uv, tris, mats) in py_scenes.items():
        out[k] = OracleScene(pos, nrm, uv, tris, mats, None, 2)
    return out


@pytest.fixture()
def rng():
    return np.random.default_rng(12345)
