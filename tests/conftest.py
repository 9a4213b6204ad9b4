import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DATA = os.path.join(ROOT, "data")
GEOM = os.path.join(DATA, "neuralGeometries")
MATCAPS = os.path.join(DATA, "matcaps")
GOLDEN = os.path.join(ROOT, "tests", "golden")
MODELS = ["car_1", "plane_1", "plane_2", "plane_3", "3a3d4a90a2db90b4203936772104a82d.obj"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def model_path(name):
    return os.path.join(GEOM, name + ".h5")


@pytest.fixture(scope="session")
def built():
    """Builds libnsdf.so and the oracle once per session."""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def gpu(built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def _oracle_facing_render_state():
    """host.py starts like the library (the reference's HEAD build: manySphere scene, reference math).  The tests compare with the
    CPU oracle unless they say otherwise, so every test starts from the tanh scene in the deterministic arithmetic."""
    try:
        from cudaneuralrender_b200 import host
    except Exception:
        yield
        return
    host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0, host.MATH_DETERMINISTIC)
    yield
    host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0, host.MATH_DETERMINISTIC)
