import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    """The real reference; only where oracle/_ref/libref.so exists."""
    from oracle_lib import Ref, build_ref
    if not Ref.available():
        try:
            if not build_ref():
                pytest.skip("oracle/_ref not built and /root/reference absent")
        except Exception as e:  # pragma: no cover
            pytest.skip(f"cannot build oracle/_ref: {e}")
    return Ref()


@pytest.fixture(scope="session")
def lamm():
    import lamm_b200
    lamm_b200.lib()
    return lamm_b200


@pytest.fixture(scope="session")
def gpu(lamm):
    """Initialised CUDA library; fails loudly (never skips) when the device is missing."""
    import torch
    assert torch.cuda.is_available(), "gpu-marked test needs a CUDA device"
    torch.cuda.set_device(0)
    lamm.init(0)
    return lamm


@pytest.fixture(autouse=True)
def _forget_host_weights(request):
    """The weight registry is keyed by host address; numpy re-uses the addresses of freed test arrays, so
    every GPU test ends by dropping the registry (what an application does when it frees a model)."""
    yield
    if request.node.get_closest_marker("gpu") is not None:
        import lamm_b200
        lamm_b200.lib().lammcu_weight_forget_all()


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def tol_c(c_ref, k):
    """C tolerance of the north star: 1e-5 * sqrt(K) * max|C| per matrix."""
    return 1e-5 * np.sqrt(k) * float(np.abs(c_ref).max())


def make_x(rng, n, k):
    """Activations with the quantizer edge cases (zero row, zero blocks, ties, halves)."""
    x = rng.standard_normal((n, k)).astype(np.float32)
    if n > 1:
        x[1] = 0.0
    if n > 2:
        x[2, : k // 2] = 0.0
    if n > 3:
        x[3, 0::7] = 2.5
        x[3, 5] = -2.5
    if n > 4:
        x[4] = np.round(x[4] * 4) / 4
        x[4, 0] = 127.0 / 2
    return x
