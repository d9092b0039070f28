"""Shared fixtures.  `-m "not gpu"` runs here (no GPU): the oracle against its self-checks and fixtures, host logic, and that
the C-ABI library loads and exports what include/b3d.h declares.  `-m gpu` tests are the parity tests proper (CUDA path through
the C-ABI versus the oracle / committed golden vectors).  Nothing here reads /root/reference."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
PRIOR_TXT = os.path.join(GOLDEN, "pose_prior.txt")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pkg(name=""):
    return importlib.import_module("3dbodyanimation_b200" + ("." + name if name else ""))


@pytest.fixture(scope="session")
def synth():
    return pkg("synth")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle_api
    oracle_api.lib()
    return oracle_api


@pytest.fixture(scope="session")
def model_np(synth):
    return synth.make_model(with_posedirs=True)


@pytest.fixture(scope="session")
def prior_np(synth):
    return synth.load_pose_prior(PRIOR_TXT)


@pytest.fixture(scope="session")
def omodel(orc, model_np):
    return orc.Model(model_np)


@pytest.fixture(scope="session")
def omodel_pd(orc, model_np):
    return orc.Model(model_np, use_posedirs=True)


@pytest.fixture(scope="session")
def ogmm(orc, prior_np):
    return orc.GMM(prior=prior_np)


@pytest.fixture(scope="session")
def api():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.init()
    return pkg("api")


@pytest.fixture(scope="session")
def gmodel(api, model_np):
    return api.AvatarModel(model_np)


@pytest.fixture(scope="session")
def ggmm(api, prior_np):
    return api.GaussianMixture(prior=prior_np)


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
