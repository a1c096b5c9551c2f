import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _cuda_ok():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """libssn.so and the oracle port, compiled once per session."""
    if os.environ.get("SSN_SKIP_BUILD"):       # sanitizer runs preload a runtime that make's subprocesses cannot carry
        return True
    import __graft_entry__ as g
    g.build(quiet=True)
    return True


# parameter sets every parity test runs (name -> kwargs for Oracle/Reference and ssn_params)
CONFIGS = {
    "default": dict(),
    "joint": dict(use_joint=1, somatic_rate=0.01),
    "joint_s": dict(use_joint=1, somatic_rate=0.001, min_somatic_qual=0),
    "nopriors_q0": dict(use_priors=0, min_somatic_qual=0),
    "LG": dict(include_loh=0, include_gor=0, min_somatic_qual=0),
    "TNr": dict(theta=0.7, n_hap=3, het_rate=0.01, min_somatic_qual=5),
}


def oracle_kwargs(cfg):
    return dict(CONFIGS[cfg])


def ssn_params(cfg):
    from somatic_sniper_b200 import api
    k = CONFIGS[cfg]
    m = dict(theta="theta", n_hap="n_hap", het_rate="het_rate", use_priors="use_priors", use_joint="use_joint_priors",
             somatic_rate="somatic_mutation_rate", min_somatic_qual="min_somatic_qual", include_loh="include_loh",
             include_gor="include_gor")
    return api.default_params(**{m[a]: v for a, v in k.items()})
