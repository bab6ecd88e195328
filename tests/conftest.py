import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu on the GPU box")


def _cuda_device_count() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    # A test marked gpu on a box without a device is an error in how the suite was invoked, not a
    # silent skip -- unless the user did not ask for gpu tests specifically (plain `pytest tests`).
    if _cuda_device_count() > 0:
        return
    asked = "gpu" in (config.getoption("-m") or "") and "not gpu" not in (config.getoption("-m") or "")
    if asked:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_cases():
    z = np.load(os.path.join(GOLDEN, "r2iq_cases.npz"))
    cases = []
    for entry in z["cases"]:
        name, sig = str(entry).split(":")
        d, tb, rnd, sb, nb, use_hist = [int(v) for v in z[f"{name}__meta"]]
        cases.append(dict(name=name, d=d, tunebin=tb, rand=bool(rnd), sideband=bool(sb), n_blocks=nb,
                          use_hist=bool(use_hist), gain=float(z[f"{name}__gain"][0]),
                          clean_in=z[f"input__{sig}"][: nb * 65536], out=z[f"{name}__out"]))
    return dict(cases=cases, history=z["history"])


@pytest.fixture(scope="session")
def golden_cases_8192():
    """FFTN_R_ADC 8192 cases (tests/golden/r2iq_cases_8192.npz): outputs of the reference compiled at the size it ships
    with; the inputs are the ones stored in r2iq_cases.npz."""
    z = np.load(os.path.join(GOLDEN, "r2iq_cases_8192.npz"))
    zin = np.load(os.path.join(GOLDEN, "r2iq_cases.npz"))
    cases = []
    for entry in z["cases"]:
        name, sig = str(entry).split(":")
        d, tb, rnd, sb, nb, use_hist = [int(v) for v in z[f"{name}__meta"]]
        cases.append(dict(name=name, d=d, tunebin=tb, rand=bool(rnd), sideband=bool(sb), n_blocks=nb,
                          use_hist=bool(use_hist), gain=float(z[f"{name}__gain"][0]),
                          clean_in=zin[f"input__{sig}"][: nb * 65536], out=z[f"{name}__out"]))
    return dict(cases=cases, history=z["history"])


def case_input(case, history):
    """Wire-format input (randomized if the case has the ADC randomizer on) and history."""
    from oracle import r2iq_oracle as orc
    x = orc.randomize(case["clean_in"]) if case["rand"] else case["clean_in"]
    h = None
    if case["use_hist"]:
        h = orc.randomize(history) if case["rand"] else history
    return x, h
