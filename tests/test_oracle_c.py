"""CPU suite, part 3: the plain-C restatement (oracle/r2iq_oracle.c) against the golden vectors."""
import os

import numpy as np

from conftest import GOLDEN, case_input
from oracle import c_lib
from oracle import r2iq_oracle as orc


def test_c_port_matches_golden(golden_cases):
    for c in golden_cases["cases"]:
        if c["d"] == 0 and c["name"] != "hf64_sine5m_d0_tuned":
            continue   # one full-rate case is enough for the slow scalar port
        x, h = case_input(c, golden_cases["history"])
        got = c_lib.r2iq(x, c["d"], c["tunebin"], gain=c["gain"], rand=c["rand"], sideband=c["sideband"], history=h)
        assert orc.rel_l2(got, c["out"]) < 5e-7, c["name"]


def test_c_port_filter_and_convert():
    z = np.load(os.path.join(GOLDEN, "filters.npz"))
    for d in range(7):
        assert orc.rel_l2(c_lib.design_filter(1.08e-8, d), z[f"hw_d{d}"]) < 2e-7
    cz = np.load(os.path.join(GOLDEN, "convert.npz"))
    assert np.array_equal(c_lib.convert(cz["x"], False), cz["plain"])
    assert np.array_equal(c_lib.convert(cz["x"], True), cz["rand"])
