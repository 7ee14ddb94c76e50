import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def bank():
    """The 200 reference scenarios (gen_assoc_mult_slice.py, seed 10) as a compiled bank."""
    import golden_inputs
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    st, us = golden_inputs.load_scenarios()
    b = scn.ScenarioBank()
    for i in range(len(st)):
        b.add_typed(st[i], us[i])
    return b
