import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (REPO, os.path.join(REPO, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(REPO, "tests", "golden")

TYPES15 = [[(m >> 3) & 1, (m >> 2) & 1, (m >> 1) & 1, m & 1, 0, 0, 0] for m in range(1, 16)]
NODR = dict(tremor_amplitude_range=[1.0, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
            max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.0, dr_actuator_range=0.0,
            matrix_noise_fraction=0.0)
TRAIN = dict(tremor_amplitude_range=[0.1, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
             max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.03,
             matrix_noise_fraction=0.1)
EVAL = dict(tremor_amplitude_range=[0.95, 1.05], first_harmonics_interval=[3.75, 6.25], second_harmonics_interval=[7.5, 12.5],
            max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.025, dr_actuator_range=0.04,
            matrix_noise_fraction=0.125)
SETTING_KEYS = tuple(NODR.keys())


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def case_settings(g, case):
    """Constructor kwargs stored with a golden case."""
    out = {}
    for k in SETTING_KEYS:
        v = g["%s/meta_%s" % (case, k)]
        out[k] = [float(x) for x in v] if v.ndim else float(v)
    return out


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def exo():
    import exo_b200
    return exo_b200
