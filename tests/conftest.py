import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'gym-ctr-reach_b200'), os.path.join(ROOT, 'oracle'), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)


@pytest.fixture(scope='session')
def golden():
    return load_golden


@pytest.fixture(scope='session')
def oracle_c():
    """The C oracle (oracle/libctr_oracle.so), built on demand.  Test infrastructure only."""
    import ctr_oracle_c
    ctr_oracle_c.build()
    return ctr_oracle_c


@pytest.fixture(scope='session')
def oracle_py():
    import ctr_oracle
    return ctr_oracle


@pytest.fixture(scope='session')
def tube_constants(oracle_c, oracle_py):
    return oracle_c.derive(oracle_py.DEFAULT_SYSTEMS)


def make_vec(num_envs, **overrides):
    """CtrReachVecEnv with the registered default kwargs + overrides (GPU tests)."""
    import ctr_reach_envs
    from ctr_reach_envs.envs import CtrReachVecEnv
    kw = ctr_reach_envs.default_kwargs()
    kw.update(overrides)
    return CtrReachVecEnv(num_envs, **kw)


@pytest.fixture
def vec_factory():
    made = []

    def factory(num_envs, **kw):
        v = make_vec(num_envs, **kw)
        made.append(v)
        return v

    yield factory
    for v in made:
        v.close()
