"""CPU tests of the host side: C-ABI surface (library loads, every symbol of include/ctr_b200.h is exported, config
struct layout), spaces / registration / tolerance schedule mirrors, sharding helpers, and the no-CPU-fallback rule."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    import sys
    sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
    import build
    build.build()
    from ctr_reach_envs import _lib
    return _lib


def _header_functions():
    src = open(os.path.join(ROOT, 'include', 'ctr_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(ctr_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_header_symbol(lib):
    names = _header_functions()
    assert len(names) >= 25
    so = ctypes.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(so, n), n
    assert sorted(lib.EXPORTS) == names            # the python binding covers the whole header
    assert lib.load().ctr_abi_version() == lib.ABI_VERSION == 2


def test_config_struct_layout_matches_header(lib, tmp_path):
    """sizeof / offsetof of ctr_config as gcc sees the header == the ctypes mirror."""
    fields = [f[0] for f in lib.CtrConfig._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "ctr_b200.h"', 'int main(void){',
            'printf("%zu\\n", sizeof(ctr_config));']
    prog += ['printf("%%zu\\n", offsetof(ctr_config, %s));' % f for f in fields]
    prog += ['printf("%d\\n", (int)CTR_F_COUNT);', 'return 0;}']
    c = tmp_path / 'layout.c'
    c.write_text('\n'.join(prog))
    exe = tmp_path / 'layout'
    subprocess.check_call(['gcc', '-I', os.path.join(ROOT, 'include'), str(c), '-o', str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out[0] == ctypes.sizeof(lib.CtrConfig)
    for f, off in zip(fields, out[1:-1]):
        assert getattr(lib.CtrConfig, f).offset == off, f
    assert out[-1] == lib.F_COUNT


def test_no_cpu_fallback(lib):
    """Without a CUDA device the product path must fail loudly (never route through the oracle)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    import ctr_reach_envs
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ctr_reach_envs.make('CTR-Reach-v0')
    cfg = lib.make_config(np.ones((1, 3, 8)))
    h = ctypes.c_void_p()
    assert lib.load().ctr_create(ctypes.byref(cfg), 16, 0, ctypes.byref(h)) != 0
    assert b'no CUDA device' in lib.load().ctr_last_error()
    src = ''
    for dp, _, fs in os.walk(os.path.join(ROOT, 'gym-ctr-reach_b200')):
        for f in fs:
            if f.endswith(('.py', '.cu', '.cuh')):
                src += open(os.path.join(dp, f)).read()
    assert 'ctr_oracle' not in src and 'import oracle' not in src      # the product never touches oracle/


def test_registration_and_default_kwargs(lib, oracle_py):
    import ctr_reach_envs
    spec = ctr_reach_envs.spec('CTR-Reach-v0')
    assert spec.entry_point == 'ctr_reach_envs.envs:CtrReachEnv' and spec.max_episode_steps == 150
    kw = spec.kwargs
    assert kw['n_substeps'] == 10 and kw['max_steps_per_episode'] == 150 and kw['select_systems'] == [0]
    assert kw['extension_action_limit'] == 0.001 and kw['rotation_action_limit'] == 5
    assert kw['goal_tolerance_parameters'] == dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=200000,
                                                   function='constant', set_tol=0)
    from ctr_reach_envs.envs.vec_env import systems_table
    table = systems_table(kw['ctr_systems_parameters'], [0, 1, 2, 3])
    assert np.array_equal(table, oracle_py.DEFAULT_SYSTEMS)            # same four tube systems as the reference
    assert np.array_equal(systems_table(kw['ctr_systems_parameters'], [2]), oracle_py.DEFAULT_SYSTEMS[2:3])


def test_spaces_match_reference_definitions(lib, oracle_py):
    from ctr_reach_envs.envs.vec_env import build_spaces
    from ctr_reach_envs.envs.goal_tolerance import GoalTolerance
    gt = GoalTolerance(dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=200000, function='constant', set_tol=0))
    L = oracle_py.DEFAULT_SYSTEMS[:, :, 0]
    obs_space, act_space = build_spaces(L[:1], gt, 1, 0.001, 5)
    assert act_space.dtype == np.float32 and act_space.shape == (6,)
    assert np.array_equal(act_space.high, np.array([1e-3] * 3 + [np.deg2rad(5)] * 3).astype(np.float32))
    assert np.array_equal(act_space.low, -act_space.high)
    o = obs_space['observation']
    assert o.shape == (13,) and o.low[12] == np.float32(0.001) and o.high[12] == np.float32(0.02)
    assert np.allclose(o.low[[2, 5, 8]], -L[0] + 1e-4)
    obs4, _ = build_spaces(L, gt, 4, 0.001, 5)
    assert obs4['observation'].shape == (14,) and obs4['observation'].high[13] == 3
    a = act_space.sample()
    assert a.dtype == np.float32 and act_space.contains(a) and not act_space.contains(a * 100 + 1)
    assert obs_space['desired_goal'].shape == (3,)


def test_goal_tolerance_matches_reference(golden, lib):
    from ctr_reach_envs.envs.goal_tolerance import GoalTolerance
    g = golden('tolerance_reference')
    for key in g.files:
        if key == 'timesteps':
            continue
        fn, n_ts = key.rsplit('_', 1)
        gt = GoalTolerance(dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=int(n_ts), function=fn, set_tol=0))
        vals = [gt.get_tol()]
        for t in g['timesteps']:
            gt.update(int(t))
            vals.append(gt.get_tol())
        assert np.array_equal(np.array(vals), g[key]), key
    pinned = GoalTolerance(dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=10, function='linear', set_tol=0.005))
    pinned.update(3)
    assert pinned.get_tol() == 0.005
    with pytest.raises(AssertionError):
        GoalTolerance(dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=10, function='cosine', set_tol=0))


def test_obs_utils_round_trip(lib):
    from ctr_reach_envs.envs import obs_utils as U
    q = np.array([-0.2, -0.1, -0.05, 0.3, -1.2, 2.9])
    assert np.allclose(U.ego2prop(U.prop2ego(q)), q)
    assert np.allclose(U.rep2joint(U.joint2rep(q)), q)
    assert np.allclose(U.joint2rep(q)[:3], [np.cos(0.3), np.sin(0.3), -0.2])


def test_shard_range_partitions(lib):
    from ctr_reach_envs.distributed import shard_range
    for total, world in ((1 << 20, 8), (1000, 3), (7, 8), (4096, 1)):
        spans = [shard_range(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1
    assert shard_range(1 << 20, 3, 8) == (3 * 131072, 4 * 131072)


def test_evaluation_csv_round_trip(tmp_path):
    """The reference's evaluation files (src/rotation_evaluation.py:66-81) read and written back byte for byte; the fixture is
    the head of a file the reference ships (oracle/make_golden_csv.py)."""
    from ctr_reach_envs import evaluation as E
    src = os.path.join(ROOT, 'tests', 'golden', 'evaluations_sample.csv')
    df = E.read_evaluations(src)
    assert list(df.columns) == E.EVAL_COLUMNS and df.shape == (25, 29)
    out = tmp_path / 'evaluations.csv'
    E.write_evaluations(str(out), df)
    assert out.read_text() == open(src).read()
    # make_frame stacks in the reference's column order (rotation_evaluation.py:66-68)
    q = np.arange(12, dtype=np.float32).reshape(2, 6)
    f = E.make_frame(np.zeros((2, 3)), np.ones((2, 3)), [5, 6], np.full((2, 3), 2.0), q, q + 100, q + 200, [0, 3])
    assert f['B_desired_2'].tolist() == [1.0, 7.0] and f['alpha_achieved_1'].tolist() == [103.0, 109.0]
    assert f['alpha_starting_3'].tolist() == [205.0, 211.0] and f['episode_length'].tolist() == [5.0, 6.0]
    assert f['system_id'].tolist() == [0.0, 3.0]
    with pytest.raises(ValueError):
        E.write_evaluations(str(out), f.drop(columns=['system_id']))
