"""ctr_jacobian_noalloc (include/ctr_b200.h): the finite-difference Jacobian with its scratch in a static device array of the
library, for callers that cannot call cudaMalloc.  No counterpart in the reference;
the existing entry point ctr_jacobian is the comparison.  This file is collected last on purpose."""
import numpy as np
import pytest


@pytest.mark.gpu
def test_jacobian_noalloc_matches_ctr_jacobian_on_two_calls_in_a_row(oracle_py, vec_factory):
    import torch
    from ctr_reach_envs import _lib
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    dev = env.device
    rng = np.random.RandomState(5)

    def batch(n):
        sysid = rng.randint(4, size=n).astype(np.int32)
        q = np.stack([oracle_py.sample_goal(rng, oracle_py.DEFAULT_SYSTEMS[s][:, 0]) for s in sysid]).astype(np.float64)
        return q, sysid

    # two calls in a row with different sizes and contents: the second must not see anything of the first in the shared array
    for n in (1500, 333):
        q, sysid = batch(n)
        J_ref, tip_ref = env.jacobian(q, system=sysid)                                   # existing entry point (handle scratch, cudaMalloc)
        J_ref = J_ref.reshape(n, 18).t().contiguous()
        tip_ref = tip_ref.t().contiguous()
        q_soa = torch.as_tensor(q, dtype=torch.float64, device=dev).t().contiguous()
        s_dev = torch.as_tensor(sysid, dtype=torch.int32, device=dev)
        J = torch.full((18, n), float('nan'), dtype=torch.float64, device=dev)
        tip = torch.full((3, n), float('nan'), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        # straight through the C-ABI: handle, device pointers, sizes
        rc = env.lib.ctr_jacobian_noalloc(env.handle, _lib.ptr(q_soa), _lib.ptr(s_dev), n, 1e-4, _lib.ptr(J), _lib.ptr(tip),
                                          env._stream())
        assert rc == 0, env.lib.ctr_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(dev)[0] >= free_before                            # nothing was allocated on the device
        # same kernels on the same inputs: the tolerance of tests/test_ik.py (tips 1e-10 m, J 2e-6) is met with room to spare
        assert np.abs(tip.cpu().numpy() - tip_ref.cpu().numpy()).max() < 1e-10
        assert np.abs(J.cpu().numpy() - J_ref.cpu().numpy()).max() < 2e-6
        assert np.array_equal(J.cpu().numpy(), J_ref.cpu().numpy()) and np.array_equal(tip.cpu().numpy(), tip_ref.cpu().numpy())
    # the python wrapper (caller-owned tensors, tip optional, no system ids = system 0)
    env0 = vec_factory(16)
    q = np.stack([oracle_py.sample_goal(rng, oracle_py.DEFAULT_SYSTEMS[0][:, 0]) for _ in range(64)]).astype(np.float64)
    q_soa = torch.as_tensor(q, dtype=torch.float64, device=dev).t().contiguous()
    J = torch.empty((18, 64), dtype=torch.float64, device=dev)
    env0.jacobian_noalloc(q_soa, J)
    J_ref, _ = env0.jacobian(q)
    assert np.array_equal(J.cpu().numpy(), J_ref.reshape(64, 18).t().cpu().numpy())


@pytest.mark.gpu
def test_jacobian_noalloc_rejects_batches_beyond_the_static_array(vec_factory):
    import torch
    from ctr_reach_envs import _lib
    env = vec_factory(16)
    n = env.JACOBIAN_NOALLOC_MAX + 1
    q = torch.zeros((6, n), dtype=torch.float64, device=env.device)
    J = torch.zeros((18, n), dtype=torch.float64, device=env.device)
    rc = env.lib.ctr_jacobian_noalloc(env.handle, _lib.ptr(q), None, n, 1e-4, _lib.ptr(J), None, env._stream())
    assert rc == 1
    assert b'CTR_JACOBIAN_NOALLOC_MAX' in env.lib.ctr_last_error()
    with pytest.raises(RuntimeError):
        env.jacobian_noalloc(q, J)
    # the largest accepted size runs (its scratch is the whole array)
    n = env.JACOBIAN_NOALLOC_MAX
    env.jacobian_noalloc(q[:, :n].contiguous(), J[:, :n].contiguous())
    torch.cuda.synchronize()
