"""HerReplay -- hindsight experience replay on the device for CtrReachVecEnv.

The reference trains with stable_baselines' HER wrapper (src/utils.py:5-7,27-28; saved_policies/**/config.yml:
goal_selection_strategy future, n_sampled_goal 4, buffer_size 500000, batch_size 256).  Here the ring lives in HBM next to the
environments, so a step feeds a learner without a host copy: store() after every env.step(), sample() returns a batch whose
goals were relabelled with the 'future' strategy and whose rewards come from the reference's compute_reward
(ctr_reach_env.py:160-170).  All work is done by libctr_b200.so (ctr_her_begin / ctr_her_store / ctr_her_sample).

Flat observations: the reference's tests flatten the dict as (observation, desired_goal, achieved_goal)
(tests/unit_tests.py:19,46) while stable_baselines' HERGoalEnvWrapper uses (observation, achieved_goal, desired_goal);
`flat_order` selects the order used by flat().
"""
import ctypes

import torch

from . import _lib

KEY_ORDER_SB = ('observation', 'achieved_goal', 'desired_goal')
KEY_ORDER_REFERENCE_TESTS = ('observation', 'desired_goal', 'achieved_goal')

_RING_INTS = (('done', torch.uint8), ('t_in_ep', torch.int32), ('ep_len', torch.int32))
_BATCH = (('obs', 'D', torch.float64), ('next_obs', 'D', torch.float64), ('achieved', 3, torch.float64),
          ('next_achieved', 3, torch.float64), ('desired', 3, torch.float64), ('action', 6, torch.float32),
          ('reward', 0, torch.float64), ('done', 0, torch.uint8), ('src', 0, torch.int64), ('goal_src', 0, torch.int64))


class HerReplay(object):
    """Time-major ring of `capacity_steps` x env.num_envs transitions in device memory.

    n_sampled_goal k: a sampled transition is relabelled with probability k / (k + 1) (the online form of storing k
    relabelled copies beside every real transition).  rewrite_error=True additionally rewrites the `desired - achieved` entries
    of the observation (obs.py:158) for relabelled samples -- an extension: the reference's wrapper only swaps desired_goal."""

    def __init__(self, env, capacity_steps, n_sampled_goal=4, rewrite_error=False, seed=0, flat_order=KEY_ORDER_SB):
        self.env, self.capacity = env, int(capacity_steps)
        self.n_sampled_goal, self.rewrite_error, self.seed = int(n_sampled_goal), bool(rewrite_error), int(seed)
        self.flat_order = tuple(flat_order)
        assert sorted(self.flat_order) == sorted(KEY_ORDER_SB)
        N, D, dev = env.num_envs, env.obs_dim, env.device
        # record major: transition (slot, env) = row slot * N + env of `rec`, 44 doubles = 352 B = 11 aligned sectors
        # (include/ctr_b200.h gives the field offsets), so that a sample gathers whole sectors
        self.ring = {'rec': torch.zeros((self.capacity * N, _lib.HER_RECORD), dtype=torch.float64, device=dev)}
        self.ring.update({k: torch.zeros(self.capacity * N, dtype=dt, device=dev) for k, dt in _RING_INTS})
        self.ring.update(cur_obs=torch.zeros((D, N), dtype=torch.float64, device=dev),
                         cur_achieved=torch.zeros((3, N), dtype=torch.float64, device=dev),
                         cur_desired=torch.zeros((3, N), dtype=torch.float64, device=dev),
                         cur_t=torch.zeros(N, dtype=torch.int32, device=dev))
        self._c = _lib.CtrHerRing(n_envs=N, capacity=self.capacity, obs_dim=D, pad_=0,
                                  **{k: v.data_ptr() for k, v in self.ring.items()})
        self.stored, self.newest, self.calls = 0, -1, 0

    def records(self):
        """Decoded view of the ring for inspection / checkpointing: dict of (capacity, N, ...) tensors."""
        N, D, c = self.env.num_envs, self.env.obs_dim, self.capacity
        rec = self.ring['rec'].reshape(c, N, _lib.HER_RECORD)
        out = dict(obs=rec[..., :D], next_obs=rec[..., D:2 * D], achieved=rec[..., 2 * D:2 * D + 3],
                   next_achieved=rec[..., 2 * D + 3:2 * D + 6], desired=rec[..., 2 * D + 6:2 * D + 9], reward=rec[..., 2 * D + 9],
                   action=rec[..., 2 * D + 10:2 * D + 13].contiguous().view(torch.float32))
        out.update({k: self.ring[k].reshape(c, N) for k, _ in _RING_INTS})
        return out

    def begin(self):
        """Call after env.reset(): the next stored transition starts from the env's current observation."""
        _lib.check(self.env.lib.ctr_her_begin(self.env.handle, ctypes.byref(self._c), self.env._stream()))

    def store(self, actions):
        """Store the transition of the env.step(actions) that was just issued (float32 CUDA tensor (N, 6) or (6, N))."""
        a, N = actions, self.env.num_envs
        if a.shape == (N, 6):
            sf, se = a.stride(1), a.stride(0)
        elif a.shape == (6, N):
            sf, se = a.stride(0), a.stride(1)
        else:
            raise ValueError("actions must have shape (num_envs, 6) or (6, num_envs)")
        slot = (self.newest + 1) % self.capacity
        _lib.check(self.env.lib.ctr_her_store(self.env.handle, ctypes.byref(self._c), _lib.ptr(a), sf, se, slot,
                                              self.env._stream()))
        self.newest, self.stored = slot, min(self.stored + 1, self.capacity)

    def alloc_batch(self, m):
        D, dev = self.env.obs_dim, self.env.device
        rows = lambda r: D if r == 'D' else r
        return {k: torch.empty(((rows(r), m) if r else (m,)), dtype=dt, device=dev) for k, r, dt in _BATCH}

    def sample(self, m, out=None, tol=None, call=None):
        """-> dict of SoA device tensors (obs/next_obs [D][m], achieved/next_achieved/desired [3][m], action [6][m], reward,
        done, src, goal_src [m]).  `call` numbers the draw (default: an internal counter), so a batch is reproducible."""
        if self.stored < 1:
            raise RuntimeError("HerReplay.sample() before any store()")
        out = self.alloc_batch(m) if out is None else out
        c = _lib.CtrHerBatch(m=m, **{k: out[k].data_ptr() for k, _, _ in _BATCH})
        call = self.calls if call is None else int(call)
        tol = float(self.env.goal_tolerance.get_tol() if tol is None else tol)
        _lib.check(self.env.lib.ctr_her_sample(self.env.handle, ctypes.byref(self._c), self.stored, self.newest,
                                               self.n_sampled_goal, int(self.rewrite_error), tol,
                                               self.seed & 0xFFFFFFFFFFFFFFFF, call, ctypes.byref(c), self.env._stream()))
        self.calls = call + 1
        return out

    def flat(self, batch, which='obs'):
        """(m, D + 6) flat observation of a sampled batch in `flat_order` (which = 'obs' | 'next_obs')."""
        part = {'observation': batch[which], 'desired_goal': batch['desired'],
                'achieved_goal': batch['achieved' if which == 'obs' else 'next_achieved']}
        return torch.cat([part[k] for k in self.flat_order], dim=0).t()
