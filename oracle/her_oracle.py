"""CPU restatement of the hindsight-replay path -- TEST INFRASTRUCTURE ONLY.

What is restated, and against what it is pinned:
  * compute_reward on the relabelled goal: the reference's own function (ctr_reach_envs/envs/ctr_reach_env.py:160-170),
    pinned by tests/golden/reward_reference.npz (generated from the unmodified reference).
  * the ring bookkeeping (what a transition is: obs before / after the step, terminal observation when the episode ends,
    episode index and length) follows what stable_baselines' HindsightExperienceReplayWrapper stores per episode.
  * the 'future' goal selection: donor = a uniformly drawn later transition of the same episode, goal = the achieved goal of
    that donor's observation, i.e. next_achieved of a transition in [k, L-1]; the last transition of an episode is never
    relabelled; relabelled transitions get done = False.  stable_baselines is THIRD PARTY, absent from /root/reference and
    from this image, and not version-pinned by the reference (setup.py:5): this part is restated from the published
    algorithm and its PARITY IS UNPINNED.  The reference's settings (future, n_sampled_goal 4) are the call-site anchor.
  * Philox4x32-10 (Salmon et al., SC'11), checked against the Random123 known-answer vectors in tests/test_her.py.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32(ctr, key):
    """ctr: uint32 [n][4], key: (k0, k1) -> uint32 [n][4] after 10 rounds."""
    c = [np.asarray(ctr[:, i], dtype=np.uint64) for i in range(4)]
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return np.stack(c, axis=1).astype(np.uint32)


def compute_reward(achieved, desired, tol, dense=False):
    """ctr_reach_env.py:168-170: d = ||a - d||_2, -(d > tol); float64 [m][3] -> [m]."""
    diff = achieved - desired
    d = np.sqrt((diff[:, 0] * diff[:, 0] + diff[:, 1] * diff[:, 1]) + diff[:, 2] * diff[:, 2])
    return -d if dense else -(d > tol).astype(np.float64)


class Ring(object):
    """Host mirror of ctr_her_ring, AoS [slot][env][...] for readability."""

    def __init__(self, n_envs, capacity, obs_dim):
        n, c, D = n_envs, capacity, obs_dim
        self.n, self.capacity, self.D = n, c, D
        self.obs, self.next_obs = np.zeros((c, n, D)), np.zeros((c, n, D))
        self.achieved, self.next_achieved, self.desired = np.zeros((c, n, 3)), np.zeros((c, n, 3)), np.zeros((c, n, 3))
        self.action = np.zeros((c, n, 6), dtype=np.float32)
        self.reward, self.done = np.zeros((c, n)), np.zeros((c, n), dtype=np.uint8)
        self.t_in_ep, self.ep_len = np.zeros((c, n), dtype=np.int32), np.zeros((c, n), dtype=np.int32)
        self.cur_t = np.zeros(n, dtype=np.int32)
        self.stored, self.newest = 0, -1

    def begin(self, obs, achieved, desired):
        self.cur_obs, self.cur_achieved, self.cur_desired = obs.copy(), achieved.copy(), desired.copy()
        self.cur_t[:] = 0

    def store(self, action, obs, achieved, desired, reward, done, terminal_obs, terminal_achieved, auto_reset=True):
        """Arguments are what the env holds AFTER the step (obs/achieved/desired already reset where done, auto_reset)."""
        s = (self.newest + 1) % self.capacity
        term = (done != 0) & bool(auto_reset)
        self.obs[s], self.achieved[s], self.desired[s] = self.cur_obs, self.cur_achieved, self.cur_desired
        self.next_obs[s] = np.where(term[:, None], terminal_obs, obs)
        self.next_achieved[s] = np.where(term[:, None], terminal_achieved, achieved)
        self.action[s], self.reward[s], self.done[s] = action, reward, (done != 0)
        t = self.cur_t + 1
        self.t_in_ep[s] = t
        self.ep_len[s] = 0
        for e in np.nonzero(done)[0]:
            for k in range(min(int(t[e]), self.capacity)):
                self.ep_len[(s - k) % self.capacity, e] = t[e]
        self.cur_t = np.where(done != 0, 0, t).astype(np.int32)
        self.cur_obs, self.cur_achieved, self.cur_desired = obs.copy(), achieved.copy(), desired.copy()
        self.newest, self.stored = s, min(self.stored + 1, self.capacity)

    def sample(self, m, n_goal, tol, seed, call, rewrite_error=False, dense=False):
        i = np.arange(m, dtype=np.uint64)
        ctr = np.stack([i & MASK, i >> np.uint64(32), np.full(m, call & 0xFFFFFFFF, dtype=np.uint64),
                        np.full(m, (call >> 32) & 0xFFFFFFFF, dtype=np.uint64)], axis=1).astype(np.uint32)
        u = philox4x32(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).astype(np.uint64)
        age = (u[:, 0] * np.uint64(self.stored)) >> np.uint64(32)
        slot = ((self.newest - age.astype(np.int64)) % self.capacity).astype(np.int64)
        env = ((u[:, 1] * np.uint64(self.n)) >> np.uint64(32)).astype(np.int64)
        k = self.t_in_ep[slot, env].astype(np.int64)
        L = self.ep_len[slot, env].astype(np.int64)
        L = np.where(L == 0, self.cur_t[env], L)
        relabel = (k < L) & (((u[:, 2] * np.uint64(n_goal + 1)) >> np.uint64(32)) < np.uint64(n_goal))
        off = ((u[:, 3] * np.maximum(L - k, 0).astype(np.uint64)) >> np.uint64(32)).astype(np.int64)
        gslot = (slot + off) % self.capacity
        goal = np.where(relabel[:, None], self.next_achieved[gslot, env], self.desired[slot, env])
        ach, nach = self.achieved[slot, env], self.next_achieved[slot, env]
        obs, nobs = self.obs[slot, env].copy(), self.next_obs[slot, env].copy()
        if rewrite_error:
            obs[relabel, 9:12] = (goal - ach)[relabel]
            nobs[relabel, 9:12] = (goal - nach)[relabel]
        reward = np.where(relabel, compute_reward(nach, goal, tol, dense), self.reward[slot, env])
        done = np.where(relabel, 0, self.done[slot, env]).astype(np.uint8)
        return dict(obs=obs, next_obs=nobs, achieved=ach, next_achieved=nach, desired=goal, action=self.action[slot, env],
                    reward=reward, done=done, src=slot * self.n + env, goal_src=np.where(relabel, gslot * self.n + env, -1))
