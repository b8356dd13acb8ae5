#!/usr/bin/env python
"""ctr_fk on 131072 reset configurations of system 0: time in the given order and in orders sorted by candidate keys (what a
work-sorted hand-out order would be worth for the one-shot FK; the env step sorts by last step's counter instead)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
q = env.buffer('joints').t().to(torch.float64).contiguous()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(qq):
    for _ in range(2):
        env.forward_kinematics(qq)
    torch.cuda.synchronize()
    a.record()
    for _ in range(10):
        env.forward_kinematics(qq)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / 10 * 1e3
tip, R, info = env.forward_kinematics(q, return_info=True)
nfev = info[:, 0].double()
seg = env.segment_tables(q)
nsec = seg[3].double()
length = (seg[1] * (torch.arange(9, device=q.device)[None, :] < seg[3][:, None])).max(dim=1).values
print("given order            %.1f us (mean nfev %.1f)" % (timed(q), nfev.mean().item()))
coarse = -nsec * 16 - torch.floor(16 * length / 0.431).clamp(0, 15) + torch.rand(n, device=q.device, dtype=torch.float64) * 0.5
for name, key in (('coarse key, random ties', coarse), ('nfev desc', -nfev), ('nsec desc', -nsec), ('nsec desc, length desc', -nsec * 10 - length), ('length desc', -length)):
    print("%-22s %.1f us" % (name, timed(q[torch.argsort(key)].contiguous())))
env.close()

if os.environ.get('FK_NEXT'):
    # what k_step's key (LAST step's RHS count) is worth against keys computed from THIS step's joints
    env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
    env.reset()
    lim = torch.tensor(env.action_space.high, device=q.device)
    gen = torch.Generator(device=q.device); gen.manual_seed(3)
    for _ in range(5):
        env.step(((torch.rand((n, 6), device=q.device, generator=gen) * 2 - 1) * lim).contiguous())
    qm = env.buffer('joints').t().to(torch.float64).contiguous()
    nfevm = env.forward_kinematics(qm, return_info=True)[2][:, 0].double()
    env.step(((torch.rand((n, 6), device=q.device, generator=gen) * 2 - 1) * lim).contiguous())
    q0 = env.buffer('joints').t().to(torch.float64).contiguous()
    nfev0 = env.forward_kinematics(q0, return_info=True)[2][:, 0].double()
    env.step(((torch.rand((n, 6), device=q.device, generator=gen) * 2 - 1) * lim).contiguous())
    q1 = env.buffer('joints').t().to(torch.float64).contiguous()
    keep = (env.buffer('t') > 0)                       # envs that were not reset by this step
    q0, q1, nfev0, nfevm = q0[keep], q1[keep], nfev0[keep], nfevm[keep]
    m = (q1.shape[0] // 640) * 640
    q0, q1, nfev0, nfevm = q0[:m], q1[:m].contiguous(), nfev0[:m], nfevm[:m]
    nfev1 = env.forward_kinematics(q1, return_info=True)[2][:, 0].double()
    seg1 = env.segment_tables(q1)
    nsec1 = seg1[3].double()
    nsec0 = env.segment_tables(q0)[3].double()
    len1 = (seg1[1] * (torch.arange(9, device=q.device)[None, :] < seg1[3][:, None])).max(dim=1).values
    print("next-step experiment on %d envs: corr(nfev0, nfev1) = %.3f, mean nfev %.1f" % (m, torch.corrcoef(torch.stack([nfev0, nfev1]))[0, 1].item(), nfev1.mean().item()))
    def windowed(key):                                 # sort inside windows of 640 like k_step
        idx = torch.argsort(key.view(-1, 640), dim=1) + (torch.arange(m // 640, device=q.device) * 640)[:, None]
        return q1[idx.view(-1)].contiguous()
    import os as _os
    for name, key in (("unsorted", None), ("last step's nfev (k_step's key), windows of 640", -nfev0), ("this step's nfev (ceiling), windows of 640", -nfev1),
                      ("mean of the last two steps' nfev", -(nfev0 + nfevm)), ("0.75 last + 0.25 the one before", -(3 * nfev0 + nfevm)),
                      ("last nfev + 12 x section-count change", -(nfev0 + 12 * (nsec1 - nsec0))), ("this step's (nsec, length)", -nsec1 * 10 - len1)):
        qq = q1 if key is None else windowed(key)
        print("    %-52s %.1f us" % (name, timed(qq)))
    env.close()
