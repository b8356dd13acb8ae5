#!/usr/bin/env python
"""Times ctr_bvp_pass (32768 configurations of system 0, no load, u_init = 0) for the library named by $CTR_B200_LIB."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
q = env.buffer('joints').t().to(torch.float64).contiguous()
for _ in range(2):
    out = env.bvp_pass(q)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    out = env.bvp_pass(q)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
st = out['steps'].double()
print("%-28s bvp_pass %d configs: %.3f ms = %.3e passes/s, mean steps %.1f max %d, tip checksum %.12f" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), n, ms, n / ms * 1e3, st.mean().item(), int(st.max().item()),
    out['tip'].sum().item()))
g = st[: n // 10 * 10].view(-1, 10)
print("    lock-step: sum of per-warp max steps x 10 / sum of steps = %.3f" % ((g.max(dim=1).values.sum() * 10) / g.sum()).item())
env.close()
if os.environ.get('BVP_LPT'):
    # upper bound of a longest-first pass order: the same configurations sorted by their measured step count / by total length
    env = CtrReachVecEnv(16, device=0, **ctr_reach_envs.default_kwargs())
    Lt = torch.tensor([0.431, 0.332, 0.174], dtype=torch.float64, device=q.device)
    Lct = torch.tensor([0.103, 0.113, 0.134], dtype=torch.float64, device=q.device)
    bt = q[:, :3]
    dtip_t = Lt + bt
    pts = torch.cat([torch.zeros(n, 1, dtype=torch.float64, device=q.device), bt, dtip_t - Lct, dtip_t], dim=1).sort(dim=1).values
    lens = 1e-5 * torch.floor(1e5 * (pts[:, 1:] - pts[:, :-1]))
    Scum = lens.cumsum(dim=1) + bt.min(dim=1).values[:, None]
    nsec = ((lens != 0) & (Scum > 0)).sum(dim=1).double()
    length = dtip_t.max(dim=1).values
    for name, key in (('steps desc', -st), ('length desc', -length), ('nsec asc', nsec), ('nsec asc, length desc', nsec * 10 - length),
                      ('random', torch.rand(n, device=q.device, dtype=torch.float64))):
        qs = q[torch.argsort(key)].contiguous()
        for _ in range(2):
            env.bvp_pass(qs)
        torch.cuda.synchronize()
        a.record()
        for _ in range(5):
            env.bvp_pass(qs)
        b.record()
        torch.cuda.synchronize()
        print("    order %-12s %.3f ms" % (name, a.elapsed_time(b) / 5))
    env.close()
if os.environ.get('BVP_JAC'):
    env = CtrReachVecEnv(16, device=0, **ctr_reach_envs.default_kwargs())
    qj = q[:32768].contiguous()
    for _ in range(2):
        J, _, tipj = env.bvp_jacobian(qj, compliance=False)
    torch.cuda.synchronize()
    a.record()
    for _ in range(5):
        J, _, tipj = env.bvp_jacobian(qj, compliance=False)
    b.record()
    torch.cuda.synchronize()
    print("    bvp_jacobian 32768 configs (7 passes each): %.3f ms, J checksum %.10f" % (a.elapsed_time(b) / 5, J.sum().item()))
    env.close()
