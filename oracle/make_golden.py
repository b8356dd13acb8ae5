"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE.  Run once here (`python oracle/make_golden.py`); the outputs are committed
because /root/reference does not exist on the GPU box.  Every array comes from the reference's
own code (imported through oracle/ref_loader.py) or from the (joint -> tip) vectors it ships in
saved_policies/**/evaluations*.csv.
"""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_loader  # noqa: E402

OUT = os.path.join(HERE, '..', 'tests', 'golden')


def main():
    pkg, gym = ref_loader.load()
    from gym import spaces as gspaces
    import ctr_reach_envs.envs.model as ref_model
    from ctr_reach_envs.envs.CTR_Python import Segment
    from ctr_reach_envs.envs.goal_tolerance import GoalTolerance
    import pandas as pd

    os.makedirs(OUT, exist_ok=True)

    # ---- capture the final ODE state (R is integrated but dropped by model.py:70) and nfev
    cap = {}
    orig = ref_model.solve_ivp

    def spy(*a, **k):
        sol = orig(*a, **k)
        cap['y'] = sol.y[:, -1].copy()
        cap['nfev'] = cap.get('nfev', 0) + sol.nfev
        cap['nsec'] = cap.get('nsec', 0) + 1
        return sol

    ref_model.solve_ivp = spy

    env = gym.make('CTR-Reach-v0', select_systems=[0, 1, 2, 3])
    model, obs_obj = env.model, env.trig_obj
    tube_L = np.array(obs_obj.tube_lengths)

    def ref_fk(q, system):
        cap.clear()
        tip = model.forward_kinematics(q, system)
        return tip.copy(), cap['y'][9:18].copy(), cap['nfev'], cap['nsec']

    # ---- (i) FK fixtures: sampled joints (float32, as the env produces) + edge sets, all 4 systems
    rng = np.random.RandomState(1234)
    gspaces.seed(1234)
    Q, SYS, TIP, ROT, NFEV, NSEC = [], [], [], [], [], []
    for system in range(4):
        qs = [obs_obj.sample_goal(system) for _ in range(400)]
        L = tube_L[system]
        lo = (-L + 1e-3).astype(np.float32)
        # edge sets: fully extended, fully retracted (nested limit), ties, large |alpha|
        edge = []
        for alpha in ([0, 0, 0], [3.1, -2.0, 1.0], [150.0, -149.0, 77.7], [-3.14159, 3.14159, 0.5]):
            edge.append(np.concatenate((np.zeros(3), alpha)))
            edge.append(np.concatenate((lo, alpha)))
            edge.append(np.concatenate(([lo[2], lo[2], lo[2]], alpha)))          # all-equal betas
            edge.append(np.concatenate(([lo[1], lo[1], -0.01], alpha)))          # tie b0 == b1
            edge.append(np.concatenate(([lo[0], -0.02, -0.02], alpha)))          # tie b1 == b2
            edge.append(np.concatenate(([-0.05, -0.03, -1e-4], alpha)))
            edge.append(np.concatenate(([-1e-5, -1e-5, 0.0], alpha)))
            edge.append(np.concatenate(([-2.5e-4, -1e-4, -3e-5], alpha)))
        qs += [np.asarray(e, dtype=np.float32) for e in edge]
        for q in qs:
            q = np.asarray(q, dtype=np.float32)
            try:
                tip, R, nfev, nsec = ref_fk(q, system)
            except Exception as exc:  # reference raises on some degenerate sets; skip those
                print('reference raised on', system, q, type(exc).__name__)
                continue
            Q.append(q), SYS.append(system), TIP.append(tip), ROT.append(R), NFEV.append(nfev), NSEC.append(nsec)
    np.savez_compressed(os.path.join(OUT, 'fk_reference.npz'), joints=np.array(Q, dtype=np.float32),
                        system=np.array(SYS, dtype=np.int32), tip=np.array(TIP), R=np.array(ROT),
                        nfev=np.array(NFEV, dtype=np.int32), nsec=np.array(NSEC, dtype=np.int32))
    print('fk_reference:', len(Q))

    # ---- (i-b) segmentation tables for a subset (padded to 6 columns)
    SQ, SS, ST = [], [], []
    for q, system in list(zip(Q, SYS))[::7]:
        t = model.current_sys_parameters[system]
        seg = Segment(t[0], t[1], t[2], q[0:3] + np.zeros(3, dtype=np.int64))
        m = len(seg.S)
        tab = np.zeros((1 + 4 * 3, 6))
        tab[0, :m] = seg.S
        tab[1:4, :m], tab[4:7, :m], tab[7:10, :m], tab[10:13, :m] = seg.EI, seg.GJ, seg.U_x, seg.U_y
        SQ.append(q), SS.append(system), ST.append(np.concatenate(([m], tab.ravel())))
    np.savez_compressed(os.path.join(OUT, 'segment_reference.npz'), joints=np.array(SQ, dtype=np.float32),
                        system=np.array(SS, dtype=np.int32), table=np.array(ST))
    print('segment_reference:', len(SQ))

    # ---- (ii) set_action traces: random + limit-pushing action sequences, float32 joints
    lim = env.action_space.high.astype(np.float32)
    J0, ACT, J1, TSYS, TCA = [], [], [], [], []
    for constrain_alpha in (False, True):
        e2 = gym.make('CTR-Reach-v0', select_systems=[0, 1, 2, 3], constrain_alpha=constrain_alpha)
        o2 = e2.trig_obj
        for system in range(4):
            for mode in range(6):
                o2.joints = o2.sample_goal(system)
                bias = np.zeros(6, dtype=np.float32)
                if mode == 1:
                    bias[:3] = lim[:3]          # push towards full extension
                if mode == 2:
                    bias[:3] = -lim[:3]         # push towards full retraction
                if mode == 3:
                    bias[:3] = [lim[0], -lim[1], lim[2]]
                if mode == 4:
                    bias[:3] = [-lim[0], lim[1], -lim[2]]
                if mode == 5:
                    bias[3:] = lim[3:]          # wind alpha up (hits +-pi when constrained)
                for _ in range(120):
                    a = rng.uniform(-lim, lim).astype(np.float32)
                    a = np.clip(a * np.float32(0.3) + bias, -lim, lim).astype(np.float32) if mode else a
                    j0 = o2.joints.copy()
                    for _s in range(10):
                        o2.set_action(a, system)
                    assert o2.joints.dtype == np.float32
                    J0.append(j0), ACT.append(a), J1.append(o2.joints.copy()), TSYS.append(system)
                    TCA.append(constrain_alpha)
    np.savez_compressed(os.path.join(OUT, 'set_action_reference.npz'), joints_before=np.array(J0), action=np.array(ACT),
                        joints_after=np.array(J1), system=np.array(TSYS, dtype=np.int32),
                        constrain_alpha=np.array(TCA), n_substeps=10, tube_lengths=tube_L)
    print('set_action_reference:', len(J0))

    # ---- (iii) full env rollouts (reset + random steps) through the reference's public API
    for name, kwargs in (('rollout_default', {}),
                         ('rollout_multi_prop', dict(select_systems=[0, 1, 2, 3], joint_representation='proprioceptive',
                                                     constrain_alpha=True))):
        e3 = gym.make('CTR-Reach-v0', **kwargs)
        np.random.seed(7)
        gspaces.seed(7)
        rec = {k: [] for k in ('action', 'joints_before', 'joints_after', 'achieved', 'desired', 'obs', 'reward',
                               'done', 'success', 'error', 't', 'system', 'tol', 'reset_obs', 'reset_joints',
                               'reset_desired', 'reset_achieved', 'reset_system')}
        tol_sched = [0.02, 0.02, 0.004]
        for ep in range(3):
            e3.goal_tolerance.current_tol = tol_sched[ep]
            o = e3.reset()
            rec['reset_obs'].append(o['observation']), rec['reset_joints'].append(e3.trig_obj.joints.copy())
            rec['reset_desired'].append(o['desired_goal']), rec['reset_achieved'].append(o['achieved_goal'])
            rec['reset_system'].append(e3.system)
            for _ in range(150):
                a = e3.action_space.sample()
                jb = e3.trig_obj.joints.copy()
                o, r, d, info = e3.step(a)
                rec['action'].append(a), rec['joints_before'].append(jb), rec['joints_after'].append(e3.trig_obj.joints.copy())
                rec['achieved'].append(o['achieved_goal']), rec['desired'].append(o['desired_goal'])
                rec['obs'].append(o['observation']), rec['reward'].append(r), rec['done'].append(d)
                rec['success'].append(info['is_success']), rec['error'].append(info['error']), rec['t'].append(e3.t)
                rec['system'].append(e3.system), rec['tol'].append(e3.get_goal_tolerance())
                if d:
                    break
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **{k: np.array(v) for k, v in rec.items()},
                            n_systems=len(e3.ctr_system_parameters),
                            representation=kwargs.get('joint_representation', 'egocentric'),
                            constrain_alpha=kwargs.get('constrain_alpha', False))
        print(name, len(rec['action']), 'steps')

    # ---- (iv) reward / flag truth table incl. d == tol and t == 150 (ctr_reach_env.py:139-140,155,168-170)
    ach = rng.uniform(-0.1, 0.1, size=(64, 3))
    des = ach + rng.uniform(-0.03, 0.03, size=(64, 3))
    des[0] = ach[0]
    des[1] = ach[1] + np.array([0.02, 0.0, 0.0])          # distance within an ulp of tol
    des[2] = ach[2] + np.array([0.0, 0.0, 0.001])
    rows = []
    for tol in (0.02, 0.001, float(np.linalg.norm(ach[1] - des[1]))):
        e3.goal_tolerance.current_tol = tol
        rw = e3.compute_reward(ach, des, {})
        err = np.linalg.norm(des - ach, axis=-1)
        rows.append(dict(tol=tol, reward=rw, success=err < tol, error=err))
    np.savez_compressed(os.path.join(OUT, 'reward_reference.npz'), achieved=ach, desired=des,
                        tol=np.array([r['tol'] for r in rows]), reward=np.array([r['reward'] for r in rows]),
                        success=np.array([r['success'] for r in rows]), error=np.array([r['error'] for r in rows]))

    # ---- (v) tolerance schedules (goal_tolerance.py)
    ts = np.array([0, 1, 10, 1000, 99999, 100000, 200000, 200001, 1500000, 1500001, 3000000])
    sched = {}
    for fn in ('constant', 'linear', 'decay'):
        for n_ts in (200000, 1500000):
            g = GoalTolerance({'inc_tol_obs': False, 'final_tol': 0.001, 'initial_tol': 0.020, 'N_ts': n_ts,
                               'function': fn, 'set_tol': 0})
            vals = [g.get_tol()]
            for t in ts:
                g.update(int(t))
                vals.append(g.get_tol())
            sched['%s_%d' % (fn, n_ts)] = np.array(vals, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, 'tolerance_reference.npz'), timesteps=ts, **sched)

    # ---- (vi) the reference's shipped (joint -> tip) vectors: 300 rows x 3 pairs per usable CSV
    root = os.path.join(ref_loader.REF_ROOT, 'ctr_reach_envs', 'saved_policies')
    files = sorted(glob.glob(os.path.join(root, '**', 'evaluations*.csv'), recursive=True))
    CQ, CS, CT, CF = [], [], [], []
    names = []
    for f in files:
        if f.endswith('tro_four_systems_sample/evaluations_2.csv'):
            continue  # system_id column mis-mapped in this file
        df = pd.read_csv(f, float_precision='round_trip').iloc[:300]
        names.append(os.path.relpath(f, root))
        for kind, pos in (('desired', 'desired_goal'), ('achieved', 'achieved_goal'), ('starting', 'starting_position')):
            q = np.column_stack([df['B_%s_%d' % (kind, i)] for i in (1, 2, 3)] +
                                [df['alpha_%s_%d' % (kind, i)] for i in (1, 2, 3)])
            tip = np.column_stack([df['%s_%s' % (pos, c)] for c in 'xyz'])
            CQ.append(q), CT.append(tip), CS.append(df['system_id'].to_numpy().astype(np.int32))
            CF.append(np.full(len(df), len(names) - 1, dtype=np.int32))
    np.savez_compressed(os.path.join(OUT, 'csv_goldens.npz'), joints=np.concatenate(CQ), tip=np.concatenate(CT),
                        system=np.concatenate(CS), file_id=np.concatenate(CF), files=np.array(names))
    print('csv_goldens:', sum(len(x) for x in CQ), 'vectors from', len(names), 'files')

    # ---- (vii) goal-sampling acceptance statistics per system (obs.py:185-207)
    acc = []
    for system in range(4):
        sp = obs_obj.joint_sample_spaces[system]
        n_ok = 0
        for _ in range(20000):
            s = sp.sample()
            b = s[:3]
            ok = all((b[i - 1] <= b[i]) and (b[i - 1] + tube_L[system][i - 1] >= tube_L[system][i] + b[i]) for i in (1, 2))
            n_ok += ok
        acc.append(n_ok / 20000.0)
    np.savez_compressed(os.path.join(OUT, 'goal_sampling_reference.npz'), acceptance=np.array(acc), tube_lengths=tube_L)
    print('acceptance', acc)


if __name__ == '__main__':
    main()
