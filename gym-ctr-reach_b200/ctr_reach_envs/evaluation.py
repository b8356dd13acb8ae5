"""Evaluation files of the reference (src/rotation_evaluation.py:21-81): one row per finished episode with the goal, the reached
tip, the episode length, the starting position, the desired / achieved / starting joints and the system id -- the format of the
saved_policies/**/evaluations*.csv files its analysis scripts (src/*plot*.py, src/rotation_evaluation.py) read.

read_evaluations / write_evaluations round-trip those files; evaluate() fills one from a CtrReachVecEnv with all environments
stepping in parallel (the reference runs one episode after the other).
"""
import numpy as np
import pandas as pd

EVAL_COLUMNS = ['desired_goal_x', 'desired_goal_y', 'desired_goal_z', 'achieved_goal_x', 'achieved_goal_y', 'achieved_goal_z',
                'episode_length', 'starting_position_x', 'starting_position_y', 'starting_position_z',
                'B_desired_1', 'B_desired_2', 'B_desired_3', 'B_achieved_1', 'B_achieved_2', 'B_achieved_3',
                'B_starting_1', 'B_starting_2', 'B_starting_3', 'alpha_desired_1', 'alpha_desired_2', 'alpha_desired_3',
                'alpha_achieved_1', 'alpha_achieved_2', 'alpha_achieved_3', 'alpha_starting_1', 'alpha_starting_2',
                'alpha_starting_3', 'system_id']          # column order of rotation_evaluation.py:69-80


def read_evaluations(path):
    """-> DataFrame with EVAL_COLUMNS; float_precision='round_trip' so that every value is the double the writer held."""
    df = pd.read_csv(path, index_col=0, float_precision='round_trip')
    if list(df.columns) != EVAL_COLUMNS:
        raise ValueError("not an evaluation file: columns differ from rotation_evaluation.py:69-80")
    return df


def make_frame(desired_goals, achieved_goals, episode_lengths, starting_positions, q_desired, q_achieved, q_starting, system_id):
    """np.column_stack in the order of rotation_evaluation.py:66-68 (everything becomes float64, as there)."""
    qd, qa, qs = (np.asarray(q, dtype=np.float64).reshape(-1, 6) for q in (q_desired, q_achieved, q_starting))
    data = np.column_stack((np.asarray(desired_goals, dtype=np.float64), np.asarray(achieved_goals, dtype=np.float64),
                            np.asarray(episode_lengths, dtype=np.float64), np.asarray(starting_positions, dtype=np.float64),
                            qd[:, :3], qa[:, :3], qs[:, :3], qd[:, 3:], qa[:, 3:], qs[:, 3:], np.asarray(system_id, dtype=np.float64)))
    return pd.DataFrame(data=data, columns=EVAL_COLUMNS)


def write_evaluations(path, df):
    if list(df.columns) != EVAL_COLUMNS:
        raise ValueError("columns differ from rotation_evaluation.py:69-80")
    df.to_csv(path)                                       # rotation_evaluation.py:81


def evaluate(env, policy, num_episodes, goal=None, system=None, max_steps=None):
    """Run `num_episodes` episodes on a CtrReachVecEnv built with auto_reset=False, env.num_envs at a time, and return the
    evaluation frame.  policy(obs_dict) -> float32 CUDA actions (N, 6).  An episode's row is taken when it finishes
    (rotation_evaluation.py:47-64: `if done or infos.get('is_success')`), from the same fields the reference reads from `infos`."""
    import torch
    if env.auto_reset:
        raise ValueError("evaluate() needs auto_reset=False: the row of a finished episode is read before it is reset")
    N = env.num_envs
    max_steps = max_steps or env.max_steps_per_episode
    rows = {k: [] for k in ('dg', 'ag', 'len', 'sp', 'qd', 'qa', 'qs', 'sys')}
    got = 0
    while got < num_episodes:
        obs = env.reset(goal=goal, system=system)
        alive = torch.ones(N, dtype=torch.bool, device=env.device)
        for _ in range(max_steps):
            obs, reward, done, info = env.step(policy(obs))
            fin = alive & (done.bool() | info['is_success'].bool())
            if bool(fin.any()):
                idx = fin.nonzero().flatten()
                b = env.buffer
                rows['dg'].append(b('desired_goal').t()[idx].cpu().numpy())
                rows['ag'].append(b('achieved_goal').t()[idx].cpu().numpy())
                rows['len'].append(b('t')[idx].cpu().numpy())
                rows['sp'].append(b('starting_position').t()[idx].cpu().numpy())
                rows['qd'].append(b('desired_joints').t()[idx].cpu().numpy())
                rows['qa'].append(b('joints').t()[idx].cpu().numpy())
                rows['qs'].append(b('starting_joints').t()[idx].cpu().numpy())
                rows['sys'].append(b('system')[idx].cpu().numpy())
                alive &= ~fin
            if not bool(alive.any()):
                break
        got = sum(len(x) for x in rows['len'])
    cat = {k: np.concatenate(v)[:num_episodes] for k, v in rows.items()}
    return make_frame(cat['dg'], cat['ag'], cat['len'], cat['sp'], cat['qd'], cat['qa'], cat['qs'], cat['sys'])
