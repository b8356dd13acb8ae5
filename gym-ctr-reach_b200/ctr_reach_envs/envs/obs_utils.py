"""Joint-representation helpers with the reference's names (envs/obs_utils.py:7-80).  Host-side numpy utilities for
callers that post-process joints; the per-step encoding itself runs inside the CUDA step kernel."""
import numpy as np

NUM_TUBES = 3


def single_joint2trig(joint):
    return np.array([np.cos(joint[1]), np.sin(joint[1]), joint[0]])


def single_trig2joint(trig):
    return np.array([trig[2], np.arctan2(trig[1], trig[0])])


def rep2joint(rep):
    rep = np.asarray(rep).reshape(NUM_TUBES, 3)
    return np.concatenate((rep[:, 2], np.arctan2(rep[:, 1], rep[:, 0])))


def joint2rep(joint):
    joint = np.asarray(joint)
    out = np.array([])
    for beta, alpha in zip(joint[:NUM_TUBES], joint[NUM_TUBES:]):
        out = np.append(out, single_joint2trig(np.array([beta, alpha])))
    return out


def ego2prop(joint):
    joint = np.asarray(joint)
    return np.concatenate((joint[:NUM_TUBES].cumsum(), joint[NUM_TUBES:].cumsum()))


def prop2ego(joint):
    joint = np.asarray(joint)
    return np.concatenate((np.diff(joint[:NUM_TUBES], prepend=0), np.diff(joint[NUM_TUBES:], prepend=0)))
