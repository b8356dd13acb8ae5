#!/usr/bin/env python
"""Stage the reference's own step-path files under oracle/_ref/ so that the UNMODIFIED reference env can be timed as the CPU arm
on the GPU box (where /root/reference does not exist).

TEST INFRASTRUCTURE ONLY.  oracle/_ref/ is git-ignored (no reference source enters the history) but not gpurun-ignored,
so it travels with the snapshot like the built .so files.  The files are byte-for-byte copies; the one numpy >= 1.24
incompatibility (envs/model.py:107, a ragged np.array) is patched at IMPORT time by oracle/ref_loader.py, exactly as when
the reference is loaded from /root/reference in the build container.  `__graft_entry__.build()` runs this when
/root/reference is present.

    python oracle/make_ref.py            # copy (idempotent); prints the staged file list
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get('CTR_REFERENCE_SRC', '/root/reference')
DST = os.path.join(HERE, '_ref')

# the files the step path executes, plus the package __init__ files their imports go through
# (envs/CTR_Python/__init__.py imports CTR_Model, envs/ctr_reach_env.py imports ctr_3d_graph: both only at import time)
FILES = [
    'ctr_reach_envs/__init__.py',
    'ctr_reach_envs/envs/__init__.py',
    'ctr_reach_envs/envs/ctr_reach_env.py',
    'ctr_reach_envs/envs/obs.py',
    'ctr_reach_envs/envs/obs_utils.py',
    'ctr_reach_envs/envs/model.py',
    'ctr_reach_envs/envs/model_utils.py',
    'ctr_reach_envs/envs/goal_tolerance.py',
    'ctr_reach_envs/envs/ctr_3d_graph.py',
    'ctr_reach_envs/envs/CTR_Python/__init__.py',
    'ctr_reach_envs/envs/CTR_Python/Tube.py',
    'ctr_reach_envs/envs/CTR_Python/Segment.py',
    'ctr_reach_envs/envs/CTR_Python/CTR_Model.py',
]


def stage(verbose=True):
    if not os.path.isdir(SRC):
        if verbose:
            print("make_ref: %s not present, nothing staged" % SRC)
        return False
    for rel in FILES:
        s, d = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if not (os.path.exists(d) and filecmp.cmp(s, d, shallow=False)):
            shutil.copyfile(s, d)
    if verbose:
        print("make_ref: %d reference files staged under %s" % (len(FILES), DST))
    return True


if __name__ == '__main__':
    sys.exit(0 if stage() else 1)
