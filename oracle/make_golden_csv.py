"""tests/golden/evaluations_sample.csv: the header and first 25 rows of one evaluation file the reference ships
(saved_policies/**/evaluations*.csv, written by src/rotation_evaluation.py:68-81), copied verbatim.  Build container only."""
import glob
import os

HERE = os.path.dirname(os.path.abspath(__file__))
src = sorted(glob.glob('/root/reference/ctr_reach_envs/saved_policies/**/evaluations*.csv', recursive=True))[0]
lines = open(src).read().splitlines(keepends=True)[:26]
open(os.path.join(HERE, '..', 'tests', 'golden', 'evaluations_sample.csv'), 'w').write(''.join(lines))
print(src, len(lines))
