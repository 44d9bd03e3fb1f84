"""A small complete grid fit for compute-sanitizer (memcheck / racecheck): 3 initialisations of the 240-segment
fixture, 1 EM iteration of 2 VI rounds, lock-step M-step, MAP decode."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from remixt_b200 import pipeline  # noqa: E402
from tests import util  # noqa: E402

exp = util.load_experiment('small')
inits = dict((i, util.init_params('small', i)) for i in (0, 9, 29))
res, best, errors = pipeline.fit_grid_batched(exp, inits, {'max_copy_number': 6, 'num_em_iter': 1, 'num_update_iter': 2})
print('best', best, 'elbo', [round(r['stats']['elbo'], 6) for r in res.values()], 'errors', errors)
