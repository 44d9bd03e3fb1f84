"""How sensitive are the REFERENCE's own free-running fits to rounding-level noise in its M-step objective?

TEST INFRASTRUCTURE ONLY.  The compiled, unmodified reference RemixtModel is wrapped so that every value
calculate_expected_log_likelihood returns is multiplied by (1 + eps * N(0,1)), eps = 1e-15 by default -- a stand-in
for "the same formula evaluated with another libm / another summation order", which is all that separates two correct
FP64 implementations.  Everything else (E-step, gradients, optimisers, random streams) is untouched.  The fits run
through this repo's host driver, which reproduces the reference's own driver bit for bit on the reference model
(tests/test_host.py).  Output: tests/golden/grid_<name>_noise.npz in the format of oracle/make_golden.py grid.

    python oracle/noise_sensitivity.py NAME [--eps 1e-15] [--procs 6]
"""
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PARAMS4 = ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')


def _fit_one(job):
    name, init_id, eps, seed = job
    from oracle import refenv
    from remixt_b200 import pipeline
    from tests import util
    ref = refenv.load_ref_bpmodel()
    noise = np.random.RandomState(10_000 + init_id)

    class Noisy(object):
        def __init__(self, *a, **k):
            object.__setattr__(self, '_m', ref.RemixtModel(*a, **k))

        def __getattr__(self, k):
            return getattr(self._m, k)

        def __setattr__(self, k, v):
            setattr(self._m, k, v)

        def calculate_expected_log_likelihood(self, sample):
            return self._m.calculate_expected_log_likelihood(sample) * (1. + eps * noise.randn())

    exp = util.load_experiment(name)
    ip = util.init_params(name, init_id)
    cfg = {'max_copy_number': 6, 'num_em_iter': 5, 'num_update_iter': 5}
    t0 = time.time()
    try:
        r = pipeline.fit(exp, ip, cfg, model_class=Noisy, random_state=np.random.RandomState(seed))
    except Exception as e:
        return {'init_id': init_id, 'error': repr(e)}
    st = r['stats']
    ids = sorted(r['brk_cn'].keys())
    return {'init_id': init_id, 'error': '', 'seconds': time.time() - t0, 'elbo': st['elbo'], 'h': np.asarray(r['h']),
            'params': np.array([st[k] for k in PARAMS4]), 'ploidy': st['ploidy'],
            'proportion_divergent': st['proportion_divergent'], 'cn': np.asarray(r['cn']).astype(np.int8),
            'brk_cn': np.array([r['brk_cn'][i] for i in ids]).astype(np.int8),
            'accept': np.array([d[1] for d in r['decisions']]).reshape(5, -1)}


def main():
    args = sys.argv[1:]
    name = args[0]
    eps = float(args[args.index('--eps') + 1]) if '--eps' in args else 1e-15
    procs = int(args[args.index('--procs') + 1]) if '--procs' in args else 6
    from oracle import make_golden
    from tests import util
    n = util.num_inits(name)
    with mp.get_context('fork').Pool(procs) as pool:
        fits = pool.map(_fit_one, [(name, i, eps, 1234) for i in range(n)], chunksize=1)
    ok = [f for f in fits if not f['error']]
    rows = [dict(init_id=f['init_id'], elbo=f['elbo'], proportion_divergent=f['proportion_divergent']) for f in ok]
    best = make_golden.reference_selection(rows, make_golden.config_for(name)) if make_golden.refenv.available() else -1
    out = {'init_ids': np.array([f['init_id'] for f in ok]), 'best_init_id': best, 'eps': eps,
           'failed_ids': np.array([f['init_id'] for f in fits if f['error']], dtype=np.int64)}
    for k in ('elbo', 'h', 'params', 'ploidy', 'proportion_divergent', 'cn', 'brk_cn', 'accept'):
        out[k] = np.array([f[k] for f in ok])
    tag = '' if eps == 1e-15 else '_%g' % eps
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'grid_%s_noise%s.npz' % (name, tag)), **out)
    print(name, 'noise %g:' % eps, len(ok), 'fits,', len(fits) - len(ok), 'failed, best', best)


if __name__ == '__main__':
    main()
