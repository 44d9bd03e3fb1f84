"""Generate golden vectors with the UNMODIFIED reference (this container only).

TEST INFRASTRUCTURE ONLY.  Nothing under ``remixt_b200/`` may import this.

For an experiment fixture ``tests/golden/exp_NAME.npz`` (made by ``oracle/make_inputs.py``):

1. ``init``:  runs the reference ``remixt.analysis.pipeline.init`` (``pipeline.py:12-109``) on the
   fixture and stores the initialisation grid INTO the fixture (keys ``init_*``), so that the
   GPU box, where the reference Python cannot travel, fits the same grid.
2. ``steps``: drives the compiled reference ``RemixtModel`` through the reference
   ``cn_model.BreakpointModel`` for one round of the five coordinate updates and records every
   public quantity after each call -> ``tests/golden/steps_NAME.npz``.
3. ``fit``:   a complete reference ``pipeline.fit`` (seeded) -> ``tests/golden/fit_NAME.npz``.

Usage: python oracle/make_golden.py NAME [init] [steps] [fit] [--em E --vi V] [--init-id I] [--tag T] [--h-scale F]

``--h-scale 1.0000000000001 --tag _perturbed`` re-runs the reference fit from an initial h_normal moved by one
part in 1e13: the spread between that fit and the unperturbed one is the reference's OWN sensitivity to
rounding-level noise (its Nelder-Mead / L-BFGS-B M-steps amplify it), which calibrates full-fit parity.
"""
import contextlib
import io
import os
import pickle
import sys
import tempfile
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refenv  # noqa: E402

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


class ArrayExperimentRef(object):
    """Stand-in for the reference Experiment with exactly the attributes ``pipeline.init`` /
    ``pipeline.fit`` read.  start/end only feed the unused ``high_quality`` column."""

    def __init__(self, x, l, adjacencies, breakpoints, breakpoint_ids, chromosome):
        self.x = x
        self.l = l
        self.adjacencies = set((int(a), int(b)) for a, b in adjacencies)
        self.breakpoints = dict()
        for bp_id, (n1, s1, n2, s2) in zip(breakpoint_ids, breakpoints):
            self.breakpoints[str(bp_id)] = frozenset([(int(n1), int(s1)), (int(n2), int(s2))])
        self.segment_chromosome_id = chromosome
        self.segment_start = np.zeros(l.shape[0], dtype=int)
        self.segment_end = np.maximum(l.astype(int) - 1, 0)
        self.segment_major_is_allele_a = np.ones(l.shape[0], dtype=int)


def load_experiment(name):
    d = np.load(os.path.join(GOLDEN, 'exp_%s.npz' % name), allow_pickle=False)
    return d, ArrayExperimentRef(d['x'], d['l'], d['adjacencies'], d['breakpoints'], d['breakpoint_ids'],
                                 d['chromosome'])


class _DictStore(dict):
    """Context-manager dict standing in for pd.HDFStore in pipeline.init (PyTables is absent)."""
    def __init__(self, *a, **k):
        dict.__init__(self)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def config_for(name):
    cfg = {'max_copy_number': 6}
    if name in ('config2', 'genome5k'):
        cfg['max_ploidy'] = 4.0
    return cfg


def run_init(name):
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import pandas as pd
        import remixt.analysis.pipeline as rp
    d, exp = load_experiment(name)
    with tempfile.TemporaryDirectory() as tmp:
        exp_fn = os.path.join(tmp, 'exp.pickle')
        with open(exp_fn, 'wb') as f:
            pickle.dump(exp, f)
        real = pd.HDFStore
        pd.HDFStore = _DictStore
        try:
            cfg = config_for(name)
            try:
                init = rp.init(os.path.join(tmp, 'init.h5'), exp_fn, cfg)
            except ValueError as e:
                print('init with', cfg, 'failed:', e, '-> retrying with max_ploidy 4.0')
                cfg['max_ploidy'] = 4.0
                init = rp.init(os.path.join(tmp, 'init.h5'), exp_fn, cfg)
        finally:
            pd.HDFStore = real
    ids = sorted(init.keys())
    out = dict(d)
    for key, dtype in (('mode_idx', np.int64), ('h_normal', float), ('h_tumour', float), ('mix_frac', float),
                       ('divergence_weight', float), ('max_depth', float)):
        out['init_' + key] = np.array([init[i][key] for i in ids], dtype=dtype)
    out['init_max_ploidy'] = np.array(cfg.get('max_ploidy', 6.0))
    np.savez_compressed(os.path.join(GOLDEN, 'exp_%s.npz' % name), **out)
    print(name, 'init grid:', len(ids), 'initialisations; max_depth', out['init_max_depth'][0])


def init_params_of(d, i):
    return dict(mode_idx=int(d['init_mode_idx'][i]), h_normal=float(d['init_h_normal'][i]),
                h_tumour=float(d['init_h_tumour'][i]), mix_frac=float(d['init_mix_frac'][i]),
                divergence_weight=float(d['init_divergence_weight'][i]), max_depth=float(d['init_max_depth'][i]))


def _snapshot(model, out, tag, big=True):
    out[tag + '/elbo'] = model.calculate_elbo()
    out[tag + '/energy'] = model.calculate_variational_energy()
    out[tag + '/entropy'] = model.calculate_variational_entropy()
    out[tag + '/hmm_log_norm_const'] = model.hmm_log_norm_const
    for a in ('p_allele_swap', 'p_outlier_total', 'p_outlier_allele', 'p_breakpoint'):
        out[tag + '/' + a] = np.asarray(getattr(model, a)).copy()
    if big:
        out[tag + '/posterior_marginals'] = np.asarray(model.posterior_marginals).copy()
        out[tag + '/framelogprob'] = np.asarray(model.framelogprob).copy()


def run_steps(name, init_id=0):
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import remixt.analysis.pipeline as rp  # noqa: F401
        import remixt.cn_model as rcn
    d, exp = load_experiment(name)
    ip = init_params_of(d, init_id)
    h_init = np.array([ip['h_normal'], ip['h_tumour'] * ip['mix_frac'], ip['h_tumour'] * (1. - ip['mix_frac'])])
    bm = rcn.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, max_copy_number=6,
                             normal_contamination=True, divergence_weight=ip['divergence_weight'],
                             min_segment_length=10000, min_proportion_genotyped=0.01, max_depth=ip['max_depth'],
                             normal_copies=np.array([[1, 1]] * exp.l.shape[0]))
    bm.num_em_iter = 0
    with contextlib.redirect_stdout(io.StringIO()):
        bm.fit(h_init)
    model = bm.model
    out = {'init_id': init_id, 'h_init': h_init}
    _snapshot(model, out, 's0')
    for r in range(2):
        for k, step in enumerate(('update_p_allele_swap', 'update_p_cn', 'update_p_breakpoint',
                                  'update_p_outlier_total', 'update_p_outlier_allele')):
            getattr(model, step)()
            _snapshot(model, out, 'r%d_%d' % (r, k), big=(step == 'update_p_cn'))
    np.random.seed(4321)
    sample = bm._create_sample()
    out['sample'] = sample
    out['ell_sample'] = model.calculate_expected_log_likelihood(sample)
    out['ell_all'] = model.calculate_expected_log_likelihood(np.ones(model.num_segments, dtype=int))
    g = np.zeros(3)
    model.calculate_expected_log_likelihood_partial_h(sample, g)
    out['grad_sample'] = g.copy()
    model.calculate_expected_log_likelihood_partial_h(np.ones(model.num_segments, dtype=int), g)
    out['grad_all'] = g.copy()
    cn = np.zeros((model.num_segments, 3, 2), dtype=int)
    model.infer_cn(cn)
    out['cn'] = cn
    np.savez_compressed(os.path.join(GOLDEN, 'steps_%s.npz' % name), **out)
    print(name, 'steps: elbo', out['s0/elbo'], '->', out['r1_4/elbo'])


def run_fit(name, init_id=0, em=5, vi=5, tag=None, h_scale=1.):
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import remixt.analysis.pipeline as rp
    d, exp = load_experiment(name)
    ip = init_params_of(d, init_id)
    cfg = {'max_copy_number': 6, 'num_em_iter': em, 'num_update_iter': vi}
    # record the M-step trajectory (h and dispersion parameters after every EM iteration, ELBO before
    # every h update) so that the E-step path can be replayed on it (tests: teacher-forced fit)
    import remixt.cn_model as rcn
    trace = {'h': [], 'params': [], 'elbo_pre_h': []}
    orig_h, orig_params = rcn.BreakpointModel.em_update_h, rcn.BreakpointModel.em_update_params

    def em_update_h(self):
        trace['elbo_pre_h'].append(self.model.calculate_elbo())
        orig_h(self)
        trace['h'].append(np.asarray(self.model.h).copy())

    def em_update_params(self):
        orig_params(self)
        trace['params'].append([getattr(self.model, k) for k in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')])

    rcn.BreakpointModel.em_update_h, rcn.BreakpointModel.em_update_params = em_update_h, em_update_params
    if h_scale != 1.:
        ip = dict(ip)
        ip['h_normal'] *= h_scale
    np.random.seed(1234)
    t0 = time.time()
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            res = rp.fit(exp, ip, cfg)
    finally:
        rcn.BreakpointModel.em_update_h, rcn.BreakpointModel.em_update_params = orig_h, orig_params
    dt = time.time() - t0
    out = {'init_id': init_id, 'num_em_iter': em, 'num_update_iter': vi, 'seconds': dt,
           'h': res['h'], 'cn': res['cn'], 'p_outlier_total': res['p_outlier_total'],
           'p_outlier_allele': res['p_outlier_allele'], 'elbo': res['stats']['elbo'],
           'elbo_diff': res['stats']['elbo_diff'], 'ploidy': res['stats']['ploidy'],
           'proportion_divergent': res['stats']['proportion_divergent'],
           'trace_h': np.array(trace['h']), 'trace_params': np.array(trace['params']),
           'trace_elbo_pre_h': np.array(trace['elbo_pre_h']), 'h_scale': h_scale}
    for k in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1'):
        out[k] = res['stats'][k]
    ids = sorted(res['brk_cn'].keys())
    out['brk_ids'] = np.array(ids)
    out['brk_cn'] = np.array([res['brk_cn'][i] for i in ids])
    np.savez_compressed(os.path.join(GOLDEN, 'fit_%s%s.npz' % (name, tag or '')), **out)
    print(name, 'fit init', init_id, ': %.1f s, elbo %.10f, h %s' % (dt, out['elbo'], res['h']))


def main():
    args = sys.argv[1:]
    name = args[0]
    em = int(args[args.index('--em') + 1]) if '--em' in args else 5
    vi = int(args[args.index('--vi') + 1]) if '--vi' in args else 5
    init_id = int(args[args.index('--init-id') + 1]) if '--init-id' in args else 0
    tag = args[args.index('--tag') + 1] if '--tag' in args else None
    if 'init' in args:
        run_init(name)
    if 'steps' in args:
        run_steps(name, init_id)
    h_scale = float(args[args.index('--h-scale') + 1]) if '--h-scale' in args else 1.
    if 'fit' in args:
        run_fit(name, init_id, em, vi, tag, h_scale)




# ---------------------------------------------------------------------------------------------------------
# grid: complete FREE-RUNNING reference fits of every initialisation of a fixture's grid, and the init_id the
# reference's own selection rule (analysis/pipeline.py:253-259) picks.  python oracle/make_golden.py NAME grid
#   [--procs P] [--seed S] [--tag T] [--h-scale F] [--inits a,b,c]
# ---------------------------------------------------------------------------------------------------------
PARAMS4 = ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')


def _grid_fit_one(job):
    name, init_id, em, vi, seed, h_scale = job[:6]
    extra = job[6] if len(job) > 6 else {}
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import remixt.analysis.pipeline as rp
        import remixt.cn_model as rcn
    d, exp = load_experiment(name)
    ip = init_params_of(d, init_id)
    if h_scale != 1.:
        ip['h_normal'] *= h_scale
    cfg = dict(config_for(name))
    cfg.update(extra)
    cfg.update({'num_em_iter': em, 'num_update_iter': vi})
    trace = {'h': [], 'params': [], 'elbo': [], 'accept': []}
    orig_h, orig_p, orig_up = rcn.BreakpointModel.update_h, rcn.BreakpointModel.update_param, \
        rcn.BreakpointModel.em_update_params

    # accept / reject of cn_model.py:526-531 and :564-569, read off the model state: a rejected step restores the
    # value held before the step (an accepted step that lands exactly on the old value is indistinguishable and
    # equivalent)
    def update_h(self):
        before = np.asarray(self.model.h).copy()
        orig_h(self)
        trace['accept'].append(not np.array_equal(before, np.asarray(self.model.h)))
        trace['h'].append(np.asarray(self.model.h).copy())

    def update_param(self, pname):
        before = getattr(self.model, pname)
        orig_p(self, pname)
        trace['accept'].append(before != getattr(self.model, pname))

    def em_update_params(self):
        orig_up(self)
        trace['params'].append([getattr(self.model, k) for k in PARAMS4])
        trace['elbo'].append(self.model.calculate_elbo())

    rcn.BreakpointModel.update_h, rcn.BreakpointModel.update_param = update_h, update_param
    rcn.BreakpointModel.em_update_params = em_update_params
    np.random.seed(seed)
    t0 = time.time()
    err = ''
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            res = rp.fit(exp, ip, cfg)
    except Exception as e:      # the reference's own workflow records a failed init and goes on
        err = repr(e)
        res = None
    finally:
        rcn.BreakpointModel.update_h, rcn.BreakpointModel.update_param = orig_h, orig_p
        rcn.BreakpointModel.em_update_params = orig_up
    out = {'init_id': init_id, 'seconds': time.time() - t0, 'error': err}
    if res is not None:
        st = res['stats']
        ids = sorted(res['brk_cn'].keys())
        out.update(elbo=st['elbo'], elbo_diff=st['elbo_diff'], h=np.asarray(res['h']),
                   params=np.array([st[k] for k in PARAMS4]), ploidy=st['ploidy'],
                   proportion_divergent=st['proportion_divergent'], cn=np.asarray(res['cn']).astype(np.int8),
                   brk_ids=np.array(ids), brk_cn=np.array([res['brk_cn'][i] for i in ids]).astype(np.int8),
                   trace_h=np.array(trace['h']), trace_params=np.array(trace['params']),
                   trace_elbo=np.array(trace['elbo']), accept=np.array(trace['accept']).reshape(em, -1),
                   p_outlier_total=np.asarray(res['p_outlier_total'])[:, 1].astype(np.float32),
                   p_outlier_allele=np.asarray(res['p_outlier_allele'])[:, 1].astype(np.float32))
    return out


class _SelectStore(object):
    """Stand-in for the HDF store of store_optimal_solution: records which solution it copies."""
    def __init__(self):
        self.read = []

    def __getitem__(self, key):
        self.read.append(key)
        return None

    def __setitem__(self, key, value):
        pass


def reference_selection(stats_rows, config):
    """init_id the reference's own store_optimal_solution picks (analysis/pipeline.py:253-264)."""
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import pandas as pd
        import remixt.analysis.pipeline as rp
    store = _SelectStore()
    rp.store_optimal_solution(pd.DataFrame(stats_rows), store, config)
    key = store.read[0]          # '/solutions/solution_{id}/cn'
    return int(key.split('solution_')[1].split('/')[0])


def run_grid(name, procs=6, em=5, vi=5, seed=1234, tag='', h_scale=1., inits=None, extra=None):
    import multiprocessing as mp
    d, _ = load_experiment(name)
    n = len(d['init_h_normal'])
    ids = list(range(n)) if inits is None else inits
    jobs = [(name, i, em, vi, seed, h_scale, dict(extra or {})) for i in ids]
    t0 = time.time()
    with mp.get_context('fork').Pool(procs) as pool:
        fits = pool.map(_grid_fit_one, jobs, chunksize=1)
    ok = [f for f in fits if not f['error']]
    rows = [dict(init_id=f['init_id'], elbo=f['elbo'], proportion_divergent=f['proportion_divergent']) for f in ok]
    best = reference_selection(rows, dict(config_for(name), **(extra or {})))
    out = {'init_ids': np.array([f['init_id'] for f in ok]), 'failed_ids': np.array([f['init_id'] for f in fits if f['error']], dtype=np.int64),
           'best_init_id': best, 'seed': seed, 'num_em_iter': em, 'num_update_iter': vi, 'h_scale': h_scale,
           'seconds': np.array([f['seconds'] for f in ok]), 'brk_ids': ok[0]['brk_ids'],
           'max_copy_number': int(dict(config_for(name), **(extra or {}))['max_copy_number'])}
    for k in ('elbo', 'elbo_diff', 'h', 'params', 'ploidy', 'proportion_divergent', 'cn', 'brk_cn', 'trace_h',
              'trace_params', 'trace_elbo', 'accept', 'p_outlier_total', 'p_outlier_allele'):
        out[k] = np.array([f[k] for f in ok])
    np.savez_compressed(os.path.join(GOLDEN, 'grid_%s%s.npz' % (name, tag)), **out)
    print(name, 'grid: %d fits (%d failed) in %.0f s wall, %.0f s cpu; best init %d, elbo %.6f' % (
        len(ok), len(fits) - len(ok), time.time() - t0, out['seconds'].sum(), best, out['elbo'][list(out['init_ids']).index(best)]))


def _main_grid():
    args = sys.argv[1:]
    def opt(flag, default, conv):
        return conv(args[args.index(flag) + 1]) if flag in args else default
    run_grid(args[0], procs=opt('--procs', 6, int), em=opt('--em', 5, int), vi=opt('--vi', 5, int),
             seed=opt('--seed', 1234, int), tag=opt('--tag', '', str), h_scale=opt('--h-scale', 1., float),
             inits=opt('--inits', None, lambda s: [int(v) for v in s.split(',')]),
             extra=({'max_copy_number': opt('--cnmax', 6, int)} if '--cnmax' in args else None))


if __name__ == '__main__':
    if 'grid' in sys.argv[2:]:
        _main_grid()
    else:
        main()
