"""Shared helpers of the test-suite (test infrastructure; may use oracle/)."""
import contextlib
import io
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def load_experiment(name):
    from remixt_b200 import pipeline
    return pipeline.ArrayExperiment.from_npz(os.path.join(GOLDEN, 'exp_%s.npz' % name))


def init_params(name, i):
    d = golden('exp_%s.npz' % name)
    return dict(mode_idx=int(d['init_mode_idx'][i]), h_normal=float(d['init_h_normal'][i]),
                h_tumour=float(d['init_h_tumour'][i]), mix_frac=float(d['init_mix_frac'][i]),
                divergence_weight=float(d['init_divergence_weight'][i]), max_depth=float(d['init_max_depth'][i]))


def num_inits(name):
    return len(golden('exp_%s.npz' % name)['init_h_normal'])


def make_driver(exp, ip, model_class=None, **kw):
    """BreakpointModel (this repo's host driver) with an un-fitted RemixtModel attached."""
    from remixt_b200 import cn_model, pipeline
    args = dict(max_copy_number=6, normal_contamination=True, divergence_weight=ip['divergence_weight'],
                min_segment_length=10000, min_proportion_genotyped=0.01, max_depth=ip['max_depth'],
                normal_copies=np.array([[1, 1]] * exp.l.shape[0]), model_class=model_class)
    args.update(kw)
    bm = cn_model.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, **args)
    bm.num_em_iter = 0
    with contextlib.redirect_stdout(io.StringIO()):
        bm.fit(pipeline.init_h(ip))
    return bm


def rel_err(a, b):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.


def assert_close(a, b, rtol=1e-6, atol=0., what=''):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    bad = ~np.isclose(a, b, rtol=rtol, atol=atol)
    if bad.any():
        i = np.argmax(np.abs(a - b) * bad)
        raise AssertionError('%s: %d of %d differ beyond rtol=%g atol=%g; worst a=%r b=%r' % (
            what, int(bad.sum()), a.size, rtol, atol, a.flat[i], b.flat[i]))


STEPS = ('update_p_allele_swap', 'update_p_cn', 'update_p_breakpoint', 'update_p_outlier_total',
         'update_p_outlier_allele')
