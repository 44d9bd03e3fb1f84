"""Host-side logic and the C-ABI surface, CPU only (no compute entry point is called without a GPU).

  * libremixt_b200.so loads and exports every symbol include/remixt_b200.h declares;
  * the state-space enumerators and the segment remap of the host driver (remixt_b200/cn_model.py)
    against the reference's own cn_model.py, imported in place where /root/reference exists;
  * the host driver's EM loop (scipy M-steps, global RNG consumption, accept/reject) driving the
    compiled UNMODIFIED reference RemixtModel reproduces the reference's full fit bit for bit;
  * solution selection.
"""
import os
import re

import numpy as np
import pytest

from tests import util

ROOT = util.ROOT


def test_cabi_library_exports_every_declared_symbol():
    from remixt_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'remixt_b200.h')).read()
    declared = set(re.findall(r'^(?:int|const char\*)\s+(rmx_\w+)\s*\(', header, flags=re.M))
    assert len(declared) >= 30
    assert declared == set(_lib.EXPORTS), declared.symmetric_difference(_lib.EXPORTS)
    for name in declared:
        assert hasattr(_lib.lib, name), name


def test_cabi_fails_loudly_without_gpu():
    """No CPU fallback: on a machine without a CUDA device creating a model raises."""
    from remixt_b200 import _lib
    if _lib.device_count() > 0:
        pytest.skip('a CUDA device is present')
    exp = util.load_experiment('small')
    with pytest.raises(RuntimeError):
        util.make_driver(exp, util.init_params('small', 0))


@pytest.mark.parametrize('M,cn_max,S,B', [(2, 6, 16, 7), (3, 6, 97, 19), (3, 8, 165, 25), (4, 6, 457, 43)])
def test_state_space_sizes(M, cn_max, S, B):
    """Sizes the reference's enumerators produce (probe)."""
    from remixt_b200 import cn_model
    assert cn_model.create_cn_states(M, 2, cn_max, 1).shape == (S, M, 2)
    assert cn_model.create_brk_states(M, cn_max, 1).shape == (B, M)


@pytest.fixture(scope='module')
def ref_cn_model():
    from oracle import refenv
    if not refenv.available():
        pytest.skip('/root/reference not present (GPU box)')
    from oracle import build_ref
    if build_ref.build(verbose=False) is None:
        pytest.skip('reference extension could not be built')
    refenv.install()
    import remixt.cn_model as rcn
    return rcn


def test_state_tables_match_reference(ref_cn_model):
    from remixt_b200 import cn_model
    ref = ref_cn_model.BreakpointModel.__new__(ref_cn_model.BreakpointModel)
    for M, cn_max in ((2, 6), (3, 6), (3, 8)):
        assert np.array_equal(cn_model.create_cn_states(M, 2, cn_max, 1), ref.create_cn_states(M, 2, cn_max, 1))
        assert np.array_equal(cn_model.create_brk_states(M, cn_max, 1), ref.create_brk_states(M, cn_max, 1))


@pytest.mark.parametrize('name', ['small', 'config1', 'genome5k'])
def test_segment_remap_matches_reference(ref_cn_model, name):
    """Dummy-segment insertion (cn_model.py:82-167) and the likelihood masks (:169-184)."""
    from remixt_b200 import cn_model
    exp = util.load_experiment(name)
    ip = util.init_params(name, 0)
    kw = dict(max_copy_number=6, normal_contamination=True, divergence_weight=ip['divergence_weight'],
              min_segment_length=10000, min_proportion_genotyped=0.01, max_depth=ip['max_depth'],
              normal_copies=np.array([[1, 1]] * exp.l.shape[0]))
    a = cn_model.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, **kw)
    b = ref_cn_model.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, **kw)
    assert a.N1 == b.N1
    for attr in ('seg_fwd_remap', 'seg_rev_remap', 'is_telomere', 'breakpoint_idx', 'breakpoint_orient', 'x1', 'l1',
                 '_total_likelihood_mask', '_allele_likelihood_mask'):
        assert np.array_equal(np.asarray(getattr(a, attr)), np.asarray(getattr(b, attr))), attr
    assert list(a.breakpoint_ids) == list(b.breakpoint_ids)


@pytest.mark.slow
def test_host_grid_driver_full_fits_on_reference_model_are_bit_exact(ref_bpmodel):
    """This repo's grid driver (pipeline.fit_grid_batched: lock-step M-step, per-fit RandomState) on the compiled
    reference RemixtModel == the reference's own free-running pipeline.fit per initialisation (golden
    grid_small.npz): the host side adds no numerical difference of its own, and the initialisation the reference
    fails on (L-BFGS-B line search, cn_model.py:520) fails here too and leaves the grid."""
    from remixt_b200 import pipeline
    g = util.golden('grid_small.npz')
    exp = util.load_experiment('small')
    failed = int(g['failed_ids'][0])
    ids = [1, failed]
    inits = dict((i, util.init_params('small', i)) for i in ids)
    cfg = {'max_copy_number': 6, 'num_em_iter': int(g['num_em_iter']), 'num_update_iter': int(g['num_update_iter'])}
    res, best, errors = pipeline.fit_grid_batched(exp, inits, cfg, random_seed=int(g['seed']),
                                                  model_class=ref_bpmodel.RemixtModel)
    assert list(errors) == [failed] and 'optimization failed' in str(errors[failed])
    assert list(res) == [1] and best == 1
    k = list(g['init_ids']).index(1)
    r = res[1]
    assert r['stats']['elbo'] == float(g['elbo'][k])
    assert np.array_equal(r['h'], g['h'][k])
    assert np.array_equal(r['cn'], g['cn'][k])
    assert np.array_equal(np.array([r['brk_cn'][str(i)] for i in g['brk_ids']]), g['brk_cn'][k])
    assert [r['stats'][n] for n in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')] == list(g['params'][k])
    assert [bool(d[1]) for d in r['decisions']] == [bool(a) for a in g['accept'][k].ravel()]
    assert r['stats']['proportion_divergent'] == float(g['proportion_divergent'][k])


def test_sub_batches_merge_on_reference_model(ref_bpmodel):
    """pipeline.fit_grid_batched(sub_batches=K): interleaved groups on their own host threads, merged in the caller's
    order with one selection over all statistics -- on the compiled reference model (no GPU): every fit as in one batch,
    the failing initialisation reported, `best` the same."""
    from remixt_b200 import pipeline
    g = util.golden('grid_small.npz')
    exp = util.load_experiment('small')
    failed = int(g['failed_ids'][0])
    ids = [1, failed, 5]
    inits = dict((i, util.init_params('small', i)) for i in ids)
    cfg = {'max_copy_number': 6, 'num_em_iter': 1, 'num_update_iter': 2}
    one = pipeline.fit_grid_batched(exp, inits, cfg, random_seed=int(g['seed']), model_class=ref_bpmodel.RemixtModel,
                                    sub_batches=1)
    prof = {}
    two = pipeline.fit_grid_batched(exp, inits, cfg, random_seed=int(g['seed']), model_class=ref_bpmodel.RemixtModel,
                                    sub_batches=2, profile=prof)
    assert prof['sub_batches'] == 2
    assert list(one[0]) == list(two[0]) and one[1] == two[1] and sorted(one[2]) == sorted(two[2])
    for i in one[0]:
        assert one[0][i]['stats'] == two[0][i]['stats']
        assert np.array_equal(one[0][i]['cn'], two[0][i]['cn']) and np.array_equal(one[0][i]['h'], two[0][i]['h'])
        assert one[0][i]['decisions'] == two[0][i]['decisions']


def test_select_solution_rule():
    """analysis/pipeline.py:253-259: best ELBO among proportion_divergent < 0.5, else among all."""
    from remixt_b200 import pipeline
    stats = {0: dict(elbo=-10., proportion_divergent=0.6), 1: dict(elbo=-20., proportion_divergent=0.1),
             2: dict(elbo=-15., proportion_divergent=0.2), 3: dict(elbo=-15., proportion_divergent=0.3)}
    assert pipeline.select_solution(stats) == 2
    assert pipeline.select_solution({0: stats[0]}) == 0
    assert pipeline.shard_ids(range(7), 1, 3) == [1, 4]


def test_decode_breakpoints_naive():
    from remixt_b200 import cn_model
    cn = np.array([[[1, 1], [2, 1]], [[1, 1], [1, 1]], [[1, 1], [2, 1]]])
    adj = [(0, 1), (1, 2)]
    brk = {'a': frozenset([(0, 1), (2, 0)])}
    out = cn_model.decode_breakpoints_naive(cn, adj, brk)
    assert np.array_equal(out['a'], [0, 1])


def test_decode_breakpoints_grid_equals_per_fit_decode():
    """pipeline.decode_breakpoints_grid (all fits at once) == BreakpointModel.decode_breakpoints per fit."""
    from oracle import bpmodel_np
    from remixt_b200 import pipeline
    exp = util.load_experiment('small')
    d = util.make_driver(exp, util.init_params('small', 0), model_class=bpmodel_np.RemixtModel)
    rng = np.random.RandomState(5)
    cn = rng.randint(0, 4, size=(4, d.N1, 3, 2)).astype(np.int64)
    cn[:, :, 0, :] = 1
    got = pipeline.decode_breakpoints_grid(d, cn, np.asarray(d.model.brk_states), d.model.transition_penalty)
    for g in range(4):
        want = d.decode_breakpoints(cn[g])
        assert list(got[g]) == list(want)
        assert all(np.array_equal(got[g][k], want[k]) for k in want)
