"""Parity of the CUDA path (through the C-ABI) with the reference, on the B200.

Three kinds of checks, in decreasing strength:
  * live: the compiled UNMODIFIED reference extension (oracle/_ref, travels to the GPU box as a
    built .so) and the CUDA model are driven side by side on the same inputs;
  * golden: committed vectors the reference produced in the build container
    (oracle/make_golden.py -> tests/golden/steps_*.npz, fit_*.npz);
  * properties at sizes the reference cannot run in test time.

Tolerances: FP64 quantities 1e-6 relative as BASELINE.json's north_star states; a tighter 1e-7 is
asserted where nothing chaotic sits in between.  The floor is libm: CUDA's and glibc's lgamma
each carry a few ulp of error on lgamma(1e5) ~ 1e6, i.e. ~1e-9 absolute on a log-likelihood, which
shows up as ~1e-9 relative on the posteriors (CPython's math.lgamma differs from glibc's by the
same amount, see oracle/bpmodel_np.py).  MAP copy number, breakpoint copy number and Viterbi
paths bit-exact.
"""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

RTOL = 1e-6       # the north_star tolerance
RTOL_TIGHT = 1e-7


@pytest.fixture(scope='module')
def cuda_bpmodel():
    from remixt_b200 import bpmodel, _lib
    assert _lib.device_count() > 0, 'no CUDA device: the GPU tests must not silently skip'
    return bpmodel


def _assert_posterior(pc, pr, tag, rtol=RTOL_TIGHT):
    util.assert_close(pc, pr, rtol=max(rtol, RTOL), atol=1e-200, what=tag + ' posterior_marginals')
    util.assert_close(np.where(pr > 1e-12, pc, 0.), np.where(pr > 1e-12, pr, 0.), rtol=rtol,
                      what=tag + ' posterior_marginals (mass)')


def _compare_state(cu, ref, tag, rtol=RTOL_TIGHT):
    for a in ('p_allele_swap', 'p_outlier_total', 'p_outlier_allele', 'p_breakpoint'):
        util.assert_close(getattr(cu, a), np.asarray(getattr(ref, a)), rtol=rtol, atol=1e-280, what=tag + ' ' + a)
    # posteriors: the north_star bar everywhere; the tight bar where there is posterior mass.  An entry of 1e-38
    # is exp(-87) relative to its segment's best state and inherits the rounding of ~20 likelihood terms of 1e7 ulp
    # scale along its path (tests/test_lltable_gpu.py::test_table_vs_exact: device and reference arithmetic each
    # carry ~5e-16 of the largest intermediate), i.e. ~1e-7 relative between two correct FP64 evaluations.
    _assert_posterior(cu.posterior_marginals, np.asarray(ref.posterior_marginals), tag, rtol)
    util.assert_close(cu.framelogprob, np.asarray(ref.framelogprob), rtol=rtol, what=tag + ' framelogprob')
    util.assert_close(cu.hmm_log_norm_const, ref.hmm_log_norm_const, rtol=rtol, what=tag + ' hmm_log_norm_const')
    for f in ('calculate_variational_energy', 'calculate_variational_entropy', 'calculate_elbo'):
        util.assert_close(getattr(cu, f)(), getattr(ref, f)(), rtol=rtol, what=tag + ' ' + f)


@pytest.mark.parametrize('name,init_id', [('small', 1), ('small', 7), ('config1', 1)])
def test_live_coordinate_updates(cuda_bpmodel, ref_bpmodel, name, init_id):
    """Every coordinate update, side by side with the compiled reference."""
    exp = util.load_experiment(name)
    ip = util.init_params(name, init_id)
    cu = util.make_driver(exp, ip).model
    ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel).model
    assert isinstance(cu, cuda_bpmodel.RemixtModel)
    _compare_state(cu, ref, 'ctor')
    # the constructor's cached transition expectation is bit-exact (same p_breakpoint, same order)
    assert np.array_equal(cu.cached_log_transmat, np.asarray(ref.cached_log_transmat))
    for r in range(2):
        for step in util.STEPS:
            getattr(cu, step)()
            getattr(ref, step)()
            _compare_state(cu, ref, 'round %d %s' % (r, step))
    util.assert_close(cu.log_transmat, np.asarray(ref.log_transmat), rtol=RTOL_TIGHT, atol=1e-12, what='log_transmat')
    util.assert_close(cu.cached_log_transmat, np.asarray(ref.cached_log_transmat), rtol=RTOL_TIGHT, atol=1e-12,
                      what='cached_log_transmat')
    _assert_posterior(cu.joint_posterior_marginals, np.asarray(ref.joint_posterior_marginals), 'joint')
    # M-step objective and gradient, subsample and full data
    rs = np.random.RandomState(7)
    sample = np.zeros(cu.num_segments, dtype=int)
    sample[rs.choice(cu.num_segments, size=min(200, cu.num_segments // 10), replace=False)] = 1
    for s in (sample, np.ones(cu.num_segments, dtype=int)):
        util.assert_close(cu.calculate_expected_log_likelihood(s), ref.calculate_expected_log_likelihood(s),
                          rtol=RTOL_TIGHT, what='E[LL]')
        g1, g2 = np.zeros(3), np.zeros(3)
        cu.calculate_expected_log_likelihood_partial_h(s, g1)
        ref.calculate_expected_log_likelihood_partial_h(s, g2)
        util.assert_close(g1, g2, rtol=RTOL_TIGHT, what='dE[LL]/dh')
    # changed parameters are seen immediately (cn_model.py:544 relies on it)
    for m in (cu, ref):
        m.negbin_r_0 = 123.
        m.betabin_M_1 = 3.5
        m.h = np.array([0.03, 0.05, 0.01])
    util.assert_close(cu.calculate_expected_log_likelihood(sample), ref.calculate_expected_log_likelihood(sample),
                      rtol=RTOL_TIGHT, what='E[LL] after setattr')
    util.assert_close(cu.calculate_elbo(), ref.calculate_elbo(), rtol=RTOL_TIGHT, what='elbo after setattr')
    # MAP decode: bit-exact
    c1 = np.zeros((cu.num_segments, 3, 2), dtype=int)
    c2 = np.zeros((cu.num_segments, 3, 2), dtype=int)
    cu.infer_cn(c1)
    ref.infer_cn(c2)
    assert np.array_equal(c1, c2)


@pytest.mark.parametrize('name', ['small', 'config1'])
def test_golden_steps(cuda_bpmodel, name):
    """Same sequence against the committed reference vectors (no reference code at run time)."""
    g = util.golden('steps_%s.npz' % name)
    exp = util.load_experiment(name)
    ip = util.init_params(name, int(g['init_id']))
    cu = util.make_driver(exp, ip).model

    def check(tag, big):
        for a in ('p_allele_swap', 'p_outlier_total', 'p_outlier_allele', 'p_breakpoint'):
            util.assert_close(getattr(cu, a), g[tag + '/' + a], rtol=RTOL_TIGHT, atol=1e-280, what=tag + ' ' + a)
        util.assert_close(cu.calculate_elbo(), g[tag + '/elbo'], rtol=RTOL_TIGHT, what=tag + ' elbo')
        util.assert_close(cu.calculate_variational_energy(), g[tag + '/energy'], rtol=RTOL_TIGHT, what=tag + ' energy')
        util.assert_close(cu.calculate_variational_entropy(), g[tag + '/entropy'], rtol=RTOL_TIGHT,
                          what=tag + ' entropy')
        util.assert_close(cu.hmm_log_norm_const, g[tag + '/hmm_log_norm_const'], rtol=RTOL_TIGHT, what=tag + ' logZ')
        if big:
            _assert_posterior(cu.posterior_marginals, g[tag + '/posterior_marginals'], tag)
            util.assert_close(cu.framelogprob, g[tag + '/framelogprob'], rtol=RTOL_TIGHT, what=tag + ' framelogprob')

    check('s0', True)
    for r in range(2):
        for k, step in enumerate(util.STEPS):
            getattr(cu, step)()
            check('r%d_%d' % (r, k), step == 'update_p_cn')
    sample = g['sample']
    util.assert_close(cu.calculate_expected_log_likelihood(sample), g['ell_sample'], rtol=RTOL_TIGHT, what='ell')
    util.assert_close(cu.calculate_expected_log_likelihood(np.ones_like(sample)), g['ell_all'], rtol=RTOL_TIGHT,
                      what='ell all')
    grad = np.zeros(3)
    cu.calculate_expected_log_likelihood_partial_h(sample, grad)
    util.assert_close(grad, g['grad_sample'], rtol=RTOL_TIGHT, what='grad')
    cu.calculate_expected_log_likelihood_partial_h(np.ones_like(sample), grad)
    util.assert_close(grad, g['grad_all'], rtol=RTOL_TIGHT, what='grad all')
    cn = np.zeros((cu.num_segments, 3, 2), dtype=int)
    cu.infer_cn(cn)
    assert np.array_equal(cn, g['cn'])


def _replay_driver(exp, ip, cfg, trace_h, trace_params, elbo_pre_h):
    """A host driver whose M-steps are REPLAYED from the reference's recorded trajectory (teacher forcing):
    the E-steps, ELBOs and the decode run on the GPU on exactly the parameters the reference had."""
    from remixt_b200 import pipeline
    bm = pipeline.build_model(exp, ip, cfg)
    state = {'i': 0, 'elbo_pre_h': []}

    def em_update_h():
        state['elbo_pre_h'].append(bm.model.calculate_elbo())
        bm.model.h = trace_h[state['i']]

    def em_update_params():
        for k, name in enumerate(('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')):
            setattr(bm.model, name, trace_params[state['i']][k])
        state['i'] += 1

    bm.em_update_h, bm.em_update_params = em_update_h, em_update_params
    return bm, state


@pytest.mark.parametrize('name', ['small', 'config1'])
def test_golden_full_fit_teacher_forced(cuda_bpmodel, name):
    """A complete fit (5 EM x 5 VI = 25 variational rounds, 5 parameter changes, MAP decode) on the reference's
    own parameter trajectory: ELBO before every M-step and at the end within 1e-7, MAP segment and
    breakpoint copy number bit-exact."""
    from remixt_b200 import pipeline
    g = util.golden('fit_%s.npz' % name)
    exp = util.load_experiment(name)
    ip = util.init_params(name, int(g['init_id']))
    cfg = {'max_copy_number': 6, 'num_em_iter': int(g['num_em_iter']), 'num_update_iter': int(g['num_update_iter'])}
    bm, state = _replay_driver(exp, ip, cfg, g['trace_h'], g['trace_params'], g['trace_elbo_pre_h'])
    bm.fit(pipeline.init_h(ip))
    util.assert_close(np.array(state['elbo_pre_h']), g['trace_elbo_pre_h'], rtol=RTOL_TIGHT, what='elbo before M-steps')
    util.assert_close(bm.prev_elbo, g['elbo'], rtol=RTOL_TIGHT, what='final elbo')
    cn, brk_cn = bm.optimal_cn()
    assert np.array_equal(cn, g['cn']), 'MAP copy number must be bit-exact'
    ids = [str(i) for i in g['brk_ids']]
    assert np.array_equal(np.array([brk_cn[i] for i in ids]), g['brk_cn']), 'breakpoint copy number'
    util.assert_close(bm.p_outlier_total, g['p_outlier_total'], rtol=RTOL, atol=1e-12, what='p_outlier_total')
    util.assert_close(bm.p_outlier_allele, g['p_outlier_allele'], rtol=RTOL, atol=1e-12, what='p_outlier_allele')


@pytest.mark.parametrize('name', ['small', 'config1'])
def test_full_fit_free_running(cuda_bpmodel, name):
    """The same fit with its own M-steps (scipy L-BFGS-B / brute + Nelder-Mead on the GPU objective).

    The reference's M-step is ill-conditioned: Nelder-Mead shrinks its simplex to 1e-4 in x on an objective
    that is flat to ~1e-11 there, so the digits of the dispersions beyond ~1e-5 relative are decided by
    rounding noise, and the EM trajectory amplifies them (Parity).  What is asserted is what is
    stable: the first M-step (identical state in) reproduces h to 1e-7 and the dispersions to the optimiser's
    own tolerance, and the free-running fit ends within the spread the reference shows against ITSELF when its
    input is moved by one part in 1e13 (fit_*_perturbed.npz), widened by a safety factor."""
    from remixt_b200 import pipeline
    g = util.golden('fit_%s.npz' % name)
    gp = util.golden('fit_%s_perturbed.npz' % name)
    exp = util.load_experiment(name)
    ip = util.init_params(name, int(g['init_id']))
    cfg = {'max_copy_number': 6, 'num_em_iter': int(g['num_em_iter']), 'num_update_iter': int(g['num_update_iter'])}
    # first EM iteration: the states entering the M-step agree to ~1e-12
    np.random.seed(1234)
    cfg1 = dict(cfg, num_em_iter=1)
    res1 = pipeline.fit(exp, ip, cfg1)
    util.assert_close(res1['h'], g['trace_h'][0], rtol=RTOL, what='h after the first M-step')
    p1 = np.array([res1['stats'][k] for k in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')])
    util.assert_close(p1, g['trace_params'][0], rtol=1e-4, what='dispersions after the first M-step (fmin xtol=1e-4)')
    # whole fit
    np.random.seed(1234)
    res = pipeline.fit(exp, ip, cfg)
    ref_spread = abs(float(gp['elbo']) - float(g['elbo'])) / abs(float(g['elbo']))
    tol = max(5e-3, 100. * ref_spread)
    util.assert_close(res['stats']['elbo'], g['elbo'], rtol=tol, what='final elbo (free running)')
    agree = np.mean(np.all(res['cn'] == g['cn'], axis=(1, 2)))
    assert agree >= 0.9, 'MAP copy number agreement %.3f' % agree
    print('%s free-running fit: elbo %.6f vs reference %.6f (rel %.2e; reference self-spread %.2e), MAP agreement %.4f'
          % (name, res['stats']['elbo'], float(g['elbo']),
             abs(res['stats']['elbo'] - float(g['elbo'])) / abs(float(g['elbo'])), ref_spread, agree))


def test_live_no_normal_contamination_male_x(cuda_bpmodel, ref_bpmodel):
    """hdel / LOH special-parameter branches (bpmodel.pyx:760-766, :818-834) and a germline
    (1,0) chromosome: normal_contamination=False zeroes the normal copies (cn_model.py:79-80)."""
    exp = util.load_experiment('small')
    ip = util.init_params('small', 1)
    cu = util.make_driver(exp, ip, normal_contamination=False).model
    ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel, normal_contamination=False).model
    _compare_state(cu, ref, 'nc=False ctor')
    for step in util.STEPS:
        getattr(cu, step)()
        getattr(ref, step)()
        _compare_state(cu, ref, 'nc=False ' + step)
    ones = np.ones(cu.num_segments, dtype=int)
    for nm, val in (('negbin_hdel_mu', 5e-6), ('negbin_hdel_r_1', 3.), ('betabin_loh_p', 2e-3), ('betabin_loh_M_0', 40.)):
        setattr(cu, nm, val)
        setattr(ref, nm, val)
        util.assert_close(cu.calculate_expected_log_likelihood(ones), ref.calculate_expected_log_likelihood(ones),
                          rtol=RTOL_TIGHT, what=nm)
    g1, g2 = np.zeros(3), np.zeros(3)
    cu.calculate_expected_log_likelihood_partial_h(ones, g1)
    ref.calculate_expected_log_likelihood_partial_h(ones, g2)
    util.assert_close(g1, g2, rtol=RTOL_TIGHT, what='grad nc=False')
    assert np.array_equal(cu.is_hdel, np.asarray(ref.is_hdel))
    assert np.array_equal(cu.is_loh, np.asarray(ref.is_loh))
    # male X: segments of the last chromosome get germline copies (1, 0) and no allele counts
    exp = util.load_experiment('small')
    on_x = exp.segment_chromosome_id == exp.segment_chromosome_id[-1]
    exp.x = exp.x.copy()
    exp.x[on_x, 0:2] = 0
    nrm = np.array([[1, 1]] * exp.l.shape[0])
    nrm[on_x, :] = [1, 0]
    cu = util.make_driver(exp, ip, normal_copies=nrm).model
    ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel, normal_copies=nrm).model
    assert np.array_equal(cu.cached_log_transmat, np.asarray(ref.cached_log_transmat))
    for step in util.STEPS:
        getattr(cu, step)()
        getattr(ref, step)()
        _compare_state(cu, ref, 'male X ' + step)
    c1 = np.zeros((cu.num_segments, 3, 2), dtype=int)
    c2 = np.zeros((cu.num_segments, 3, 2), dtype=int)
    cu.infer_cn(c1)
    ref.infer_cn(c2)
    assert np.array_equal(c1, c2)


def test_live_transition_model_1_and_no_breakpoints(cuda_bpmodel, ref_bpmodel):
    exp = util.load_experiment('small')
    ip = util.init_params('small', 4)
    for kw in (dict(transition_model=1), dict(disable_breakpoints=True)):
        cu = util.make_driver(exp, ip, **kw).model
        ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel, **kw).model
        for step in util.STEPS:
            getattr(cu, step)()
            getattr(ref, step)()
            _compare_state(cu, ref, '%s %s' % (kw, step))


def test_module_functions(cuda_bpmodel, ref_bpmodel):
    """sum_product / max_product / sum_product_2paramtrans on explicit arrays, and sum_product against
    exhaustive enumeration of all state sequences of a 3-segment chain (the reference's own test pattern,
    remixt/tests/test_model.py:58-93, 207-242)."""
    rs = np.random.RandomState(3)
    for N, S in ((3, 4), (40, 17), (7, 97)):
        F = rs.randn(N, S) * 3.
        T = -np.abs(rs.randn(N - 1, S, S)) * 4.
        a1, b1, a2, b2 = (np.zeros((N, S)) for _ in range(4))
        cuda_bpmodel.sum_product(F, T, a1, b1)
        ref_bpmodel.sum_product(F, T, a2, b2)
        util.assert_close(a1, a2, rtol=1e-12, what='alphas')
        util.assert_close(b1, b2, rtol=1e-12, atol=1e-13, what='betas')
        p1, p2 = np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64)
        l1 = cuda_bpmodel.max_product(F, T, p1)
        l2 = ref_bpmodel.max_product(F, T, p2)
        assert np.array_equal(p1, p2)
        assert l1 == l2
        cuda_bpmodel.sum_product_2paramtrans(F, a1, b1, 0.3, 0.7)
        ref_bpmodel.sum_product_2paramtrans(F, a2, b2, 0.3, 0.7)
        util.assert_close(a1, a2, rtol=1e-12, what='alphas 2p')
        util.assert_close(b1, b2, rtol=1e-12, atol=1e-13, what='betas 2p')
    # exhaustive enumeration, 3 segments x 4 states
    N, S = 3, 4
    F = rs.randn(N, S)
    T = rs.randn(N - 1, S, S)
    a, b = np.zeros((N, S)), np.zeros((N, S))
    cuda_bpmodel.sum_product(F, T, a, b)
    import itertools
    logp = {}
    for seq in itertools.product(range(S), repeat=N):
        v = sum(F[n, seq[n]] for n in range(N)) + sum(T[n, seq[n], seq[n + 1]] for n in range(N - 1))
        logp[seq] = v
    Z = np.log(sum(np.exp(v) for v in logp.values()))
    util.assert_close(np.log(np.exp(a[-1]).sum()), Z, rtol=1e-12, what='norm const')
    for n in range(N):
        marg = np.zeros(S)
        for seq, v in logp.items():
            marg[seq[n]] += np.exp(v - Z)
        post = np.exp(a[n] + b[n] - Z)
        util.assert_close(post, marg, rtol=1e-11, what='marginal %d' % n)


def test_errors(cuda_bpmodel):
    """Same exception classes as the reference's argument checks (bpmodel.pyx:509-529)."""
    exp = util.load_experiment('small')
    ip = util.init_params('small', 1)
    bm = util.make_driver(exp, ip)
    m = bm.model
    N, S = m.num_segments, m.num_cn_states
    args = [3, N, m.num_breakpoints, True, np.asarray(m.cn_states), np.asarray(m.brk_states),
            np.array([0.04, 0.03, 0.02]), bm.l1, np.ascontiguousarray(bm.x1[:, 2]), np.ascontiguousarray(bm.x1[:, 0:2]),
            bm.is_telomere, bm.breakpoint_idx, bm.breakpoint_orient, 10., 1e-7]
    cuda_bpmodel.RemixtModel(*args)
    for idx, bad in ((6, np.array([0.1, 0.1])), (10, bm.is_telomere[:-1]), (11, bm.breakpoint_idx[:-1])):
        a = list(args)
        a[idx] = bad
        with pytest.raises(ValueError):
            cuda_bpmodel.RemixtModel(*a)
    a = list(args)
    a[2] = m.num_breakpoints + 1
    with pytest.raises(ValueError):
        cuda_bpmodel.RemixtModel(*a)
    a = list(args)
    a[4] = np.asarray(m.cn_states).astype(np.int32)
    with pytest.raises(ValueError):
        cuda_bpmodel.RemixtModel(*a)
    # an expected allele ratio of exactly 1 with allele reads present raises (bpmodel.pyx:335)
    cn_states = np.asarray(m.cn_states).copy()
    cn_states[:, :, 0, :] = [1, 0]
    a = list(args)
    a[4] = cn_states
    mm = cuda_bpmodel.RemixtModel(*a)
    with pytest.raises(ValueError):
        mm.update_p_cn()


def test_determinism_and_properties_genome(cuda_bpmodel):
    """Whole-genome-shaped input (23 chromosome chains, ~300 breakpoints): run-to-run bit
    reproducibility, normalisation, ELBO monotone under coordinate ascent (the reference's own
    check, cn_model.py:430-442), and chain independence."""
    exp = util.load_experiment('genome5k')
    ip = util.init_params('genome5k', 1)
    runs = []
    for rep in range(2):
        m = util.make_driver(exp, ip).model
        elbos = [m.calculate_elbo()]
        for r in range(3):
            for step in util.STEPS:
                getattr(m, step)()
                elbos.append(m.calculate_elbo())
        runs.append((np.array(elbos), m.posterior_marginals, m.p_breakpoint, m.p_outlier_total))
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b), 'results must be bit-reproducible run to run'
    elbos, post, pb, pot = runs[0]
    assert np.all(np.diff(elbos[1:]) > -1e-6 * np.abs(elbos[1:-1])), elbos
    util.assert_close(post.sum(axis=1), np.ones(post.shape[0]), rtol=1e-12, what='posterior rows')
    util.assert_close(pb.sum(axis=1), np.ones(pb.shape[0]), rtol=1e-12, what='p_breakpoint rows')
    # the fused loop equals the step-by-step loop bit for bit
    m1 = util.make_driver(exp, ip).model
    m1.variational_update(3)
    assert np.array_equal(m1.posterior_marginals, post)
    assert m1.calculate_elbo() == elbos[-1]


def test_compact_cn_states_equals_full(cuda_bpmodel):
    """``cn_states`` given as its two factors (CompactCnStates) or as the reference's (N1,S,M,2) array: same model."""
    exp = util.load_experiment('small')
    ip = util.init_params('small', 4)
    drv = util.make_driver(exp, ip)
    h = np.asarray(drv.model.h)
    full = drv.model_args(h) + (ip['divergence_weight'],)
    comp = drv.model_args(h, compact=True) + (ip['divergence_weight'],)
    assert isinstance(full[4], np.ndarray) and isinstance(comp[4], cuda_bpmodel.CompactCnStates)
    assert np.array_equal(np.asarray(comp[4]), full[4]) and comp[4].max() == full[4].max()
    assert np.array_equal(np.asarray(comp[4][3:9]), full[4][3:9])
    a, b = cuda_bpmodel.RemixtModel(*full), cuda_bpmodel.RemixtModel(*comp)
    for m in (a, b):
        drv.attach(m)       # likelihood masks (dummy segments have no data)
        m.variational_update(2)
    assert a.calculate_elbo() == b.calculate_elbo()
    assert np.array_equal(a.posterior_marginals, b.posterior_marginals)
    assert np.array_equal(a.is_loh, b.is_loh) and np.array_equal(a.cn_states_total, b.cn_states_total)
