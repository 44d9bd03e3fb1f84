"""The oracle (oracle/bpmodel_np.py, a numpy restatement of remixt/bpmodel.pyx) against
  * the committed golden vectors the compiled UNMODIFIED reference produced (tests/golden/steps_*.npz,
    generator oracle/make_golden.py), and
  * the compiled reference itself (oracle/_ref), side by side, where it has been built.
CPU only.  Tolerance 1e-9 relative on FP64 quantities (summation order differs), Viterbi bit-exact.
"""
import itertools

import numpy as np
import pytest

from oracle import bpmodel_np
from tests import util

RTOL = 1e-9


def _drive_and_check(m, g, rounds):
    def check(tag, big):
        for a in ('p_allele_swap', 'p_outlier_total', 'p_outlier_allele', 'p_breakpoint'):
            util.assert_close(getattr(m, a), g[tag + '/' + a], rtol=RTOL, atol=1e-280, what=tag + ' ' + a)
        util.assert_close(m.calculate_elbo(), g[tag + '/elbo'], rtol=RTOL, what=tag + ' elbo')
        util.assert_close(m.calculate_variational_energy(), g[tag + '/energy'], rtol=RTOL, what=tag + ' energy')
        util.assert_close(m.calculate_variational_entropy(), g[tag + '/entropy'], rtol=RTOL, what=tag + ' entropy')
        util.assert_close(m.hmm_log_norm_const, g[tag + '/hmm_log_norm_const'], rtol=RTOL, what=tag + ' logZ')
        if big:
            util.assert_close(m.posterior_marginals, g[tag + '/posterior_marginals'], rtol=RTOL, atol=1e-200,
                              what=tag + ' posterior')
            util.assert_close(m.framelogprob, g[tag + '/framelogprob'], rtol=RTOL, what=tag + ' framelogprob')
    check('s0', True)
    for r in range(rounds):
        for k, step in enumerate(util.STEPS):
            getattr(m, step)()
            check('r%d_%d' % (r, k), step == 'update_p_cn')


def test_oracle_vs_golden_small():
    g = util.golden('steps_small.npz')
    exp = util.load_experiment('small')
    ip = util.init_params('small', int(g['init_id']))
    m = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel).model
    _drive_and_check(m, g, rounds=2)
    sample = g['sample']
    util.assert_close(m.calculate_expected_log_likelihood(sample), g['ell_sample'], rtol=RTOL, what='ell')
    util.assert_close(m.calculate_expected_log_likelihood(np.ones_like(sample)), g['ell_all'], rtol=RTOL, what='ell all')
    grad = np.zeros(3)
    m.calculate_expected_log_likelihood_partial_h(sample, grad)
    util.assert_close(grad, g['grad_sample'], rtol=RTOL, what='grad')
    m.calculate_expected_log_likelihood_partial_h(np.ones_like(sample), grad)
    util.assert_close(grad, g['grad_all'], rtol=RTOL, what='grad all')
    cn = np.zeros((m.num_segments, 3, 2), dtype=int)
    m.infer_cn(cn)
    assert np.array_equal(cn, g['cn'])


@pytest.mark.slow
def test_oracle_vs_golden_config1_first_round():
    """BASELINE config 1 (1000 segments, one chromosome): constructor state and the first update_p_cn."""
    g = util.golden('steps_config1.npz')
    exp = util.load_experiment('config1')
    ip = util.init_params('config1', int(g['init_id']))
    m = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel).model
    util.assert_close(m.calculate_elbo(), g['s0/elbo'], rtol=RTOL, what='s0 elbo')
    m.update_p_allele_swap()
    util.assert_close(m.p_allele_swap, g['r0_0/p_allele_swap'], rtol=RTOL, what='swap')
    m.update_p_cn()
    util.assert_close(m.posterior_marginals, g['r0_1/posterior_marginals'], rtol=RTOL, atol=1e-200, what='posterior')
    util.assert_close(m.hmm_log_norm_const, g['r0_1/hmm_log_norm_const'], rtol=RTOL, what='logZ')
    m.update_p_breakpoint()
    util.assert_close(m.p_breakpoint, g['r0_2/p_breakpoint'], rtol=RTOL, atol=1e-280, what='p_breakpoint')
    util.assert_close(m.calculate_elbo(), g['r0_2/elbo'], rtol=RTOL, what='elbo')


def test_oracle_vs_compiled_reference_live(ref_bpmodel):
    """Side by side with oracle/_ref on inputs the golden files do not cover: no normal contamination
    (hdel / LOH parameters), transition_model 1, and the module functions."""
    exp = util.load_experiment('small')
    ip = util.init_params('small', 3)
    for kw in (dict(normal_contamination=False), dict(transition_model=1)):
        a = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel, **kw).model
        b = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel, **kw).model
        assert np.array_equal(a.cached_log_transmat, np.asarray(b.cached_log_transmat))
        for step in util.STEPS:
            getattr(a, step)()
            getattr(b, step)()
            util.assert_close(a.calculate_elbo(), b.calculate_elbo(), rtol=RTOL, what='%s %s' % (kw, step))
        util.assert_close(a.posterior_marginals, np.asarray(b.posterior_marginals), rtol=RTOL, atol=1e-200, what='post')
        util.assert_close(a.p_breakpoint, np.asarray(b.p_breakpoint), rtol=RTOL, atol=1e-280, what='p_breakpoint')
        ones = np.ones(a.num_segments, dtype=int)
        g1, g2 = np.zeros(3), np.zeros(3)
        a.calculate_expected_log_likelihood_partial_h(ones, g1)
        b.calculate_expected_log_likelihood_partial_h(ones, g2)
        util.assert_close(g1, g2, rtol=RTOL, what='grad')
        c1 = np.zeros((a.num_segments, 3, 2), dtype=int)
        c2 = np.zeros((a.num_segments, 3, 2), dtype=int)
        a.infer_cn(c1)
        b.infer_cn(c2)
        assert np.array_equal(c1, c2)
    rs = np.random.RandomState(5)
    N, S = 30, 11
    F, T = rs.randn(N, S) * 3., -np.abs(rs.randn(N - 1, S, S)) * 4.
    a1, b1, a2, b2 = (np.zeros((N, S)) for _ in range(4))
    bpmodel_np.sum_product(F, T, a1, b1)
    ref_bpmodel.sum_product(F, T, a2, b2)
    util.assert_close(a1, a2, rtol=1e-12, what='alphas')
    util.assert_close(b1, b2, rtol=1e-12, atol=1e-13, what='betas')
    bpmodel_np.sum_product_2paramtrans(F, a1, b1, 0.3, 0.7)
    ref_bpmodel.sum_product_2paramtrans(F, a2, b2, 0.3, 0.7)
    util.assert_close(a1, a2, rtol=1e-12, what='alphas 2p')
    p1, p2 = np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64)
    assert bpmodel_np.max_product(F, T, p1) == ref_bpmodel.max_product(F, T, p2)
    assert np.array_equal(p1, p2)


def test_oracle_sum_product_vs_exhaustive_enumeration():
    """The reference's own test pattern (remixt/tests/test_model.py:58-93, 207-242): posterior marginals
    and the normalising constant of a 3-segment chain against enumeration of all state sequences."""
    rs = np.random.RandomState(11)
    N, S = 3, 4
    F, T = rs.randn(N, S), rs.randn(N - 1, S, S)
    a, b = np.zeros((N, S)), np.zeros((N, S))
    bpmodel_np.sum_product(F, T, a, b)
    logp = {}
    for seq in itertools.product(range(S), repeat=N):
        logp[seq] = sum(F[n, seq[n]] for n in range(N)) + sum(T[n, seq[n], seq[n + 1]] for n in range(N - 1))
    Z = np.log(sum(np.exp(v) for v in logp.values()))
    util.assert_close(bpmodel_np.logsum(a[-1]), Z, rtol=1e-12, what='norm const')
    for n in range(N):
        marg = np.zeros(S)
        for seq, v in logp.items():
            marg[seq[n]] += np.exp(v - Z)
        util.assert_close(np.exp(a[n] + b[n] - Z), marg, rtol=1e-11, what='marginal')
    best = max(logp, key=logp.get)
    path = np.zeros(N, dtype=np.int64)
    lp = bpmodel_np.max_product(F, T, path)
    assert tuple(path) == best
    util.assert_close(lp, logp[best], rtol=1e-12, what='viterbi logprob')


def test_oracle_digamma_and_lgamma():
    from scipy import special
    x = np.concatenate([np.linspace(1e-3, 30., 200), np.array([1e-7, 5e-7, 123.4, 5000.5])])
    util.assert_close(bpmodel_np.digamma(x), special.digamma(x), rtol=1e-8, what='digamma')
    util.assert_close(bpmodel_np.lgamma(x), special.gammaln(x), rtol=1e-12, atol=1e-13, what='lgamma')
    with pytest.raises(ValueError):
        bpmodel_np.digamma(np.array([0.]))
