"""The likelihood table itself (K1), entry by entry: device vs the oracle's literal evaluation and vs a
40-digit evaluation (mpmath) of the reference's closed forms (bpmodel.pyx:238-272, :304-347).

Two correct FP64 evaluations of these terms differ by a few ulp of their LARGEST intermediate
(lgamma(x + r) ~ x log x), not of the result, so the tolerance is stated on that scale."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

ULPS = 1e-15      # ~5 ulp of the largest intermediate term


def _scales(m):
    x = np.asarray(m.x, dtype=float)
    y = np.asarray(m.y, dtype=float)
    na = y[:, 0] + y[:, 1]
    sc_t = (x + 1e3) * np.log(x + 1e3)
    sc_a = (na + 1e3) * np.log(na + 1e3)
    return sc_t, sc_a


def _oracle_table(ref):
    lt = ref._ll_total()            # (N,S,2)
    la = ref._ll_allele()           # (N,S,2,2)
    N, S = lt.shape[:2]
    out = np.empty((6, N, S))
    out[0], out[1] = lt[:, :, 0], lt[:, :, 1]
    for v in range(2):
        for w in range(2):
            out[2 + 2 * v + w] = la[:, :, v, w]
    return out


def _pair(name, init_id, **kw):
    from oracle import bpmodel_np
    exp = util.load_experiment(name)
    ip = util.init_params(name, init_id)
    cu = util.make_driver(exp, ip, **kw).model
    ref = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel, **kw).model
    return cu, ref


def _check(cu, ref, what):
    dev = cu.log_likelihood_table
    ora = _oracle_table(ref)
    sc_t, sc_a = _scales(ref)
    tol = np.empty_like(dev)
    tol[:2] = (ULPS * sc_t + 1e-13)[np.newaxis, :, np.newaxis]
    tol[2:] = (ULPS * sc_a + 1e-13)[np.newaxis, :, np.newaxis]
    err = np.abs(dev - ora)
    bad = err > tol
    assert not bad.any(), '%s: %d entries beyond tolerance, worst %r (tol %r) dev %r oracle %r' % (
        what, int(bad.sum()), err[bad].max(), tol[bad][err[bad].argmax()], dev[bad][err[bad].argmax()],
        ora[bad][err[bad].argmax()])
    return dev, ora


@pytest.mark.parametrize('name,init_id', [('small', 1), ('small', 7), ('config1', 1)])
def test_table_vs_oracle(name, init_id):
    cu, ref = _pair(name, init_id)
    _check(cu, ref, 'default parameters')
    # other dispersions / haploid depths: small shape parameters (a, b below the Stirling threshold)
    for m in (cu, ref):
        m.negbin_r_0, m.negbin_r_1 = 37.5, 1.25
        m.betabin_M_0, m.betabin_M_1 = 12.0, 0.75
        m.h = np.array([0.02, 0.07, 0.005])
    _check(cu, ref, 'changed parameters')
    # masks switched off on some segments
    tm = np.asarray(ref.total_likelihood_mask).copy()
    am = np.asarray(ref.allele_likelihood_mask).copy()
    tm[::3] = 0
    am[1::4] = 0
    for m in (cu, ref):
        m.total_likelihood_mask = tm
        m.allele_likelihood_mask = am
    dev, _ = _check(cu, ref, 'masks')
    assert np.all(dev[:2, ::3] == 0.) and np.all(dev[2:, 1::4] == 0.)


def test_table_special_parameters():
    """normal_contamination=False: homozygous-deletion and LOH states use their own parameters (slow branch)."""
    cu, ref = _pair('small', 1, normal_contamination=False)
    _check(cu, ref, 'hdel / loh')


def test_table_low_counts():
    """Allele counts of a few reads: lgamma arguments below the wide table (recurrence branch)."""
    from oracle import bpmodel_np
    exp = util.load_experiment('small')
    rs = np.random.RandomState(3)
    exp.x = exp.x.copy()
    a, b = rs.poisson(4., size=exp.x.shape[0]), rs.poisson(2., size=exp.x.shape[0])
    exp.x[:, 0], exp.x[:, 1] = np.maximum(a, b), np.minimum(a, b)      # column 0 is the major count
    exp.x[5, 0] = exp.x[5, 1] = 0.
    exp.x[:, 2] = np.maximum(exp.x[:, 2] // 1000, exp.x[:, 0] + exp.x[:, 1])
    ip = util.init_params('small', 1)
    cu = util.make_driver(exp, ip).model
    ref = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel).model
    for m in (cu, ref):
        m.h = np.asarray(ref.h) / 1000.
    _check(cu, ref, 'low counts')


def test_table_vs_exact():
    """Device error vs a 40-digit evaluation, next to the reference arithmetic's own error."""
    mp = pytest.importorskip('mpmath')
    mp.mp.dps = 40
    cu, ref = _pair('small', 7)
    dev = cu.log_likelihood_table
    ora = _oracle_table(ref)
    sc_t, sc_a = _scales(ref)
    h = [mp.mpf(float(v)) for v in np.asarray(ref.h)]
    rs = np.random.RandomState(11)
    N, S = dev.shape[1:]
    worst_dev = worst_ora = 0.
    for _ in range(300):
        n, s = rs.randint(N), rs.randint(S)
        tot = [int(ref.cn_states[n, s, m, 0] + ref.cn_states[n, s, m, 1]) for m in range(3)]
        a0 = [int(ref.cn_states[n, s, m, 0]) for m in range(3)]
        depth = sum(h[m] * tot[m] for m in range(3))
        if ref.total_likelihood_mask[n]:
            x = mp.mpf(float(ref.x[n]))
            mu = depth * mp.mpf(float(ref.l[n]))
            for u, r in enumerate((ref.negbin_r_0, ref.negbin_r_1)):
                r = mp.mpf(float(r))
                p = mu / (r + mu)
                t = mp.loggamma(x + r) - mp.loggamma(x + 1) - mp.loggamma(r) + x * mp.log(p) + r * mp.log(1 - p)
                worst_dev = max(worst_dev, abs(float(dev[u, n, s] - t)) / sc_t[n])
                worst_ora = max(worst_ora, abs(float(ora[u, n, s] - t)) / sc_t[n])
        if ref.allele_likelihood_mask[n]:
            p = sum(h[m] * a0[m] for m in range(3)) / depth
            y = [mp.mpf(float(ref.y[n, 0])), mp.mpf(float(ref.y[n, 1]))]
            nn = y[0] + y[1]
            for v, M in enumerate((ref.betabin_M_0, ref.betabin_M_1)):
                M = mp.mpf(float(M))
                for w in range(2):
                    k = y[w]
                    t = (mp.loggamma(nn + 1) - mp.loggamma(k + 1) - mp.loggamma(nn - k + 1) + mp.loggamma(k + M * p)
                         + mp.loggamma(nn - k + M * (1 - p)) - mp.loggamma(nn + M) + mp.loggamma(M)
                         - mp.loggamma(M * p) - mp.loggamma(M * (1 - p)))
                    worst_dev = max(worst_dev, abs(float(dev[2 + 2 * v + w, n, s] - t)) / sc_a[n])
                    worst_ora = max(worst_ora, abs(float(ora[2 + 2 * v + w, n, s] - t)) / sc_a[n])
    print('error / largest intermediate: device %.3g, reference arithmetic %.3g' % (worst_dev, worst_ora))
    assert worst_dev < ULPS
    assert worst_dev < 4. * max(worst_ora, 1e-16)
