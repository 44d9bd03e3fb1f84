"""CPU restatement (numpy, FP64) of the reference's inference core ``remixt/bpmodel.pyx``.

TEST INFRASTRUCTURE ONLY: the checker the CUDA path is compared with.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
may import this module; nothing under ``remixt_b200/`` does.

It restates the reference ALGORITHM literally -- it materialises the three ``(N-1, S, S)`` arrays the
reference materialises (``bpmodel.pyx:558-561``) and evaluates every likelihood term where the
reference evaluates it -- vectorised over segments and states instead of looped.  Summation order
inside a numpy reduction differs from the reference's scalar loops, so agreement is ~1e-13
relative, not bitwise; Viterbi paths are bit-exact.

Parity pin: the reference ships no golden vectors for this path (its unit tests import modules that
do not exist), so the pin is (a) the committed vectors in ``tests/golden/``
that the compiled UNMODIFIED reference produced (``oracle/make_golden.py``), checked by
``tests/test_oracle.py``, and (b) the compiled reference itself (``oracle/_ref``) run side by side
where it is built.

Every function cites the reference lines (``remixt/bpmodel.pyx``) it follows.
"""
import math

import numpy as np



def _libm_lgamma():
    # the reference calls libm's lgamma (``from libc.math cimport lgamma``, bpmodel.pyx:8); CPython's
    # math.lgamma is a different (Lanczos) implementation whose few-ulp error on lgamma(1e5) ~ 1e6
    # shows up at 1e-9 in the outlier posteriors, so bind the C library's function itself
    import ctypes
    import ctypes.util
    try:
        libm = ctypes.CDLL(ctypes.util.find_library('m') or 'libm.so.6')
        fn = libm.lgamma
        fn.restype = ctypes.c_double
        fn.argtypes = [ctypes.c_double]
        fn(1.5)
        return fn
    except Exception:
        return math.lgamma


_lgamma = np.frompyfunc(_libm_lgamma(), 1, 1)


def lgamma(a):
    """libm lgamma, elementwise (``bpmodel.pyx:8``)."""
    return np.asarray(_lgamma(np.asarray(a, dtype=float)), dtype=float)


def logsum(X, axis=-1):
    """``_logsum`` (``bpmodel.pyx:77-82``): max-shifted log-sum-exp."""
    vmax = np.max(X, axis=axis, keepdims=True)
    out = np.log(np.sum(np.exp(X - vmax), axis=axis, keepdims=True)) + vmax
    return np.squeeze(out, axis=axis)


def exp_normalize(X, axes=(-1,)):
    """``_exp_normalize`` / ``_exp_normalize2`` (``bpmodel.pyx:120-141``): exp(x - LSE) then divide by its sum."""
    axes = tuple(axes)
    vmax = np.max(X, axis=axes, keepdims=True)
    lse = np.log(np.sum(np.exp(X - vmax), axis=axes, keepdims=True)) + vmax
    Y = np.exp(X - lse)
    return Y / np.sum(Y, axis=axes, keepdims=True)


def entropy(X):
    """``_entropy`` (``bpmodel.pyx:110-115``): sum of x log x over x > 0."""
    X = np.asarray(X, dtype=float).ravel()
    pos = X > 0.
    return float(np.sum(X[pos] * np.log(X[pos])))


def digamma(x):
    """AS 103 digamma as the reference implements it (``bpmodel.pyx:162-235``), elementwise."""
    x = np.array(x, dtype=float, copy=True)
    if np.any(x <= 0.0):
        raise ValueError('x <= 0.0 in digamma')
    out = np.zeros_like(x)
    small = x <= 0.000001
    out[small] = -0.57721566490153286060 - 1.0 / x[small] + 1.6449340668482264365 * x[small]
    xs = x[~small]
    value = np.zeros_like(xs)
    while True:
        low = xs < 8.5
        if not low.any():
            break
        value[low] -= 1.0 / xs[low]
        xs[low] += 1.0
    r = 1.0 / xs
    value += np.log(xs) - 0.5 * r
    r = r * r
    value -= r * (1.0 / 12.0 - r * (1.0 / 120.0 - r * (1.0 / 252.0 - r * (1.0 / 240.0 - r * (1.0 / 132.0)))))
    out[~small] = value
    return out


def negbin_log_likelihood(x, mu, r):
    """``bpmodel.pyx:238-272``; p outside [0, 1] silently becomes 0.5 (``:263-264``)."""
    with np.errstate(divide='ignore', invalid='ignore'):
        p = mu / (r + mu)
        p = np.where((p < 0.) | (p > 1.), 0.5, p)
        ll = lgamma(x + r) - lgamma(x + 1) - lgamma(r) + x * np.log(p) + r * np.log(1 - p)
    if np.any(np.isnan(ll)):
        raise ValueError('ll is nan in negbin_log_likelihood')
    return ll


def negbin_log_likelihood_partial_mu(x, mu, r):
    """``bpmodel.pyx:275-301``."""
    with np.errstate(divide='ignore', invalid='ignore'):
        partial_mu = x / mu - (r + x) / (r + mu)
    if np.any(np.isnan(partial_mu)):
        raise ValueError('partial_mu is nan in negbin_log_likelihood_partial_mu')
    return partial_mu


def betabin_log_likelihood(k, n, p, M):
    """``bpmodel.pyx:304-347``."""
    if np.any(p <= 0.) or np.any((1 - p) <= 0.):
        raise ValueError('p <= 0 or (1 - p) <= 0. for expected allele ratio p')
    ll = (lgamma(n + 1) - lgamma(k + 1) - lgamma(n - k + 1)
          + lgamma(k + M * p) + lgamma(n - k + M * (1 - p))
          - lgamma(n + M)
          - lgamma(M * p) - lgamma(M * (1 - p))
          + lgamma(M))
    if np.any(np.isnan(ll)):
        raise ValueError('ll is nan in betabin_log_likelihood')
    return ll


def betabin_log_likelihood_partial_p(k, n, p, M):
    """``bpmodel.pyx:350-394``."""
    if np.any(p <= 0.) or np.any((1 - p) <= 0.):
        raise ValueError('p <= 0 or (1 - p) <= 0. for expected allele ratio p')
    partial_p = (M * digamma(k + M * p)
                 + (-M) * digamma(n - k + M * (1 - p))
                 - M * digamma(M * p)
                 - (-M) * digamma(M * (1 - p)))
    if np.any(np.isnan(partial_p)):
        raise ValueError('partial_p is nan in betabin_log_likelihood_partial_p')
    return partial_p


# ---------------------------------------------------------------------------------------------
# module functions
# ---------------------------------------------------------------------------------------------
def sum_product(framelogprob, log_transmat, alphas, betas):
    """Sum-product messages on a chain, log domain (``bpmodel.pyx:1213-1246``)."""
    F = np.asarray(framelogprob)
    T = np.asarray(log_transmat)
    N = F.shape[0]
    alphas[0, :] = F[0, :]
    for n in range(1, N):
        alphas[n, :] = logsum(alphas[n - 1, :, np.newaxis] + T[n - 1], axis=0) + F[n, :]
    betas[N - 1, :] = 0.0
    for n in range(N - 2, -1, -1):
        betas[n, :] = logsum(T[n] + (F[n + 1, :] + betas[n + 1, :])[np.newaxis, :], axis=1)


def _logsumpair(x, y):
    vmax = np.maximum(x, y)
    return np.log(np.exp(x - vmax) + np.exp(y - vmax)) + vmax


def sum_product_2paramtrans(framelogprob, alphas, betas, e0, e1):
    """Two-parameter (stay / switch) transition messages (``bpmodel.pyx:1249-1292``)."""
    F = np.asarray(framelogprob)
    N = F.shape[0]
    log_e0, log_e1 = math.log(e0), math.log(e1)
    alphas[0, :] = F[0, :]
    for n in range(1, N):
        notrans = log_e0 + logsum(alphas[n - 1, :])
        alphas[n, :] = _logsumpair(notrans, log_e1 + alphas[n - 1, :]) + F[n, :]
    betas[N - 1, :] = 0.0
    for n in range(N - 2, -1, -1):
        work = F[n + 1, :] + betas[n + 1, :]
        notrans = log_e0 + logsum(work)
        betas[n, :] = _logsumpair(notrans, log_e1 + F[n + 1, :] + betas[n + 1, :])


def max_product(framelogprob, log_transmat, state_sequence):
    """Viterbi decode (``bpmodel.pyx:1296-1333``).  ``_max`` / ``_argmax`` (``:21``, ``:48``) keep the
    FIRST maximum (strict ``>``), which is what ``numpy.argmax`` does; an all-NaN / all -inf buffer
    gives index 0 in both."""
    F = np.asarray(framelogprob)
    T = np.asarray(log_transmat)
    N = F.shape[0]
    lattice = np.zeros_like(F)
    lattice[0] = F[0]
    for n in range(1, N):
        lattice[n] = np.max(lattice[n - 1, :, np.newaxis] + T[n - 1], axis=0) + F[n]
    pos = int(np.argmax(lattice[-1, :]))
    state_sequence[-1] = pos
    logprob = lattice[-1, pos]
    for n in range(N - 2, -1, -1):
        pos = int(np.argmax(lattice[n, :] + T[n, :, state_sequence[n + 1]]))
        state_sequence[n] = pos
    return float(logprob)


# ---------------------------------------------------------------------------------------------
# RemixtModel
# ---------------------------------------------------------------------------------------------
class RemixtModel(object):
    """numpy restatement of ``cdef class RemixtModel`` (``bpmodel.pyx:397-1210``): same constructor,
    attribute names and method names; every array attribute is a plain numpy array."""

    def __init__(self, num_clones, num_segments, num_breakpoints, normal_contamination, cn_states, brk_states,
                 h_init, l, x, y, is_telomere, breakpoint_idx, breakpoint_orient, transition_penalty,
                 divergence_weight, shard=False):
        # bpmodel.pyx:478-491
        self.num_clones = int(num_clones)
        self.num_segments = int(num_segments)
        self.num_breakpoints = int(num_breakpoints)
        self.normal_contamination = bool(normal_contamination)
        self.cn_states = np.asarray(cn_states)
        self.brk_states = np.asarray(brk_states)
        self.h = np.array(h_init, dtype=float)
        self.l = np.asarray(l, dtype=float)
        self.x = np.asarray(x, dtype=float)
        self.y = np.asarray(y, dtype=float)
        self.num_alleles = 2
        self.cn_max = int(max(self.cn_states.max(), self.brk_states.max()))
        self.num_cn_states = self.cn_states.shape[1]
        self.num_brk_states = self.brk_states.shape[0]
        N, S, M, B, K = (self.num_segments, self.num_cn_states, self.num_clones, self.num_brk_states,
                         self.num_breakpoints)
        self.total_likelihood_mask = np.ones((N,), dtype=np.int64)
        self.allele_likelihood_mask = np.ones((N,), dtype=np.int64)
        # derived tables, bpmodel.pyx:497-507
        cs = self.cn_states
        self.cn_states_total = cs.sum(axis=-1)
        self.num_alleles_subclonal = np.sum(cs[:, :, 1:, :].max(axis=-2) != cs[:, :, 1:, :].min(axis=-2), axis=-1)
        self.is_hdel = np.all(cs == 0, axis=(-2, -1)) * 1
        self.is_loh = np.any(cs.sum(axis=-2) == 0, axis=-1) * 1
        # validation, bpmodel.pyx:509-529
        if cs.shape[0] != N or cs.shape[2] != M or cs.shape[3] != 2:
            raise ValueError('cn_states must have shape (num_segments, num_cn_states, num_clones, num_alleles)')
        if self.brk_states.shape[1] != M:
            raise ValueError('cn_states must have shape (num_brk_states, num_clones)')
        if self.h.shape[0] != M:
            raise ValueError('h must have length equal to num_clones')
        for name, arr in (('is_telomere', is_telomere), ('breakpoint_idx', breakpoint_idx),
                          ('breakpoint_orient', breakpoint_orient)):
            if np.asarray(arr).shape[0] != N:
                raise ValueError('%s must have length equal to num_segments' % name)
        # shard=True: one chromosome group of a sharded fit (all K breakpoints, its own breakends); test-only
        if (np.asarray(breakpoint_idx).max() + 1 > K) if shard else (np.asarray(breakpoint_idx).max() + 1 != K):
            raise ValueError('breakpoint_idx must have maximum of num_breakpoints positive indices')
        self.is_telomere = np.asarray(is_telomere)
        self.breakpoint_idx = np.asarray(breakpoint_idx)
        self.breakpoint_orient = np.asarray(breakpoint_orient)
        self.transition_penalty = abs(float(transition_penalty))
        self.divergence_weight = abs(float(divergence_weight))
        self.transition_model = 0                                   # zero-initialised cdef int
        # p_breakpoint favours states with at most one copy per clone, bpmodel.pyx:546-554
        self.p_breakpoint = np.zeros((K, B))
        self.p_breakpoint[:, self.brk_states.max(axis=1) <= 1] = 1.
        if K > 0:
            self.p_breakpoint /= np.sum(self.p_breakpoint, axis=-1)[:, np.newaxis]
        # bpmodel.pyx:556-567
        self.hmm_log_norm_const = 0.
        self.framelogprob = np.ones((N, S))
        self.log_transmat = np.zeros((N - 1, S, S))
        self.cached_log_transmat = np.zeros((N - 1, S, S))
        self.posterior_marginals = np.ones((N, S))
        self.posterior_marginals /= np.sum(self.posterior_marginals, axis=-1)[:, np.newaxis]
        self.joint_posterior_marginals = np.ones((N - 1, S, S))
        self.joint_posterior_marginals /= np.sum(self.joint_posterior_marginals, axis=(-2, -1))[:, np.newaxis, np.newaxis]
        # bpmodel.pyx:569-584
        self.p_allele_swap = np.ones((N, 2)) * 0.5
        self.prior_outlier_total = 0.01
        self.prior_outlier_allele = 0.01
        self.p_outlier_total = np.ones((N, 2))
        self.p_outlier_allele = np.ones((N, 2))
        self.p_outlier_total[:, 0] = 1. - self.prior_outlier_total
        self.p_outlier_total[:, 1] = self.prior_outlier_total
        self.p_outlier_allele[:, 0] = 1. - self.prior_outlier_allele
        self.p_outlier_allele[:, 1] = self.prior_outlier_allele
        # bpmodel.pyx:587-597
        self.negbin_r_0, self.negbin_r_1 = 500., 10.
        self.negbin_hdel_mu, self.negbin_hdel_r_0, self.negbin_hdel_r_1 = 1e-5, 10., 1.
        self.betabin_M_0, self.betabin_M_1 = 500., 10.
        self.betabin_loh_p, self.betabin_loh_M_0, self.betabin_loh_M_1 = 1e-3, 10., 1.
        # bpmodel.pyx:604
        self.calculate_log_transmat(self.cached_log_transmat)

    # ------------------------------------------------------------------ transitions
    def calc_transition(self, cn_diff):
        """``bpmodel.pyx:607-616``."""
        cn_diff = np.asarray(cn_diff, dtype=float)
        if self.transition_model == 0:
            return np.abs(cn_diff)
        return (cn_diff != 0) * 1.

    def _expected_brk_transition(self, n, m):
        """The ``_p_d`` table of ``bpmodel.pyx:658-664`` for row n, clone m: index d + cn_max + 1."""
        d = np.arange(-self.cn_max - 1, self.cn_max + 2)
        k = self.breakpoint_idx[n]
        pd = np.zeros(d.shape[0])
        for s_b in range(self.num_brk_states):          # reference accumulation order over s_b
            pd += self.p_breakpoint[k, s_b] * self.calc_transition(d - self.breakpoint_orient[n] * self.brk_states[s_b, m])
        return pd

    def calculate_log_transmat(self, log_transmat):
        """``bpmodel.pyx:640-684``: same accumulation order per entry (clones, then the allele-flip term)."""
        log_transmat[:] = 0.
        pen = self.transition_penalty
        tot, cs = self.cn_states_total, self.cn_states
        for n in range(self.num_segments - 1):
            if self.is_telomere[n] > 0:
                continue
            T = np.zeros((self.num_cn_states, self.num_cn_states))
            dtot = tot[n, :, np.newaxis, :] - tot[n + 1, np.newaxis, :, :]            # (S,S,M)
            if self.breakpoint_idx[n] < 0:
                for m in range(self.num_clones):
                    T += -pen * self.calc_transition(dtot[:, :, m])
            else:
                for m in range(self.num_clones):
                    pd = self._expected_brk_transition(n, m)
                    T += -pen * pd[dtot[:, :, m] + self.cn_max + 1]
            change = np.zeros((2, self.num_cn_states, self.num_cn_states))
            for flip in range(2):
                for m in range(self.num_clones):
                    for allele in range(2):
                        other = 1 - allele if flip == 1 else allele
                        change[flip] += self.calc_transition(cs[n, :, np.newaxis, m, allele] - cs[n + 1, np.newaxis, :, m, other])
                    change[flip] -= self.calc_transition(dtot[:, :, m])
            T += -pen * np.minimum(change[0], change[1])
            log_transmat[n] = T

    # ------------------------------------------------------------------ likelihood terms
    def _rows(self, rows):
        return np.arange(self.num_segments) if rows is None else np.asarray(rows)

    def _expected_total_reads(self, rows=None):
        """(R,S) ``calculate_expected_total_reads`` (``bpmodel.pyx:686-698``), clone order of the reference."""
        rows = self._rows(rows)
        mu = np.zeros((rows.shape[0], self.num_cn_states))
        for m in range(self.num_clones):
            mu += self.h[m] * self.cn_states_total[rows, :, m]
        return mu * self.l[rows, np.newaxis]

    def _depths(self, rows=None):
        rows = self._rows(rows)
        minor = np.zeros((rows.shape[0], self.num_cn_states))
        total = np.zeros((rows.shape[0], self.num_cn_states))
        for m in range(self.num_clones):
            minor += self.h[m] * self.cn_states[rows, :, m, 0]
            total += self.h[m] * self.cn_states_total[rows, :, m]
        return minor, total

    def _log_prior_cn(self):
        """(N,S) ``calculate_log_prior_cn`` (``bpmodel.pyx:746-749``)."""
        return -1.0 * self.num_alleles_subclonal * self.l[:, np.newaxis] * self.divergence_weight

    def _ll_total(self, rows=None):
        """(R,S,2) ``calculate_log_likelihood_total`` (``bpmodel.pyx:751-776``) for u = 0, 1."""
        rows = self._rows(rows)
        out = np.zeros((rows.shape[0], self.num_cn_states, 2))
        mu = self._expected_total_reads(rows)
        x = np.broadcast_to(self.x[rows, np.newaxis], mu.shape)
        use = np.broadcast_to(self.total_likelihood_mask[rows, np.newaxis] != 0, mu.shape)
        hdel = (self.is_hdel[rows] == 1) & (not self.normal_contamination)
        for u, (r, r_hdel) in enumerate(((self.negbin_r_0, self.negbin_hdel_r_0), (self.negbin_r_1, self.negbin_hdel_r_1))):
            sel = use & ~hdel
            out[sel, u] = negbin_log_likelihood(x[sel], mu[sel], r)
            sel = use & hdel
            if sel.any():
                out[sel, u] = negbin_log_likelihood(x[sel], self.negbin_hdel_mu, r_hdel)
        return out

    def _allele_p(self, rows):
        """Expected allele ratio with the hdel / LOH substitutions of ``bpmodel.pyx:813-834``;
        returns (p, loh_special, evaluated) over (R,S)."""
        minor, total = self._depths(rows)
        use = np.broadcast_to(self.allele_likelihood_mask[rows, np.newaxis] != 0, minor.shape)
        hdel = self.is_hdel[rows] == 1
        if np.any(use & ~hdel & (total <= 0)):
            raise ValueError('total_depth <= 0 for a copy number state')
        with np.errstate(divide='ignore', invalid='ignore'):
            p = np.where(hdel, 0., minor / total)
        loh = (self.is_loh[rows] == 1) & (not self.normal_contamination)
        if np.any(use & loh & (p != 0.) & (p != 1.)):
            raise ValueError('unexpected expected allele ratio p for loh state')
        p = np.where(loh & (p == 0.), self.betabin_loh_p, np.where(loh & (p == 1.), 1. - self.betabin_loh_p, p))
        return p, loh, use

    def _ll_allele(self, rows=None):
        """(R,S,2,2) ``calculate_log_likelihood_allele`` (``bpmodel.pyx:809-853``) for (v, w)."""
        rows = self._rows(rows)
        out = np.zeros((rows.shape[0], self.num_cn_states, 2, 2))
        p, loh, use = self._allele_p(rows)
        n_reads = np.broadcast_to((self.y[rows, 0] + self.y[rows, 1])[:, np.newaxis], p.shape)
        use = use & (n_reads != 0)               # zero allelic reads -> 0 AFTER p was computed (:844-845)
        for v, (M, M_loh) in enumerate(((self.betabin_M_0, self.betabin_loh_M_0), (self.betabin_M_1, self.betabin_loh_M_1))):
            for w in range(2):
                k = np.broadcast_to(self.y[rows, w][:, np.newaxis], p.shape)
                for sel, Mv in ((use & ~loh, M), (use & loh, M_loh)):
                    if sel.any():
                        out[sel, v, w] = betabin_log_likelihood(k[sel], n_reads[sel], p[sel], Mv)
        return out

    # scalar accessors (cpdef, bpmodel.pyx:686-896)
    def calculate_expected_total_reads(self, n, s):
        return float(self._expected_total_reads(np.array([n]))[0, s])

    def calculate_expected_allele_ratio(self, n, s):
        minor, total = self._depths(np.array([n]))
        if total[0, s] <= 0:
            raise ValueError('total_depth <= 0 for s: {}'.format(s))
        return float(minor[0, s] / total[0, s])

    def calculate_log_prior_cn(self, n, s):
        return float(self._log_prior_cn()[n, s])

    def calculate_log_likelihood_total(self, n, s, u):
        return float(self._ll_total(np.array([n]))[0, s, u])

    def calculate_log_likelihood_allele(self, n, s, v, w):
        return float(self._ll_allele(np.array([n]))[0, s, v, w])

    # ------------------------------------------------------------------ coordinate updates
    def update_framelogprob(self):
        """``bpmodel.pyx:898-919``."""
        F = np.zeros((self.num_segments, self.num_cn_states))
        llt, lla = self._ll_total(), self._ll_allele()
        for u in range(2):
            F += self.p_outlier_total[:, u, np.newaxis] * llt[:, :, u]
        for v in range(2):
            for w in range(2):
                F += self.p_outlier_allele[:, v, np.newaxis] * self.p_allele_swap[:, w, np.newaxis] * lla[:, :, v, w]
        F += self._log_prior_cn()
        self.framelogprob = F

    def update_p_cn(self):
        """``bpmodel.pyx:921-962``."""
        N, S = self.num_segments, self.num_cn_states
        alphas, betas = np.empty((N, S)), np.empty((N, S))
        self.update_framelogprob()
        assert not np.any(np.isnan(self.framelogprob))
        self.calculate_log_transmat(self.log_transmat)
        sum_product(self.framelogprob, self.log_transmat, alphas, betas)
        assert not np.any(np.isnan(alphas))
        assert not np.any(np.isnan(betas))
        self.hmm_log_norm_const = float(logsum(alphas[-1, :]))
        self.posterior_marginals = exp_normalize(alphas + betas, axes=(-1,))
        assert not np.any(np.isnan(self.posterior_marginals))
        if N > 1:
            lj = (alphas[:-1, :, np.newaxis] + self.log_transmat
                  + (self.framelogprob[1:, :] + betas[1:, :])[:, np.newaxis, :])
            self.joint_posterior_marginals = exp_normalize(lj, axes=(-2, -1))
        assert not np.any(np.isnan(self.joint_posterior_marginals))

    def update_p_breakpoint(self):
        """``bpmodel.pyx:964-985`` with ``add_log_breakpoint_p_expectation_cn`` (``:619-637``)."""
        self.apply_breakpoint_logq(self.breakpoint_logq())

    def apply_breakpoint_logq(self, log_p):
        """The tail of ``update_p_breakpoint`` (``bpmodel.pyx:982-985``)."""
        if self.num_breakpoints > 0:
            self.p_breakpoint = exp_normalize(np.asarray(log_p, dtype=float), axes=(-1,))
        self.calculate_log_transmat(self.cached_log_transmat)

    def breakpoint_logq(self):
        """The accumulation of ``update_p_breakpoint`` (``bpmodel.pyx:966-980``) over this model's breakend rows."""
        log_p = np.zeros((self.num_breakpoints, self.num_brk_states))
        tot = self.cn_states_total
        d = np.arange(-self.cn_max - 1, self.cn_max + 2)
        for n in range(self.num_segments - 1):
            if self.breakpoint_idx[n] < 0:
                continue
            for m in range(self.num_clones):
                dd = tot[n, :, np.newaxis, m] - tot[n + 1, np.newaxis, :, m]
                p_d = np.bincount((dd + self.cn_max + 1).ravel(), weights=self.joint_posterior_marginals[n].ravel(),
                                  minlength=d.shape[0])
                for s_b in range(self.num_brk_states):
                    f = self.calc_transition(d - self.breakpoint_orient[n] * self.brk_states[s_b, m])
                    log_p[self.breakpoint_idx[n], s_b] += np.sum(-self.transition_penalty * p_d * f)
        return log_p

    def update_p_outlier_total(self):
        """``bpmodel.pyx:987-1003``."""
        llt = self._ll_total()
        lp = np.empty((self.num_segments, 2))
        lp[:, 0] = math.log(1. - self.prior_outlier_total)
        lp[:, 1] = math.log(self.prior_outlier_total)
        lp += np.sum(self.posterior_marginals[:, :, np.newaxis] * llt, axis=1)
        self.p_outlier_total = exp_normalize(lp)

    def update_p_outlier_allele(self):
        """``bpmodel.pyx:1005-1023``."""
        lla = self._ll_allele()
        lp = np.empty((self.num_segments, 2))
        lp[:, 0] = math.log(1. - self.prior_outlier_allele)
        lp[:, 1] = math.log(self.prior_outlier_allele)
        w = self.posterior_marginals[:, :, np.newaxis, np.newaxis] * self.p_allele_swap[:, np.newaxis, np.newaxis, :]
        lp += np.sum(w * lla, axis=(1, 3))
        self.p_outlier_allele = exp_normalize(lp)

    def update_p_allele_swap(self):
        """``bpmodel.pyx:1025-1042``."""
        lla = self._ll_allele()
        w = self.posterior_marginals[:, :, np.newaxis, np.newaxis] * self.p_outlier_allele[:, np.newaxis, :, np.newaxis]
        self.p_allele_swap = exp_normalize(np.sum(w * lla, axis=(1, 2)))

    # ------------------------------------------------------------------ objective
    def calculate_variational_entropy(self):
        """``bpmodel.pyx:1044-1058``."""
        e = -self.hmm_log_norm_const
        e += np.sum(self.posterior_marginals * self.framelogprob)
        e += np.sum(self.joint_posterior_marginals * self.log_transmat)
        e += entropy(self.p_breakpoint)
        e += entropy(self.p_outlier_total)
        e += entropy(self.p_outlier_allele)
        e += entropy(self.p_allele_swap)
        return float(e)

    def _expected_ll_terms(self, rows=None):
        rows = self._rows(rows)
        post = self.posterior_marginals[rows]
        llt, lla = self._ll_total(rows), self._ll_allele(rows)
        t = np.sum(post[:, :, np.newaxis] * self.p_outlier_total[rows][:, np.newaxis, :] * llt)
        a = np.sum(post[:, :, np.newaxis, np.newaxis] * self.p_outlier_allele[rows][:, np.newaxis, :, np.newaxis]
                   * self.p_allele_swap[rows][:, np.newaxis, np.newaxis, :] * lla)
        return float(t), float(a)

    def calculate_variational_energy(self):
        """``bpmodel.pyx:1060-1117``."""
        energy = np.sum(self.posterior_marginals * self._log_prior_cn())
        t, a = self._expected_ll_terms()
        energy += t
        energy += np.sum(self.p_outlier_total[:, 0] * math.log(1. - self.prior_outlier_total)
                         + self.p_outlier_total[:, 1] * math.log(self.prior_outlier_total))
        energy += a
        energy += np.sum(self.p_outlier_allele[:, 0] * math.log(1. - self.prior_outlier_allele)
                         + self.p_outlier_allele[:, 1] * math.log(self.prior_outlier_allele))
        energy += np.sum(self.joint_posterior_marginals * self.cached_log_transmat)
        return float(energy)

    def calculate_elbo(self):
        """``bpmodel.pyx:1119-1123``."""
        return self.calculate_variational_energy() - self.calculate_variational_entropy()

    def calculate_expected_log_likelihood(self, sample):
        """``bpmodel.pyx:1125-1157``."""
        rows = np.flatnonzero(np.asarray(sample) != 0)
        if rows.size == 0:
            return 0.
        t, a = self._expected_ll_terms(rows)
        return t + a

    def calculate_expected_log_likelihood_partial_h(self, sample, partial_h):
        """``bpmodel.pyx:1159-1195`` with the ``_partial_h`` helpers (``:700, :726, :778, :855``)."""
        partial_h[:] = 0.
        rows = np.flatnonzero(np.asarray(sample) != 0)
        if rows.size == 0:
            return
        post = self.posterior_marginals[rows]
        shape = post.shape
        special = not self.normal_contamination
        # total reads
        mu = self._expected_total_reads(rows)
        x = np.broadcast_to(self.x[rows, np.newaxis], shape)
        use = np.broadcast_to(self.total_likelihood_mask[rows, np.newaxis] != 0, shape) & ~((self.is_hdel[rows] == 1) & special)
        for u, r in enumerate((self.negbin_r_0, self.negbin_r_1)):
            dmu = np.zeros(shape)
            dmu[use] = negbin_log_likelihood_partial_mu(x[use], mu[use], r)
            wgt = post * self.p_outlier_total[rows, u][:, np.newaxis] * dmu
            for m in range(self.num_clones):
                partial_h[m] += np.sum(wgt * (self.l[rows, np.newaxis] * self.cn_states_total[rows, :, m]))
        # allele reads
        minor, total = self._depths(rows)
        n_reads = np.broadcast_to((self.y[rows, 0] + self.y[rows, 1])[:, np.newaxis], shape)
        use = np.broadcast_to(self.allele_likelihood_mask[rows, np.newaxis] != 0, shape) & ~((self.is_loh[rows] == 1) & special)
        if np.any(use & (total <= 0)):
            raise ValueError('total_depth <= 0 for a copy number state')
        use = use & (n_reads != 0)
        with np.errstate(divide='ignore', invalid='ignore'):
            p = minor / total
        for v, M in enumerate((self.betabin_M_0, self.betabin_M_1)):
            for w in range(2):
                k = np.broadcast_to(self.y[rows, w][:, np.newaxis], shape)
                dp = np.zeros(shape)
                if use.any():
                    dp[use] = betabin_log_likelihood_partial_p(k[use], n_reads[use], p[use], M)
                wgt = post * self.p_outlier_allele[rows, v][:, np.newaxis] * self.p_allele_swap[rows, w][:, np.newaxis] * dp
                for m in range(self.num_clones):
                    with np.errstate(divide='ignore', invalid='ignore'):
                        ratio_partial = (self.cn_states[rows, :, m, 0] * total - minor * self.cn_states_total[rows, :, m]) / (total * total)
                    partial_h[m] += np.sum(np.where(use, wgt * ratio_partial, 0.))

    # ------------------------------------------------------------------ decode
    def infer_cn(self, cn):
        """``bpmodel.pyx:1197-1210``.  The allele-swap branch re-assigns the loop variable and indexes
        both sides with it, so the output is ``cn_states[n, path[n]]`` whatever ``p_allele_swap`` is."""
        path = np.zeros((self.num_segments,), dtype=np.int64)
        max_product(self.framelogprob, self.log_transmat, path)
        self.state_sequence = path
        cn[:] = self.cn_states[np.arange(self.num_segments), path]
