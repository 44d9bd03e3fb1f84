"""Host-side driver of the B200 inference core: the mirror of the reference's
``remixt/cn_model.py`` (class ``BreakpointModel``, ``cn_model.py:29-628``).

Same constructor keywords, same ``fit(h_init)`` / ``optimal_cn()`` / property surface, same
sequence of calls into the ``RemixtModel`` object and the same consumption of the global
``numpy.random`` stream, so that a fit driven from here and a fit driven by the reference's own
``cn_model.py`` visit identical states when both sit on the same ``RemixtModel`` implementation.
The numerical work is all in ``RemixtModel`` (CUDA by default); what stays on the host is what
the reference keeps on the host: the segment remap, the scipy optimisers of the M-step and the
breakpoint decode.

``model_class`` (extension) picks the ``RemixtModel`` implementation; the parity tests pass the
compiled reference class there to check this driver independently of the kernels.
"""
import collections
import contextlib
import datetime
import itertools

import numpy as np
import scipy.optimize

from . import bpmodel as _bpmodel


def _now():
    return datetime.datetime.now().time().isoformat()


_STATE_TABLES = {}     # the enumerations below are pure functions of small integers; every fit of a grid asks again


def create_cn_states(num_clones, num_alleles, cn_max, cn_diff_max):
    """Allele specific copy number states of one segment, ``(S, M, 2)`` (``cn_model.py:228-253``), memoised."""
    key = ('cn', int(num_clones), int(num_alleles), int(cn_max), int(cn_diff_max))
    if key not in _STATE_TABLES:
        _STATE_TABLES[key] = _create_cn_states(num_clones, num_alleles, cn_max, cn_diff_max)
    return _STATE_TABLES[key].copy()


def _create_cn_states(num_clones, num_alleles, cn_max, cn_diff_max):
    """Allele specific copy number states of one segment, ``(S, M, 2)`` (``cn_model.py:228-253``).

    Order and representatives follow the reference exactly: candidates are visited in
    ``itertools.product`` order; a state and its allele-swapped twin share one slot whose
    position is that of the first of the two visited and whose value is the last visited.
    """
    tumour_vars = (num_clones - 1) * num_alleles
    slots = {}
    for combo in itertools.product(range(cn_max + 1), repeat=tumour_vars):
        state = np.array((1, 1) + combo).reshape((num_clones, num_alleles))
        tumour = state[1:, :]
        if (tumour.sum(axis=1) > cn_max).any():
            continue
        if ((tumour.max(axis=0) - tumour.min(axis=0)) > cn_diff_max).any():
            continue
        key = frozenset([tuple(tumour.flatten()), tuple(tumour[:, ::-1].flatten())])
        slots[key] = state
    return np.array(list(slots.values()))


def create_brk_states(num_clones, cn_max, cn_diff_max):
    """Breakpoint copy number states, ``(B, M)`` with clone 0 fixed at 0 (``cn_model.py:255-276``), memoised."""
    key = ('brk', int(num_clones), int(cn_max), int(cn_diff_max))
    if key not in _STATE_TABLES:
        _STATE_TABLES[key] = _create_brk_states(num_clones, cn_max, cn_diff_max)
    return _STATE_TABLES[key].copy()


def _create_brk_states(num_clones, cn_max, cn_diff_max):
    """Breakpoint copy number states, ``(B, M)`` with clone 0 fixed at 0 (``cn_model.py:255-276``)."""
    states = []
    for combo in itertools.product(range(cn_max + 1), repeat=num_clones - 1):
        tumour = np.array(combo)
        if (tumour > cn_max).any():
            continue
        if tumour.max() - tumour.min() > cn_diff_max:
            continue
        states.append(np.array((0,) + combo))
    return np.array(states)


def _numerical_gradient(x, f, args=()):
    # forward differences; only reached on the optimiser's failure branch (cn_model.py:513-518)
    x = np.asarray(x, dtype=float)
    f0 = f(x, *args)
    g = np.zeros_like(x)
    for i in range(x.shape[0]):
        step = 1.4901161193847656e-08 * max(abs(x[i]), 0.1)
        xs = x.copy()
        xs[i] += step
        g[i] = (f(xs, *args) - f0) / step
    return g


class BreakpointModel(object):

    likelihood_param_bounds = {
        'negbin_r_0': (10., 2000.),
        'negbin_r_1': (1., 2000.),
        'betabin_M_0': (10., 2000.),
        'betabin_M_1': (1., 2000.),
        'negbin_hdel_mu': (1e-9, 1e-4),
        'negbin_hdel_r_0': (10., 2000.),
        'negbin_hdel_r_1': (1., 200.),
        'betabin_loh_p': (1e-5, 1e-2),
        'betabin_loh_M_0': (10., 2000.),
        'betabin_loh_M_1': (1., 200.),
    }

    def __init__(self, x, l, adjacencies, breakpoints, **kwargs):
        """Create a copy number model (arguments and keywords as ``cn_model.py:31-52``).

        Args:
            x (numpy.array): observed major, minor, total reads, ``(N, 3)``
            l (numpy.array): segment lengths ``(N,)``
            adjacencies (iterable of tuple): pairs of adjacent segments
            breakpoints (dict): breakpoint id -> frozenset of two ``(segment, side)`` breakends
        """
        assert np.all(x[:, 1] <= x[:, 0])

        self.N = x.shape[0]
        self.breakpoint_ids, self.breakpoints = zip(*breakpoints.items()) if len(breakpoints) else ((), ())

        self.max_copy_number = kwargs.get('max_copy_number', 6)
        self.max_copy_number_diff = kwargs.get('max_copy_number_diff', 1)
        self.normal_contamination = kwargs.get('normal_contamination', True)
        self.is_female = kwargs.get('is_female', True)
        self.divergence_weight = kwargs.get('divergence_weight', 1e6)
        self.min_segment_length = kwargs.get('min_segment_length', 10000)
        self.min_proportion_genotyped = kwargs.get('min_proportion_genotyped', 0.01)
        self.max_depth = kwargs.get('max_depth')
        self.transition_log_prob = kwargs.get('transition_log_prob', 10.)
        self.transition_model = kwargs.get('transition_model', 0)
        self.disable_breakpoints = kwargs.get('disable_breakpoints', False)
        self.breakpoint_init = kwargs.get('breakpoint_init', None)
        self.normal_copies = kwargs['normal_copies'] if 'normal_copies' in kwargs else np.ones((self.N, 2), dtype=int)
        self.do_h_update = kwargs.get('do_h_update', True)
        self.model_class = kwargs.get('model_class', None) or _bpmodel.RemixtModel
        self.model_kwargs = kwargs.get('model_kwargs', {})
        self.verbose = kwargs.get('verbose', False)
        # extension: a numpy.random.RandomState to draw the M-step subsamples from; None = the
        # global numpy.random stream, as the reference (RandomState(s) and seed(s) give one stream)
        self.random_state = kwargs.get('random_state', None)

        if self.max_depth is None:
            raise ValueError('must specify max depth')

        if not self.normal_contamination:
            self.normal_copies = self.normal_copies * 0

        # extension: ``segmentation=`` another BreakpointModel built from the SAME (x, l, adjacencies,
        # breakpoints); its dummy-segment layout is adopted instead of being rebuilt (the fits of an
        # initialisation grid differ in max_depth / divergence_weight only)
        donor = kwargs.get('segmentation', None)
        if donor is not None and donor.N == self.N and donor.breakpoint_ids == self.breakpoint_ids:
            self._adopt_segmentation(donor)
        else:
            self._build_segmentation(x, l, adjacencies)

        # likelihood masks for poorly modelled segments (cn_model.py:169-184)
        # everything but the depth cut depends on the segmentation and two thresholds only: the fits of an
        # initialisation grid (which differ in max_depth) take it from the donor
        key = (float(self.min_segment_length), float(self.min_proportion_genotyped))
        cached = getattr(donor, '_mask_inputs', None) if donor is not None and donor.N1 == self.N1 else None
        if cached is not None and cached[0] == key:
            _, long_enough, genotyped_ok, depth = cached
        else:
            long_enough = self.l1 >= self.min_segment_length
            genotyped = self.x1[:, :2].sum(axis=1).astype(float) / (self.x1[:, 2].astype(float) + 1e-16)
            depth = self.x1[:, 2].astype(float) / (self.l1.astype(float) + 1e-16)
            genotyped_ok = genotyped >= self.min_proportion_genotyped
        self._mask_inputs = (key, long_enough, genotyped_ok, depth)
        not_amplified = depth <= self.max_depth
        self._total_likelihood_mask = long_enough & not_amplified
        self._allele_likelihood_mask = long_enough & genotyped_ok & not_amplified

        if self.disable_breakpoints:
            for f in ('num_breakpoints', 'breakpoint_idx', 'breakpoint_orient'):
                setattr(self, '_seg_' + f, getattr(self, f))
            self.num_breakpoints = 0
            self.breakpoint_idx = -np.ones(self.breakpoint_idx.shape, dtype=int)
            self.breakpoint_orient = np.zeros(self.breakpoint_orient.shape, dtype=int)

        self.check_elbo = False
        # (name, accepted, E[LL] before, E[LL] after) of every M-step acceptance test, in call order (cn_model.py:526-531, :564-569)
        self.decisions = []
        self.prev_elbo = None
        self.prev_elbo_diff = None
        self.num_em_iter = 1
        self.num_update_iter = 1

        self.likelihood_params = ['negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1']
        if not self.normal_contamination:
            self.likelihood_params += ['negbin_hdel_mu', 'negbin_hdel_r_0', 'negbin_hdel_r_1',
                                       'betabin_loh_p', 'betabin_loh_M_0', 'betabin_loh_M_1']

    def _build_segmentation(self, x, l, adjacencies):
        """Insert zero-length dummy segments so that at most one breakend sits between any two
        adjacent model segments (``cn_model.py:82-167``).

        A breakend ``(n, side)`` sits on the boundary to the right of segment ``n`` (side 1,
        orientation +1) or to its left, i.e. to the right of ``n - 1`` (side 0, orientation -1).
        Each boundary with breakends becomes a run of model segments, one per breakend, the first
        of which is the original segment; a boundary that is not a wild-type adjacency gets one
        more zero-length segment to act as the telomere.
        """
        adjacent = adjacencies if isinstance(adjacencies, (set, frozenset)) else set(map(tuple, adjacencies))

        at_boundary = collections.defaultdict(set)
        for bp_idx, breakpoint in enumerate(self.breakpoints):
            for be_idx, (n, side) in enumerate(breakpoint):
                if side == 1:
                    at_boundary[n].add((bp_idx, be_idx, +1))
                elif side == 0:
                    at_boundary[n - 1].add((bp_idx, be_idx, -1))

        rev, tel, bidx, borient = [], [], [], []
        fwd = np.zeros(self.N, dtype=int)
        for n in range(-1, self.N):
            if n >= 0:
                fwd[n] = len(rev)
            linked = (n, n + 1) in adjacent
            if n in at_boundary:
                # iteration order of this set decides which breakend lands on which model
                # segment; it has to be the plain set order the reference iterates in
                for bp_idx, be_idx, orient in at_boundary[n]:
                    rev.append(n); tel.append(0); bidx.append(bp_idx); borient.append(orient)
                if not linked:
                    rev.append(n); tel.append(1); bidx.append(-1); borient.append(0)
            elif n >= 0:
                rev.append(n); tel.append(0 if linked else 1); bidx.append(-1); borient.append(0)

        self.N1 = len(rev)
        self.seg_fwd_remap = fwd
        self.seg_rev_remap = np.array(rev, dtype=int)
        self.seg_is_original = np.zeros(self.N1, dtype=bool)
        self.seg_is_original[fwd] = True
        self.num_breakpoints = len(self.breakpoints)
        self.is_telomere = np.array(tel, dtype=int)
        self.breakpoint_idx = np.array(bidx, dtype=int)
        self.breakpoint_orient = np.array(borient, dtype=int)

        assert not np.any((self.breakpoint_idx >= 0) & (self.is_telomere == 1))
        assert np.all(np.bincount(self.breakpoint_idx[self.breakpoint_idx >= 0]) == 2)

        self.x1 = np.zeros((self.N1, x.shape[1]), dtype=float)
        self.l1 = np.zeros((self.N1,), dtype=float)
        self.x1[fwd, :] = x
        self.l1[fwd] = l

    _SEG_FIELDS = ('N1', 'seg_fwd_remap', 'seg_rev_remap', 'seg_is_original', 'num_breakpoints', 'is_telomere',
                   'breakpoint_idx', 'breakpoint_orient', 'x1', 'l1')

    def _adopt_segmentation(self, donor):
        # arrays are shared read-only between the fits (disable_breakpoints rebinds, never writes in place)
        for f in self._SEG_FIELDS:
            setattr(self, f, getattr(donor, '_seg_' + f, getattr(donor, f)))

    # kept as methods for API compatibility with the reference class
    def create_cn_states(self, num_clones, num_alleles, cn_max, cn_diff_max):
        return create_cn_states(num_clones, num_alleles, cn_max, cn_diff_max)

    def create_brk_states(self, num_clones, cn_max, cn_diff_max):
        return create_brk_states(num_clones, cn_max, cn_diff_max)

    def get_likelihood_param_values(self):
        return dict((name, getattr(self.model, name)) for name in self.likelihood_params)

    # the reference's ``cdef public`` attributes that are plain values or per-segment / per-breakpoint arrays
    # (bpmodel.pyx:398-456); the three (N-1, S, S) arrays and ``framelogprob`` / ``posterior_marginals`` are listed
    # separately because a device model materialises them on demand (gigabytes at whole-genome size)
    _MODEL_DATA = ('num_clones', 'num_segments', 'num_breakpoints', 'num_alleles', 'cn_max', 'num_cn_states',
                   'num_brk_states', 'normal_contamination', 'transition_penalty', 'divergence_weight',
                   'transition_model', 'hmm_log_norm_const', 'prior_outlier_total', 'prior_outlier_allele',
                   'negbin_r_0', 'negbin_r_1', 'negbin_hdel_mu', 'negbin_hdel_r_0', 'negbin_hdel_r_1',
                   'betabin_M_0', 'betabin_M_1', 'betabin_loh_p', 'betabin_loh_M_0', 'betabin_loh_M_1',
                   'h', 'l', 'x', 'y', 'brk_states', 'is_telomere', 'breakpoint_idx', 'breakpoint_orient',
                   'p_breakpoint', 'p_allele_swap', 'p_outlier_total', 'p_outlier_allele',
                   'total_likelihood_mask', 'allele_likelihood_mask')
    _MODEL_DATA_LARGE = ('framelogprob', 'posterior_marginals', 'log_transmat', 'cached_log_transmat',
                         'joint_posterior_marginals')

    def get_model_data(self, large=False):
        """The model's public state as a dict (``cn_model.py:286-297``, which walks ``dir(model)``).  On a device model
        that walk would materialise three (N-1, S, S) arrays (2.3 GB each at whole-genome size) and trip over
        properties that are undefined before ``infer_cn``; here the attributes are named, and the large ones are
        included only on request (``large=True``)."""
        data = {}
        for a in self._MODEL_DATA + (self._MODEL_DATA_LARGE if large else ()):
            if not hasattr(type(self.model), a) and not hasattr(self.model, a):
                continue
            value = getattr(self.model, a)
            data[a] = np.asarray(value) if type(value).__name__ == '_memoryviewslice' else value
        return data

    def _state_weights(self, indicator):
        return (np.asarray(self.model.posterior_marginals) * (np.asarray(indicator) == 1)).sum(axis=-1)

    def get_param_sample_weight(self, name):
        """Per-segment resampling weight for the stochastic M-step of ``name`` (``cn_model.py:323-352``)."""
        model = self.model
        if name in ('negbin_r_0', 'negbin_r_1'):
            weights = np.asarray(model.p_outlier_total)[:, int(name[-1])]
        elif name in ('betabin_M_0', 'betabin_M_1'):
            weights = np.asarray(model.p_outlier_allele)[:, int(name[-1])]
        elif name == 'negbin_hdel_mu':
            weights = self._state_weights(model.is_hdel)
        elif name in ('negbin_hdel_r_0', 'negbin_hdel_r_1'):
            weights = self._state_weights(model.is_hdel) * np.asarray(model.p_outlier_total)[:, int(name[-1])]
        elif name == 'betabin_loh_p':
            weights = self._state_weights(model.is_loh)
        elif name in ('betabin_loh_M_0', 'betabin_loh_M_1'):
            weights = self._state_weights(model.is_loh) * np.asarray(model.p_outlier_allele)[:, int(name[-1])]
        else:
            raise KeyError(name)
        norm = weights.sum()
        if norm > 0.:
            return weights / norm
        if self.verbose:
            print('nothing for ' + name)
        return None

    def model_args(self, h_init, compact=False):
        """Positional arguments of the ``RemixtModel`` constructor for this segmentation
        (``cn_model.py:359-384``); ``h_init`` is ``(M,)``, or ``(G, M)`` for a ``RemixtBatch``."""
        M = np.asarray(h_init).shape[-1]
        table = create_cn_states(M, 2, self.max_copy_number, self.max_copy_number_diff)
        brk_states = create_brk_states(M, self.max_copy_number, self.max_copy_number_diff)
        # (N1, S, M, 2): the shared state table with clone 0 replaced by the segment's germline
        # copies, rows taken through the remap (index -1 wraps to the last segment, cn_model.py:364)
        normal = np.asarray(self.normal_copies)[self.seg_rev_remap]
        if compact:     # the same array as its two factors, for model classes that take them (bpmodel.CompactCnStates)
            from remixt_b200.bpmodel import CompactCnStates
            cn_states = CompactCnStates(table, normal)
        else:
            cn_states = np.empty((self.N1,) + table.shape, dtype=table.dtype)
            cn_states[:] = table[np.newaxis]
            cn_states[:, :, 0, :] = normal[:, np.newaxis, :]
        return (M, self.N1, self.num_breakpoints, self.normal_contamination,
                cn_states, brk_states, h_init,
                self.l1, self.x1[:, 2], self.x1[:, 0:2],
                self.is_telomere, self.breakpoint_idx, self.breakpoint_orient,
                self.transition_log_prob)

    def _make_model(self, h_init):
        compact = bool(getattr(self.model_class, 'accepts_compact_cn_states', False))
        return self.model_class(*(self.model_args(h_init, compact=compact) + (self.divergence_weight,)),
                                **self.model_kwargs)

    def attach(self, model):
        """Adopt ``model`` (a fresh ``RemixtModel``, or a view of one fit of a ``RemixtBatch``) and apply the
        per-fit settings ``fit`` applies after construction (``cn_model.py:386-404``)."""
        self.model = model
        self.model.total_likelihood_mask = self._total_likelihood_mask.astype(int)
        self.model.allele_likelihood_mask = self._allele_likelihood_mask.astype(int)

        if self.breakpoint_init is not None:
            brk_states = np.array(self.model.brk_states)
            p_breakpoint = np.ones((self.model.num_breakpoints, self.model.num_brk_states))
            for k, bp in enumerate(self.breakpoints):
                match = np.all(brk_states == self.breakpoint_init[bp], axis=1)
                p_breakpoint[k, match] = 1000.
            p_breakpoint /= np.sum(p_breakpoint, axis=-1)[:, np.newaxis]
            self.model.p_breakpoint = p_breakpoint

        self.model.transition_model = self.transition_model

    def prepare(self, h_init):
        """Everything ``fit`` does before its EM loop (``cn_model.py:354-407``): build the ``RemixtModel``,
        set masks and transition model, evaluate the first ELBO."""
        self.attach(self._make_model(h_init))
        if self.prev_elbo is None:
            self.prev_elbo = self.model.calculate_elbo()

    def record_iteration(self, i, elbo):
        """Bookkeeping at the end of an EM iteration (``cn_model.py:418-428``)."""
        self.prev_elbo_diff = elbo - self.prev_elbo
        self.prev_elbo = elbo
        if self.verbose:
            print('[{}] completed iteration {}'.format(_now(), i))
            print('[{}]     elbo: {:.10f}'.format(_now(), self.prev_elbo))
            print('[{}]     elbo diff: {:.10f}'.format(_now(), self.prev_elbo_diff))
            print('[{}]     h = {}'.format(_now(), np.asarray(self.model.h)))
            for name, value in self.get_likelihood_param_values().items():
                print('[{}]     {} = {}'.format(_now(), name, value))

    def fit(self, h_init):
        """Fit the model with a series of updates (``cn_model.py:354-428``)."""
        self.prepare(h_init)

        for i in range(self.num_em_iter):
            self.variational_update(self.num_update_iter)

            if self.do_h_update:
                self.em_update_h()

            self.em_update_params()

            self.record_iteration(i, self.model.calculate_elbo())

    @contextlib.contextmanager
    def elbo_check(self, name, threshold=-1e-6):
        if self.verbose:
            print('[{}] optimizing {}'.format(_now(), name))
        if not self.check_elbo:
            yield
            return
        elbo_before = self.model.calculate_elbo()
        yield
        elbo_after = self.model.calculate_elbo()
        if self.verbose:
            print('[{}]     elbo: {:.10f}'.format(_now(), elbo_after))
            print('[{}]     elbo diff: {:.10f}'.format(_now(), elbo_after - elbo_before))
        if elbo_after - elbo_before < threshold:
            raise Exception('elbo error for step {}!'.format(name))

    def variational_update(self, n_iter=1):
        """``n_iter`` rounds of the coordinate updates, in the reference's order (``cn_model.py:444-460``)."""
        if not self.check_elbo and hasattr(self.model, 'variational_update'):
            self.model.variational_update(n_iter)     # same updates, one device synchronisation
            return
        for _ in range(n_iter):
            self._variational_update_stepwise()

    def _variational_update_stepwise(self):
        for name, step in (('update_p_allele_swap', 'update_p_allele_swap'), ('p_cn', 'update_p_cn'),
                           ('p_breakpoint', 'update_p_breakpoint'), ('p_outlier_total', 'update_p_outlier_total'),
                           ('p_outlier_allele', 'update_p_outlier_allele')):
            with self.elbo_check(name):
                getattr(self.model, step)()

    def em_update_h(self):
        with self.elbo_check('h'):
            self.update_h()

    def em_update_params(self):
        for name in self.likelihood_params:
            with self.elbo_check(name):
                self.update_param(name)

    def _create_sample(self, weights=None):
        # consumes the global numpy.random stream exactly as cn_model.py:475-480
        sample_size = int(min(200, self.model.num_segments / 10))
        rs = self.random_state if self.random_state is not None else np.random
        chosen = rs.choice(self.model.num_segments, size=sample_size, replace=False, p=weights)
        sample = np.zeros((self.model.num_segments,), dtype=int)
        sample[chosen] = 1
        return sample

    def _all_segments(self):
        return np.ones((self.model.num_segments,), dtype=int)

    def update_h(self):
        """M-step for the haploid depths: L-BFGS-B on a segment subsample, accepted only if the
        full-data expected log likelihood does not drop (``cn_model.py:482-531``)."""
        model = self.model

        def nll(h, model, sample):
            model.h = h
            return -model.calculate_expected_log_likelihood(sample)

        def nll_grad(h, model, sample):
            model.h = h
            g = np.zeros((model.num_clones,))
            model.calculate_expected_log_likelihood_partial_h(sample, g)
            return -g

        h_before = model.h
        ell_before = model.calculate_expected_log_likelihood(self._all_segments())
        sample = self._create_sample()

        result = scipy.optimize.minimize(
            nll, model.h, method='L-BFGS-B', jac=nll_grad,
            bounds=[(1e-8, 10.)] * model.num_clones, args=(model, sample))

        if not result.success:
            message = result.message.decode() if isinstance(result.message, bytes) else str(result.message)
            if message == 'ABNORMAL_TERMINATION_IN_LNSRCH':
                analytic = nll_grad(result.x, model, sample)
                numerical = _numerical_gradient(result.x, nll, args=(model, sample))
                if not np.allclose(analytic, numerical, atol=2.):
                    raise ValueError('gradiant error, analytic: {}, numerical: {}\n'.format(analytic, numerical))
            else:
                raise ValueError('optimization failed\n{}'.format(result))

        model.h = result.x
        ell_after = model.calculate_expected_log_likelihood(self._all_segments())
        self.decisions.append(('h', not (ell_after < ell_before), ell_before, ell_after))
        if ell_after < ell_before:
            if self.verbose:
                print('[{}] h rejected, elbo before: {}, after: {}'.format(_now(), ell_before, ell_after))
            model.h = h_before
        else:
            model.h = result.x

    def update_param(self, name):
        """M-step for one dispersion parameter: brute-force grid + simplex polish on a weighted
        segment subsample (``cn_model.py:533-569``).  As in the reference, the acceptance test runs
        with whatever value the optimiser evaluated last, not with its returned optimum."""
        model = self.model
        bounds = self.likelihood_param_bounds[name]
        weights = self.get_param_sample_weight(name)

        def nll(value, model, name, bounds, sample):
            assert value.shape == (1,)
            value = float(value[0])
            if value < bounds[0] or value > bounds[1]:
                return np.inf
            setattr(model, name, value)
            return -model.calculate_expected_log_likelihood(sample)

        value_before = getattr(model, name)
        ell_before = model.calculate_expected_log_likelihood(self._all_segments())
        sample = self._create_sample(weights)

        result = scipy.optimize.brute(nll, args=(model, name, bounds, sample), ranges=[bounds], full_output=True)
        assert result[0].shape == (1,)
        result_value = float(result[0][0])

        ell_after = model.calculate_expected_log_likelihood(self._all_segments())
        self.decisions.append((name, not (ell_after < ell_before), ell_before, ell_after))
        if ell_after < ell_before:
            if self.verbose:
                print('[{}] {} rejected, elbo before: {}, after: {}'.format(_now(), name, ell_before, ell_after))
            setattr(model, name, value_before)
        else:
            setattr(model, name, result_value)

    def decode_breakpoints(self, cn):
        """Breakpoint copy number given the MAP segment copy number of the model segments (``cn_model.py:578-596``)."""
        model = self.model
        brk_states = np.asarray(model.brk_states)
        bidx = np.asarray(model.breakpoint_idx)
        borient = np.asarray(model.breakpoint_orient)
        penalty = model.transition_penalty
        total = cn.sum(axis=-1)
        log_p = np.zeros((model.num_breakpoints, model.num_brk_states))
        for n in np.flatnonzero(bidx[:model.num_segments - 1] >= 0):
            for m in range(model.num_clones):
                d = total[n, m] - total[n + 1, m]
                log_p[bidx[n], :] += -penalty * abs(d - borient[n] * brk_states[:, m])

        brk_cn = dict()
        for k in range(model.num_breakpoints):
            brk_cn[self.breakpoint_ids[k]] = brk_states[log_p[k, :].argmax()]
        return brk_cn

    def optimal_cn(self):
        """MAP clone copy number per original segment and per breakpoint (``cn_model.py:571-598``)."""
        model = self.model
        cn = np.zeros((model.num_segments, model.num_clones, model.num_alleles), dtype=int)
        model.infer_cn(cn)
        return cn[self.seg_fwd_remap], self.decode_breakpoints(cn)

    def breakpoint_prob(self):
        return dict(zip(self.breakpoints, np.asarray(self.model.p_breakpoint)))

    @property
    def h(self):
        return np.asarray(self.model.h)

    @property
    def p_breakpoint(self):
        return np.asarray(self.model.p_breakpoint)

    @property
    def p_outlier_total(self):
        return np.asarray(self.model.p_outlier_total)[self.seg_fwd_remap]

    @property
    def p_outlier_allele(self):
        return np.asarray(self.model.p_outlier_allele)[self.seg_fwd_remap]

    @property
    def total_likelihood_mask(self):
        return np.asarray(self.model.total_likelihood_mask)[self.seg_fwd_remap]

    @property
    def allele_likelihood_mask(self):
        return np.asarray(self.model.allele_likelihood_mask)[self.seg_fwd_remap]


def decode_breakpoints_naive(cn, adjacencies, breakpoints):
    """Breakpoint copy number from total copy number steps alone (``cn_model.py:631-687``):
    the copies a breakend carries are what its segment has in excess of the segment across the
    wild-type adjacency (floored at zero); a breakpoint carries the smaller of its two ends."""
    total = cn.sum(axis=-1)
    across = {}
    for a, b in adjacencies:
        across[(a, 1)] = b
        across[(b, 0)] = a

    def excess(breakend):
        n, _ = breakend
        other = total[across[breakend], :] if breakend in across else 0
        return np.maximum(total[n, :] - other, 0)

    brk_cn = {}
    for breakpoint_id, breakpoint in breakpoints.items():
        end_1, end_2 = breakpoint
        brk_cn[breakpoint_id] = np.minimum(excess(end_1), excess(end_2))
    return brk_cn
