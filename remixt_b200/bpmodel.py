"""Drop-in replacement for the reference's Cython module ``remixt.bpmodel``.

Same public surface as ``remixt/bpmodel.pyx``: class :class:`RemixtModel` (constructor
``bpmodel.pyx:461-476``, ``cdef public`` attributes ``:398-456``, ``cpdef`` methods
``:640-1210``) and the module functions :func:`sum_product` (``:1213``),
:func:`sum_product_2paramtrans` (``:1249``) and :func:`max_product` (``:1296``).

Every method forwards through the C-ABI of ``libremixt_b200.so`` to sm_100a CUDA kernels;
the model's arrays live in HBM and attribute reads copy them back on demand.  There is no
CPU implementation behind this class.

Differences from the reference that a caller can observe, all deliberate:

* Array attributes are returned as fresh ``numpy`` arrays (device -> host copies), not as
  memoryviews onto caller-owned buffers; assignments copy host -> device immediately.
* ``cn_states`` must be one state table shared by all segments plus a per-segment normal row
  (``cn_states[n, s, 1:, :]`` independent of ``n`` and ``cn_states[n, s, 0, :]`` independent of
  ``s``), which is how ``cn_model.BreakpointModel.fit`` builds it (``cn_model.py:359-364``).
* ``log_transmat`` / ``cached_log_transmat`` / ``joint_posterior_marginals`` are never stored;
  reading them materialises ``(N-1, S, S)`` arrays on demand (small problems / tests).
"""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import lib, check

#: CUDA device ordinal new models are created on (override per model with ``device=``).
default_device = int(os.environ.get('REMIXT_B200_DEVICE', os.environ.get('LOCAL_RANK', '0')))


def set_default_device(device):
    global default_device
    default_device = int(device)


class CompactCnStates(object):
    """``cn_states`` ``(N1, S, M, 2)`` held as the two factors the reference builds it from (``cn_model.py:360-364``):
    the state table ``(S, M, 2)`` shared by every segment and the per-segment normal copies ``(N1, 2)`` that replace
    clone 0.  Converts to the full int64 array on demand (``np.asarray``); the CUDA models take the factors as they
    are (``rmx_desc.cn_table`` / ``normal_copies``), which spares building, scanning and copying 140 MB per
    whole-genome fit."""
    ndim = 4
    dtype = np.dtype(np.int64)

    def __init__(self, table, normal_copies):
        self.table = np.ascontiguousarray(table, dtype=np.int64)
        self.normal_copies = np.ascontiguousarray(normal_copies, dtype=np.int64)
        if self.table.ndim != 3 or self.normal_copies.ndim != 2 or self.normal_copies.shape[1] != self.table.shape[2]:
            raise ValueError('table must be (S, M, num_alleles) and normal_copies (N1, num_alleles)')
        self.shape = (self.normal_copies.shape[0],) + self.table.shape

    def __array__(self, dtype=None, copy=None):
        a = np.empty(self.shape, dtype=np.int64)
        a[:] = self.table[np.newaxis]
        a[:, :, 0, :] = self.normal_copies[:, np.newaxis, :]
        return a if dtype is None else a.astype(dtype)

    def max(self):
        return max(int(self.table[:, 1:, :].max()) if self.table.shape[1] > 1 else 0, int(self.normal_copies.max()))

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return CompactCnStates(self.table, self.normal_copies[idx])
        return np.asarray(self)[idx]


def _typed(a, dtype, ndim, name):
    # Cython's typed buffer arguments reject other dtypes / ranks (bpmodel.pyx:466-474)
    if not isinstance(a, np.ndarray):
        raise TypeError('Argument %r has incorrect type (expected numpy.ndarray, got %s)' % (name, type(a).__name__))
    if a.dtype != dtype:
        raise ValueError("Buffer dtype mismatch, expected '%s' but got '%s'" % (np.dtype(dtype).name, a.dtype.name))
    if a.ndim != ndim:
        raise ValueError('Buffer has wrong number of dimensions (expected %d, got %d)' % (ndim, a.ndim))
    return np.ascontiguousarray(a)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class _ModelBase(object):
    accepts_compact_cn_states = True

    """Shared plumbing of :class:`RemixtModel` (one fit) and :class:`RemixtBatch` (G fits in one set of
    launches).  ``_gn`` is the number of fits the handle covers; with ``_single`` the leading fit
    dimension is dropped so the object looks exactly like the reference's ``RemixtModel``."""

    _handle = None
    _single = True
    _gn = 1
    _parent = None

    def _setup(self, num_clones, num_segments, num_breakpoints, normal_contamination,
               cn_states, brk_states, h_init, l, x, y, is_telomere, breakpoint_idx,
               breakpoint_orient, transition_penalty, divergence_weight, device, num_fits, shard=False):
        compact = isinstance(cn_states, CompactCnStates)
        if not compact:
            cn_states = _typed(cn_states, np.int64, 4, 'cn_states')
        brk_states = _typed(brk_states, np.int64, 2, 'brk_states')
        h_init = _typed(h_init, np.float64, 1 if num_fits is None else 2, 'h_init')
        l = _typed(l, np.float64, 1, 'l')
        x = _typed(x, np.float64, 1, 'x')
        y = _typed(y, np.float64, 2, 'y')
        is_telomere = _typed(is_telomere, np.int64, 1, 'is_telomere')
        breakpoint_idx = _typed(breakpoint_idx, np.int64, 1, 'breakpoint_idx')
        breakpoint_orient = _typed(breakpoint_orient, np.int64, 1, 'breakpoint_orient')

        self.num_clones = int(num_clones)
        self.num_segments = int(num_segments)
        self.num_breakpoints = int(num_breakpoints)
        self.normal_contamination = bool(normal_contamination)
        self.num_alleles = 2
        self.cn_states = cn_states
        self.brk_states = brk_states
        self.cn_max = int(max(cn_states.max(), brk_states.max()))       # bpmodel.pyx:489
        self.num_cn_states = cn_states.shape[1]
        self.num_brk_states = brk_states.shape[0]

        # shape validation, messages as bpmodel.pyx:509-529
        if ((cn_states.shape[0] != num_segments) or (cn_states.shape[2] != num_clones) or
                (cn_states.shape[3] != self.num_alleles)):
            raise ValueError('cn_states must have shape (num_segments, num_cn_states, num_clones, num_alleles)')
        if brk_states.shape[1] != num_clones:
            raise ValueError('cn_states must have shape (num_brk_states, num_clones)')
        if h_init.shape[-1] != num_clones:
            raise ValueError('h must have length equal to num_clones')
        if is_telomere.shape[0] != num_segments:
            raise ValueError('is_telomere must have length equal to num_segments')
        if breakpoint_idx.shape[0] != num_segments:
            raise ValueError('breakpoint_idx must have length equal to num_segments')
        if breakpoint_orient.shape[0] != num_segments:
            raise ValueError('breakpoint_orient must have length equal to num_segments')
        # a shard of a chromosome-sharded fit (remixt_b200/sharded.py) sees the breakends of its own chromosomes only
        if (breakpoint_idx.max() + 1 > num_breakpoints) if shard else (breakpoint_idx.max() + 1 != num_breakpoints):
            raise ValueError('breakpoint_idx must have maximum of num_breakpoints positive indices')
        if l.shape[0] != num_segments or x.shape[0] != num_segments or y.shape != (num_segments, 2):
            raise ValueError('l, x, y must have num_segments rows')

        self.is_telomere = is_telomere
        self.breakpoint_idx = breakpoint_idx
        self.breakpoint_orient = breakpoint_orient
        self.device = default_device if device is None else int(device)

        desc = _lib.RmxDesc()
        desc.num_clones = self.num_clones
        desc.num_segments = self.num_segments
        desc.num_breakpoints = self.num_breakpoints
        desc.normal_contamination = 1 if self.normal_contamination else 0
        desc.num_cn_states = self.num_cn_states
        desc.num_brk_states = self.num_brk_states
        desc.device = self.device
        desc.flags = _lib.DESC_SHARD if shard else 0
        if compact:
            desc.cn_states = None
            desc.cn_table = _ptr(cn_states.table)
            desc.normal_copies = _ptr(cn_states.normal_copies)
        else:
            desc.cn_states = _ptr(cn_states)
            desc.cn_table = None
            desc.normal_copies = None
        desc.brk_states = _ptr(brk_states)
        desc.h_init = _ptr(h_init)
        desc.l = _ptr(l)
        desc.x = _ptr(x)
        desc.y = _ptr(y)
        desc.is_telomere = _ptr(is_telomere)
        desc.breakpoint_idx = _ptr(breakpoint_idx)
        desc.breakpoint_orient = _ptr(breakpoint_orient)
        desc.transition_penalty = float(transition_penalty)
        handle = C.c_void_p()
        if num_fits is None:
            desc.divergence_weight = float(divergence_weight)
            rc = lib.rmx_create(C.byref(desc), C.byref(handle))
        else:
            dw = np.ascontiguousarray(np.broadcast_to(np.asarray(divergence_weight, dtype=np.float64), (num_fits,)))
            if h_init.shape[0] != num_fits:
                raise ValueError('h_init must have one row per fit')
            desc.divergence_weight = 0.
            rc = lib.rmx_batch_create(C.byref(desc), int(num_fits), _ptr(h_init), _ptr(dw), C.byref(handle))
        check(rc, None)
        self._handle = handle
        self._derived = {}

    def close(self):
        """Releases the device state now (its blocks go back to the pool) instead of when the object is collected;
        the per-fit views of a batch hold a reference cycle with it, so collection can be late."""
        for view in list(getattr(self, '_views', {}).values()):
            view.close()
        if getattr(self, '_views', None):
            self._views.clear()
        h, self._handle = self._handle, None
        if h is not None and lib is not None:
            try:
                lib.rmx_destroy(h)
            except Exception:
                pass

    def __del__(self):
        self.close()

    # ------------------------------------------------------------------ attribute plumbing
    def _out(self, a):
        return a[0] if self._single else a

    def _get(self, array_id, shape, dtype=np.float64):
        out = np.empty((self._gn,) + tuple(shape), dtype=dtype)
        check(lib.rmx_get_array(self._handle, array_id, _ptr(out), out.size), self._handle)
        return self._out(out)

    def _set(self, array_id, value, shape, dtype=np.float64):
        a = np.asarray(value)
        full = (self._gn,) + tuple(shape)
        if a.shape != tuple(shape) and a.shape != full:
            raise ValueError('array of shape %s expected, got %s' % (tuple(shape), a.shape))
        # one attribute's worth of elements on a batch is broadcast to its fits by the library
        a = np.ascontiguousarray(a, dtype=dtype)
        check(lib.rmx_set_array(self._handle, array_id, _ptr(a), a.size), self._handle)

    def _get_scalar(self, name):
        v = np.empty((self._gn,), dtype=np.float64)
        check(lib.rmx_get_scalar(self._handle, _lib.SCALAR_IDS[name], _ptr(v)), self._handle)
        return float(v[0]) if self._single else v

    def _scalars_out(self, fn, *args):
        v = np.empty((self._gn,), dtype=np.float64)
        check(fn(self._handle, *(args + (_ptr(v),))), self._handle)
        return float(v[0]) if self._single else v

    def _set_scalar(self, name, value):
        check(lib.rmx_set_scalar(self._handle, _lib.SCALAR_IDS[name], float(value)), self._handle)

    def _get_derived(self, which, shape):
        if which not in self._derived:
            out = np.empty(shape, dtype=np.int64)
            check(lib.rmx_get_derived(self._handle, which, _ptr(out), out.size), self._handle)
            self._derived[which] = out
        return self._derived[which]

    @property
    def _NS(self):
        return (self.num_segments, self.num_cn_states)

    @property
    def _NSS(self):
        return (self.num_segments - 1, self.num_cn_states, self.num_cn_states)

    h = property(lambda self: self._get(_lib.A_H, (self.num_clones,)),
                 lambda self, v: self._set(_lib.A_H, v, (self.num_clones,)))
    p_breakpoint = property(
        lambda self: self._get(_lib.A_P_BREAKPOINT, (self.num_breakpoints, self.num_brk_states)),
        lambda self, v: self._set(_lib.A_P_BREAKPOINT, v, (self.num_breakpoints, self.num_brk_states)))
    p_allele_swap = property(lambda self: self._get(_lib.A_P_ALLELE_SWAP, (self.num_segments, 2)),
                             lambda self, v: self._set(_lib.A_P_ALLELE_SWAP, v, (self.num_segments, 2)))
    p_outlier_total = property(lambda self: self._get(_lib.A_P_OUTLIER_TOTAL, (self.num_segments, 2)),
                               lambda self, v: self._set(_lib.A_P_OUTLIER_TOTAL, v, (self.num_segments, 2)))
    p_outlier_allele = property(lambda self: self._get(_lib.A_P_OUTLIER_ALLELE, (self.num_segments, 2)),
                                lambda self, v: self._set(_lib.A_P_OUTLIER_ALLELE, v, (self.num_segments, 2)))
    posterior_marginals = property(lambda self: self._get(_lib.A_POSTERIOR_MARGINALS, self._NS),
                                   lambda self, v: self._set(_lib.A_POSTERIOR_MARGINALS, v, self._NS))
    framelogprob = property(lambda self: self._get(_lib.A_FRAMELOGPROB, self._NS),
                            lambda self, v: self._set(_lib.A_FRAMELOGPROB, v, self._NS))
    total_likelihood_mask = property(
        lambda self: self._get(_lib.A_TOTAL_LIKELIHOOD_MASK, (self.num_segments,), np.int64),
        lambda self, v: self._set(_lib.A_TOTAL_LIKELIHOOD_MASK, v, (self.num_segments,), np.int64))
    allele_likelihood_mask = property(
        lambda self: self._get(_lib.A_ALLELE_LIKELIHOOD_MASK, (self.num_segments,), np.int64),
        lambda self, v: self._set(_lib.A_ALLELE_LIKELIHOOD_MASK, v, (self.num_segments,), np.int64))
    log_transmat = property(lambda self: self._get(_lib.A_LOG_TRANSMAT, self._NSS))
    cached_log_transmat = property(lambda self: self._get(_lib.A_CACHED_LOG_TRANSMAT, self._NSS))
    joint_posterior_marginals = property(lambda self: self._get(_lib.A_JOINT_POSTERIOR_MARGINALS, self._NSS))
    l = property(lambda self: self._get(_lib.A_L, (self.num_segments,)))
    x = property(lambda self: self._get(_lib.A_X, (self.num_segments,)))
    y = property(lambda self: self._get(_lib.A_Y, (self.num_segments, 2)))

    cn_states_total = property(
        lambda self: self._get_derived(0, (self.num_segments, self.num_cn_states, self.num_clones)))
    num_alleles_subclonal = property(lambda self: self._get_derived(1, self._NS))
    is_hdel = property(lambda self: self._get_derived(2, self._NS))
    is_loh = property(lambda self: self._get_derived(3, self._NS))

    @property
    def breakpoint_side(self):
        # bpmodel.pyx:537-544 (computed by the reference, used nowhere)
        side = np.zeros((self.num_segments,), dtype=np.int64)
        seen = np.zeros((max(self.num_breakpoints, 1),), dtype=np.int64)
        for n in np.flatnonzero(self.breakpoint_idx >= 0):
            k = self.breakpoint_idx[n]
            side[n] = seen[k]
            seen[k] += 1
        return side

    @property
    def transition_model(self):
        v = self._get_scalar('transition_model')
        return int(v) if self._single else v.astype(int)

    @transition_model.setter
    def transition_model(self, value):
        self._set_scalar('transition_model', int(value))

    # ------------------------------------------------------------------ coordinate updates
    def update_framelogprob(self):
        check(lib.rmx_update(self._handle, _lib.U_FRAMELOGPROB), self._handle)

    def update_p_cn(self):
        check(lib.rmx_update(self._handle, _lib.U_CN), self._handle)

    def update_p_breakpoint(self):
        check(lib.rmx_update(self._handle, _lib.U_BREAKPOINT), self._handle)

    def update_p_outlier_total(self):
        check(lib.rmx_update(self._handle, _lib.U_OUTLIER_TOTAL), self._handle)

    def update_p_outlier_allele(self):
        check(lib.rmx_update(self._handle, _lib.U_OUTLIER_ALLELE), self._handle)

    def update_p_allele_swap(self):
        check(lib.rmx_update(self._handle, _lib.U_ALLELE_SWAP), self._handle)

    def variational_update(self, n_iter=1):
        """``n_iter`` rounds of the five coordinate updates in the order of ``cn_model.py:444-460``
        with a single host synchronisation (extension; the reference drives them one by one)."""
        check(lib.rmx_variational_update(self._handle, int(n_iter)), self._handle)

    # ------------------------------------------------------------------ objective
    def calculate_variational_entropy(self):
        return self._scalars_out(lib.rmx_variational_entropy)

    def calculate_variational_energy(self):
        return self._scalars_out(lib.rmx_variational_energy)

    def calculate_elbo(self):
        return self._scalars_out(lib.rmx_elbo)

    def _sample(self, sample):
        s = np.asarray(sample)
        if s.shape != (self.num_segments,) and s.shape != (self._gn, self.num_segments):
            raise ValueError('sample must have length equal to num_segments')
        return np.ascontiguousarray(np.broadcast_to(s, (self._gn, self.num_segments)), dtype=np.int64)

    def calculate_expected_log_likelihood(self, sample):
        s = self._sample(sample)
        return self._scalars_out(lib.rmx_expected_log_likelihood, _ptr(s))

    def calculate_expected_log_likelihood_partial_h(self, sample, partial_h):
        s = self._sample(sample)
        out = np.zeros((self._gn, self.num_clones), dtype=np.float64)
        check(lib.rmx_expected_log_likelihood_partial_h(self._handle, _ptr(s), _ptr(out)), self._handle)
        partial_h[...] = self._out(out)

    def calculate_expected_log_likelihood_grid(self, sample, name, values):
        """E[LL](sample) for each value of the named dispersion parameter, one synchronisation
        (extension used by the batched M-step; see ``rmx_expected_log_likelihood_grid``).  ``values`` is
        ``(n,)`` (every fit evaluates the same grid) or ``(gn, n)``."""
        s = self._sample(sample)
        vals = np.asarray(values, dtype=np.float64)
        n = vals.shape[-1]
        vals = np.ascontiguousarray(np.broadcast_to(vals, (self._gn, n)))
        out = np.zeros(vals.shape, dtype=np.float64)
        check(lib.rmx_expected_log_likelihood_grid(self._handle, _ptr(s), _lib.SCALAR_IDS[name], _ptr(vals),
                                                   n, _ptr(out)), self._handle)
        return self._out(out)

    # ------------------------------------------------------------------ batched M-step (RemixtBatch handles)
    def mstep_stage_sample(self, sample=None):
        """Uploads the M-step sample of every fit once (``rmx_mstep_stage_sample``); ``sample`` is the ``(gn, N1)``
        mask of ``calculate_expected_log_likelihood`` or ``None`` for every segment."""
        if sample is None:
            check(lib.rmx_mstep_stage_sample(self._handle, None), self._handle)
        else:
            s = self._sample(sample)
            check(lib.rmx_mstep_stage_sample(self._handle, _ptr(s)), self._handle)

    def mstep_stage_rows(self, rows, counts):
        """The same from row lists (``rmx_mstep_stage_rows``): ``rows`` ``(gn, R)`` segment indices, fit ``g`` uses
        its first ``counts[g]``."""
        r = np.ascontiguousarray(rows, dtype=np.int32).reshape(self._gn, -1)
        c = np.ascontiguousarray(counts, dtype=np.int32).reshape(self._gn)
        check(lib.rmx_mstep_stage_rows(self._handle, _ptr(r), _ptr(c), r.shape[1]), self._handle)

    def mstep_evaluate(self, active=None, h=None, name=None, values=None, grad=False):
        """E[LL] of the staged sample for the active fits with one synchronisation (``rmx_mstep_evaluate``):
        ``h`` ``(gn, M)`` is assigned first; with ``name`` the dispersion parameter takes ``values`` ``(gn, n)`` in
        turn.  Returns ``ell (gn, n)`` and, with ``grad``, ``(ell, grad (gn, M))``."""
        gn = self._gn
        act = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        hv = None if h is None else np.ascontiguousarray(h, dtype=np.float64).reshape(gn, self.num_clones)
        if name is None:
            sid, vals, n = -1, None, 1
        else:
            vals = np.ascontiguousarray(np.asarray(values, dtype=np.float64).reshape(gn, -1))
            sid, n = _lib.SCALAR_IDS[name], vals.shape[1]
        ell = np.zeros((gn, n), dtype=np.float64)
        g = np.zeros((gn, self.num_clones), dtype=np.float64) if grad else None
        check(lib.rmx_mstep_evaluate(self._handle, None if act is None else _ptr(act), None if hv is None else _ptr(hv),
                                     sid, None if vals is None else _ptr(vals), n, _ptr(ell),
                                     None if g is None else _ptr(g)), self._handle)
        return (ell, g) if grad else ell

    def mstep_param_search(self, name, lo, hi, grid, active=None, trace=0):
        """``scipy.optimize.brute(grid) + fmin`` of ``update_param(name)`` (``cn_model.py:553-558``) for every active fit
        inside the library (``rmx_mstep_param_search``).  Returns ``(optimum (gn,), (rounds, evaluations))`` and, with
        ``trace`` > 0, also the per-fit fmin evaluations as a list of ``(k, 2)`` arrays of (x, f)."""
        gn = self._gn
        act = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        grid = np.ascontiguousarray(grid, dtype=np.float64).ravel()
        result = np.zeros(gn, dtype=np.float64)
        evals = np.zeros(gn, dtype=np.int32)
        rounds = np.zeros(2, dtype=np.int64)
        tr = np.zeros((gn, int(trace), 2), dtype=np.float64) if trace else None
        check(lib.rmx_mstep_param_search(self._handle, None if act is None else _ptr(act), _lib.SCALAR_IDS[name],
                                         float(lo), float(hi), _ptr(grid), int(grid.shape[0]), _ptr(result), _ptr(evals),
                                         _ptr(rounds), None if tr is None else _ptr(tr), int(trace)), self._handle)
        stats = (int(rounds[0]), int(rounds[1]))
        if trace:
            return result, stats, [tr[g, :min(int(evals[g]), int(trace))].copy() for g in range(gn)]
        return result, stats

    def set_scalar_array(self, name, values, active=None):
        """``setattr(model, name, values[g])`` for every (active) fit of the handle."""
        v = np.ascontiguousarray(values, dtype=np.float64).reshape(self._gn)
        act = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        check(lib.rmx_set_scalar_array(self._handle, _lib.SCALAR_IDS[name], None if act is None else _ptr(act), _ptr(v)),
              self._handle)

    # ------------------------------------------------------------------ decode / inspection
    def infer_cn(self, cn):
        shape = (self.num_segments, self.num_clones, self.num_alleles)
        if not self._single:
            shape = (self._gn,) + shape
        if not (isinstance(cn, np.ndarray) and cn.dtype == np.int64 and cn.flags.c_contiguous and cn.shape == shape):
            raise ValueError('cn must be a C-contiguous int64 array of shape %s' % (shape,))
        check(lib.rmx_infer_cn(self._handle, _ptr(cn)), self._handle)

    @property
    def state_sequence(self):
        return self._get(_lib.A_STATE_PATH, (self.num_segments,), np.int64)

    # ------------------------------------------------------------------ multi-GPU shards (remixt_b200/sharded.py)
    def comm_init(self, unique_id, rank, nranks):
        """Attaches an NCCL communicator to this shard (``rmx_comm_init``): from here on ``update_p_breakpoint``,
        ``hmm_log_norm_const``, the ELBO and the M-step objectives are those of the whole fit, reduced in stream."""
        buf = np.frombuffer(bytes(unique_id), dtype=np.uint8).copy()
        if buf.size != 128:
            raise ValueError('unique_id must be the 128 bytes of nccl_unique_id()')
        check(lib.rmx_comm_init(self._handle, _ptr(buf), int(rank), int(nranks)), self._handle)

    def comm_destroy(self):
        check(lib.rmx_comm_destroy(self._handle), self._handle)

    # update_p_breakpoint split around its reduction (chromosome-sharded fits, remixt_b200/sharded.py)
    def breakpoint_logq(self):
        """(K, B) partial ``log q`` of ``update_p_breakpoint`` (``bpmodel.pyx:971-980``) over this model's own
        breakend rows, before the exp-normalise."""
        out = np.zeros((self._gn, self.num_breakpoints, self.num_brk_states), dtype=np.float64)
        check(lib.rmx_breakpoint_logq(self._handle, _ptr(out)), self._handle)
        return self._out(out)

    def apply_breakpoint_logq(self, logq):
        """``p_breakpoint = exp_normalize(logq)`` and the cached transition expectation
        (``bpmodel.pyx:982-985``); ``logq`` is the sum of the shards' :meth:`breakpoint_logq`."""
        shape = (self._gn, self.num_breakpoints, self.num_brk_states)
        a = np.ascontiguousarray(np.broadcast_to(np.asarray(logq, dtype=np.float64).reshape(
            shape if np.size(logq) == np.prod(shape) else shape[1:]), shape))
        check(lib.rmx_breakpoint_apply_logq(self._handle, _ptr(a)), self._handle)

    @property
    def log_likelihood_table(self):
        """(6, N1, S): ``calculate_log_likelihood_total(n, s, u)`` in planes u and
        ``calculate_log_likelihood_allele(n, s, v, w)`` in planes 2 + 2 v + w (``bpmodel.pyx:751,809``), as the
        device caches them for the current h / dispersions / masks.  Not a reference attribute; diagnostic."""
        return self._get(_lib.A_LL_TABLE, (6, self.num_segments, self.num_cn_states))

    def calculate_log_transmat(self, log_transmat):
        out = np.empty(self._NSS, dtype=np.float64)
        check(lib.rmx_calculate_log_transmat(self._handle, _ptr(out)), self._handle)
        log_transmat[:] = out

    def _eval(self, what, n, s, u=0, v=0, w=0):
        out = np.zeros((8,), dtype=np.float64)
        check(lib.rmx_eval(self._handle, what, int(n), int(s), int(u), int(v), int(w), _ptr(out)), self._handle)
        return out

    def calculate_expected_total_reads(self, n, s):
        return float(self._eval(0, n, s)[0])

    def calculate_expected_total_reads_partial_h(self, n, s, partial_h):
        partial_h[:] = self._eval(1, n, s)[:self.num_clones]

    def calculate_expected_allele_ratio(self, n, s):
        return float(self._eval(2, n, s)[0])

    def calculate_expected_allele_ratio_partial_h(self, n, s, partial_h):
        partial_h[:] = self._eval(3, n, s)[:self.num_clones]

    def calculate_log_prior_cn(self, n, s):
        return float(self._eval(4, n, s)[0])

    def calculate_log_likelihood_total(self, n, s, u):
        return float(self._eval(5, n, s, u=u)[0])

    def calculate_log_likelihood_total_partial_h(self, n, s, u, partial_h):
        partial_h[:] = self._eval(6, n, s, u=u)[:self.num_clones]

    def calculate_log_likelihood_allele(self, n, s, v, w):
        return float(self._eval(7, n, s, v=v, w=w)[0])

    def calculate_log_likelihood_allele_partial_h(self, n, s, v, w, partial_h):
        partial_h[:] = self._eval(8, n, s, v=v, w=w)[:self.num_clones]

    # ------------------------------------------------------------------ instrumentation
    def launch_count(self):
        n = C.c_int64()
        check(lib.rmx_launch_count(self._handle, C.byref(n)), self._handle)
        return n.value

    def chain_timing(self, enable=True):
        """Returns (total_ms, launches) of the chain-message kernel since the last call and
        switches CUDA-event timing of that kernel on or off."""
        ms, n = C.c_double(), C.c_int64()
        check(lib.rmx_chain_timing(self._handle, 1 if enable else 0, C.byref(ms), C.byref(n)), self._handle)
        return ms.value, n.value

    def kernel_timing(self, enable=True):
        """``{kernel class: (total_ms, launches)}`` of the batch's launches since the last call (CUDA events on
        its stream); switches the timing on or off."""
        ms = np.zeros(len(_lib.KERNEL_CLASSES))
        n = np.zeros(len(_lib.KERNEL_CLASSES), dtype=np.int64)
        check(lib.rmx_kernel_timing(self._handle, 1 if enable else 0, _ptr(ms), _ptr(n)), self._handle)
        return dict((k, (float(ms[i]), int(n[i]))) for i, k in enumerate(_lib.KERNEL_CLASSES))

    def synchronize(self):
        check(lib.rmx_synchronize(self._handle), self._handle)

    def event_record(self, slot):
        check(lib.rmx_event_record(self._handle, int(slot)), self._handle)

    def event_elapsed(self, slot0, other, slot1):
        """Milliseconds between event ``slot0`` of this model and event ``slot1`` of ``other``."""
        ms = C.c_double()
        check(lib.rmx_event_elapsed(self._handle, int(slot0), other._handle, int(slot1), C.byref(ms)), self._handle)
        return ms.value


class RemixtModel(_ModelBase):
    """Structured variational inference core of ReMixT on one B200 (``bpmodel.pyx:397``): one fit."""

    def __init__(self, num_clones, num_segments, num_breakpoints, normal_contamination,
                 cn_states, brk_states, h_init, l, x, y, is_telomere, breakpoint_idx,
                 breakpoint_orient, transition_penalty, divergence_weight, device=None, shard=False):
        self._setup(num_clones, num_segments, num_breakpoints, normal_contamination, cn_states, brk_states,
                    h_init, l, x, y, is_telomere, breakpoint_idx, breakpoint_orient, transition_penalty,
                    divergence_weight, device, None, shard=shard)


class RemixtBatch(_ModelBase):
    """G fits that share data, state tables and breakpoint structure -- the fits of the reference's
    initialisation grid (``analysis/pipeline.py:42-103``) -- advanced by the same kernel launches.

    Same constructor as :class:`RemixtModel` except ``h_init`` is ``(G, M)`` and ``divergence_weight``
    ``(G,)``.  Every method acts on all G fits: array attributes have a leading dimension G, scalar
    results are ``(G,)`` arrays.  ``batch.fit(g)`` is a :class:`RemixtModel` view of fit ``g`` (the same
    device state) on which the host driver of one fit runs unchanged.
    """
    _single = False

    def __init__(self, num_clones, num_segments, num_breakpoints, normal_contamination,
                 cn_states, brk_states, h_init, l, x, y, is_telomere, breakpoint_idx,
                 breakpoint_orient, transition_penalty, divergence_weight, device=None, shard=False):
        h_init = np.ascontiguousarray(h_init, dtype=np.float64)
        if h_init.ndim != 2:
            raise ValueError('h_init of a batch must be (num_fits, num_clones)')
        self._gn = h_init.shape[0]
        self._setup(num_clones, num_segments, num_breakpoints, normal_contamination, cn_states, brk_states,
                    h_init, l, x, y, is_telomere, breakpoint_idx, breakpoint_orient, transition_penalty,
                    divergence_weight, device, self._gn, shard=shard)
        self._views = {}

    @property
    def num_fits(self):
        return self._gn

    def fit(self, g):
        """:class:`RemixtModel` view of fit ``g``."""
        if g not in self._views:
            if not 0 <= g < self._gn:
                raise IndexError(g)
            handle = C.c_void_p()
            check(lib.rmx_batch_view(self._handle, int(g), 1, C.byref(handle)), self._handle)
            view = RemixtModel.__new__(RemixtModel)
            for k in ('num_clones', 'num_segments', 'num_breakpoints', 'normal_contamination', 'num_alleles',
                      'cn_states', 'brk_states', 'cn_max', 'num_cn_states', 'num_brk_states', 'is_telomere',
                      'breakpoint_idx', 'breakpoint_orient', 'device'):
                setattr(view, k, getattr(self, k))
            view._handle = handle
            view._derived = self._derived
            view._parent = self          # keeps the batch object (not only the device batch) alive
            self._views[g] = view
        return self._views[g]

    def __len__(self):
        return self._gn

    def __getitem__(self, g):
        return self.fit(g)


def nccl_unique_id(nccl_library=None):
    """128 bytes identifying a new NCCL communicator (``rmx_comm_unique_id``); rank 0 makes it and hands it to the
    other ranks.  ``nccl_library``: path of ``libnccl.so.2`` when the process has not loaded it yet (a process that
    imported torch has)."""
    if nccl_library is not None:
        check(lib.rmx_comm_load(str(nccl_library).encode()), None)
    buf = np.zeros(128, dtype=np.uint8)
    check(lib.rmx_comm_unique_id(_ptr(buf)), None)
    return buf.tobytes()


def _scalar_property(name):
    return property(lambda self: self._get_scalar(name), lambda self, v: self._set_scalar(name, v))


for _name in ('negbin_r_0', 'negbin_r_1', 'negbin_hdel_mu', 'negbin_hdel_r_0', 'negbin_hdel_r_1',
              'betabin_M_0', 'betabin_M_1', 'betabin_loh_p', 'betabin_loh_M_0', 'betabin_loh_M_1',
              'prior_outlier_total', 'prior_outlier_allele', 'transition_penalty', 'divergence_weight',
              'hmm_log_norm_const'):
    setattr(_ModelBase, _name, _scalar_property(_name))


# ---------------------------------------------------------------------- module functions
def _f64(a, ndim, name):
    a = np.asarray(a)
    if a.dtype != np.float64 or a.ndim != ndim:
        raise ValueError('%s must be a float64 array with %d dimensions' % (name, ndim))
    return np.ascontiguousarray(a)


def sum_product(framelogprob, log_transmat, alphas, betas):
    """Sum-product messages on a chain (``bpmodel.pyx:1213-1246``); fills ``alphas`` / ``betas``."""
    F = _f64(framelogprob, 2, 'framelogprob')
    T = _f64(log_transmat, 3, 'log_transmat')
    N, S = F.shape
    if T.shape != (N - 1, S, S):
        raise ValueError('log_transmat must have shape (n_observations - 1, n_components, n_components)')
    a = np.empty((N, S))
    b = np.empty((N, S))
    check(lib.rmx_sum_product(default_device, N, S, _ptr(F), _ptr(T), _ptr(a), _ptr(b)), None)
    np.asarray(alphas)[...] = a
    np.asarray(betas)[...] = b


def sum_product_2paramtrans(framelogprob, alphas, betas, e0, e1):
    """Two-parameter-transition sum-product (``bpmodel.pyx:1249-1292``)."""
    F = _f64(framelogprob, 2, 'framelogprob')
    N, S = F.shape
    a = np.empty((N, S))
    b = np.empty((N, S))
    check(lib.rmx_sum_product_2paramtrans(default_device, N, S, _ptr(F), _ptr(a), _ptr(b), float(e0), float(e1)),
          None)
    np.asarray(alphas)[...] = a
    np.asarray(betas)[...] = b


def max_product(framelogprob, log_transmat, state_sequence):
    """Viterbi decode (``bpmodel.pyx:1296-1333``); fills ``state_sequence``, returns the log probability."""
    F = _f64(framelogprob, 2, 'framelogprob')
    T = _f64(log_transmat, 3, 'log_transmat')
    N, S = F.shape
    if T.shape != (N - 1, S, S):
        raise ValueError('log_transmat must have shape (n_observations - 1, n_components, n_components)')
    path = np.empty((N,), dtype=np.int64)
    lp = C.c_double()
    check(lib.rmx_max_product(default_device, N, S, _ptr(F), _ptr(T), _ptr(path), C.byref(lp)), None)
    np.asarray(state_sequence)[...] = path
    return lp.value
