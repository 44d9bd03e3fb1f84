"""Lock-step M-step of an initialisation grid.

The reference runs, per fit and EM iteration, ``update_h`` (L-BFGS-B on a segment sample, ``cn_model.py:482-531``)
and ``update_param`` for every dispersion parameter (``scipy.optimize.brute`` = 20-point grid + Nelder-Mead polish,
``cn_model.py:533-569``): about 290 objective evaluations, each a Python call into the model.  For a grid of G fits
that is 290 G host round trips per EM iteration against an E-step of a few milliseconds per fit.

Here the G optimisers advance in LOCK STEP: every round gathers the next evaluation point of every fit that is
still running and hands them to the device as ONE call (``RemixtBatch.mstep_evaluate``: one launch sequence, one
synchronisation).  Per fit, the sequence of evaluation points, objective values, accept / reject decisions and
the consumption of its ``RandomState`` are exactly those of the per-fit driver (``cn_model.BreakpointModel.update_h``
/ ``update_param``), including the reference's quirks:

* the acceptance test of ``update_param`` runs with the LAST value the optimiser evaluated, not with its optimum
  (``cn_model.py:561-563``); values outside the parameter's bounds are never assigned (``:541-543``);
* ``brute`` evaluates its grid in ascending order and polishes with ``fmin`` from the grid minimum.

``neldermead_1d`` is a re-statement of ``scipy.optimize.fmin`` (``_minimize_neldermead`` with default options,
one variable) as a generator that yields the next point and receives its value; ``tests/test_mstep_cpu.py`` checks
it evaluation by evaluation against scipy.  L-BFGS-B stays scipy's own code: one host thread per fit runs the
unmodified ``scipy.optimize.minimize`` and blocks in its objective until the round has been evaluated.

The evaluator interface (what a *backend* provides) is small so that the same driver runs on the CUDA batch
(:class:`BatchBackend`) and, for tests, on a list of per-fit model objects such as the compiled reference
(:class:`ModelListBackend`).
"""
import collections
import os
import threading
import time

import numpy as np
import scipy.optimize


# ---------------------------------------------------------------------------------------------------------------
# Nelder-Mead in one variable as a coroutine (scipy.optimize.fmin defaults: xtol = ftol = 1e-4, maxiter = maxfun = 200)
# ---------------------------------------------------------------------------------------------------------------
def neldermead_1d(x0, xatol=1e-4, fatol=1e-4, maxiter=200, maxfun=200):
    """Generator form of ``scipy.optimize.fmin(func, [x0])`` for one variable: yields the next evaluation point
    (float), expects its objective value through ``send``; returns ``(x, fval, iterations, funcalls, warnflag)``
    as the ``StopIteration`` value.  Follows ``scipy/optimize/_optimize.py:_minimize_neldermead`` statement by
    statement (rho = 1, chi = 2, psi = sigma = 0.5)."""
    calls = 0
    s0 = float(x0)
    s1 = 1.05 * s0 if s0 != 0. else 0.00025
    f0 = f1 = float('inf')
    exhausted = False
    # initial simplex values
    for k in range(2):
        if calls >= maxfun:
            exhausted = True
            break
        calls += 1
        v = yield (s0 if k == 0 else s1)
        if k == 0:
            f0 = v
        else:
            f1 = v
    if _second_first(f0, f1):
        s0, s1, f0, f1 = s1, s0, f1, f0
    iterations = 1
    converged = False
    while calls < maxfun and iterations < maxiter:
        aborted = False
        while True:     # one iteration; `break` leaves it (the except / fall-through of the try block)
            if abs(s1 - s0) <= xatol and abs(f0 - f1) <= fatol:
                converged = True
                break
            xbar = s0
            xr = 2 * xbar - 1 * s1
            if calls >= maxfun:
                aborted = True
                break
            calls += 1
            fxr = yield xr
            doshrink = False
            if fxr < f0:
                xe = 3 * xbar - 2 * s1
                if calls >= maxfun:
                    aborted = True
                    break
                calls += 1
                fxe = yield xe
                if fxe < fxr:
                    s1, f1 = xe, fxe
                else:
                    s1, f1 = xr, fxr
            else:
                # with one variable fsim[-2] is fsim[0], so `fxr < fsim[-2]` cannot hold here: contraction
                if fxr < f0:
                    s1, f1 = xr, fxr
                elif fxr < f1:
                    xc = 1.5 * xbar - 0.5 * s1
                    if calls >= maxfun:
                        aborted = True
                        break
                    calls += 1
                    fxc = yield xc
                    if fxc <= fxr:
                        s1, f1 = xc, fxc
                    else:
                        doshrink = True
                else:
                    xcc = 0.5 * xbar + 0.5 * s1
                    if calls >= maxfun:
                        aborted = True
                        break
                    calls += 1
                    fxcc = yield xcc
                    if fxcc < f1:
                        s1, f1 = xcc, fxcc
                    else:
                        doshrink = True
                if doshrink:
                    s1 = s0 + 0.5 * (s1 - s0)
                    if calls >= maxfun:
                        aborted = True
                        break
                    calls += 1
                    f1 = yield s1
            break
        if converged:
            break
        if not aborted:
            iterations += 1
        if _second_first(f0, f1):
            s0, s1, f0, f1 = s1, s0, f1, f0
    warnflag = 1 if calls >= maxfun else (2 if iterations >= maxiter else 0)
    return s0, _nanmin(f0, f1), iterations, calls, warnflag


def _second_first(f0, f1):
    # np.argsort([f0, f1]) == [1, 0]: NaN sorts last, ties keep their order
    if f1 != f1:
        return False
    if f0 != f0:
        return True
    return f1 < f0


def _nanmin(a, b):
    # np.min propagates NaN
    if a != a or b != b:
        return float('nan')
    return min(a, b)


# ---------------------------------------------------------------------------------------------------------------
# L-BFGS-B as a coroutine over scipy's own compiled routine
# ---------------------------------------------------------------------------------------------------------------
def lbfgsb_coroutine(x0, bounds, maxcor=10, ftol=2.220446049250313e-09, gtol=1e-5, maxfun=15000, maxiter=15000,
                     maxls=20):
    """``scipy.optimize.minimize(fun, x0, method='L-BFGS-B', jac=jac, bounds=bounds)`` as a generator: yields the
    next point, expects ``(f, g)`` through ``send``, returns an ``OptimizeResult`` as the ``StopIteration`` value.

    The algorithm is scipy's own compiled routine (``scipy.optimize._lbfgsb.setulb``, reverse communication), driven
    by the loop of ``scipy/optimize/_lbfgsb_py.py:_minimize_lbfgsb`` including what its ``ScalarFunction`` wrapper
    does around the user's callables (clipping ``x0`` into the bounds, evaluating the start point once, not
    re-evaluating a point it already holds).  ``setulb`` is a private interface: :func:`lbfgsb_coroutine_ok` checks
    this driver against ``scipy.optimize.minimize`` itself, and :class:`GridMStep` falls back to one blocking
    scipy call per fit on host threads when the check fails."""
    from scipy.optimize import OptimizeResult, _lbfgsb
    from scipy.optimize import _lbfgsb_py as lb
    m = maxcor
    factr = ftol / np.finfo(float).eps
    x0 = np.asarray(x0).ravel()
    n, = x0.shape
    lo = np.array([b[0] for b in bounds], dtype=float)
    hi = np.array([b[1] for b in bounds], dtype=float)
    x0 = np.clip(x0, lo, hi)
    int_dtype = np.int64 if getattr(lb, 'HAS_ILP64', False) else np.int32
    nbd = np.full(n, 2, dtype=int_dtype)        # both bounds finite (the caller's bounds are)
    assert np.all(np.isfinite(lo)) and np.all(np.isfinite(hi))
    low_bnd, upper_bnd = lo.copy(), hi.copy()
    x = np.array(x0, dtype=np.float64)
    f = np.array(0.0, dtype=np.float64)
    g = np.zeros((n,), dtype=np.float64)
    wa = np.zeros(2 * m * n + 5 * n + 11 * m * m + 8 * m, np.float64)
    iwa = np.zeros(3 * n, dtype=int_dtype)
    task = np.zeros(2, dtype=int_dtype)
    ln_task = np.zeros(2, dtype=int_dtype)
    lsave = np.zeros(4, dtype=int_dtype)
    isave = np.zeros(44, dtype=int_dtype)
    dsave = np.zeros(29, dtype=np.float64)
    # ScalarFunction.__init__ evaluates fun and grad at x0
    held_x = np.array(x0, dtype=float)
    held_f, held_g = yield held_x.copy()
    held_g = np.atleast_1d(held_g)
    nfev = 1
    n_iterations = 0
    held_key = held_x.tobytes()
    setulb = _lbfgsb.setulb
    while True:
        if g.dtype != np.float64:
            g = g.astype(np.float64)
        setulb(m, x, low_bnd, upper_bnd, nbd, f, g, factr, gtol, wa, iwa, task, lsave, isave, dsave, maxls, ln_task)
        if task[0] == 3:
            key = x.tobytes()
            if key != held_key and not np.array_equal(x, held_x):   # bytes differ and values differ (-0. == 0.)
                held_x = x.copy()
                held_key = key
                held_f, held_g = yield held_x.copy()
                if getattr(held_g, 'ndim', 0) != 1:
                    held_g = np.atleast_1d(held_g)
                nfev += 1
            f, g = held_f, held_g
        elif task[0] == 1:
            n_iterations += 1
            if n_iterations >= maxiter:
                task[0] = 5
                task[1] = 504
            elif nfev > maxfun:
                task[0] = 5
                task[1] = 502
        else:
            break
    if task[0] == 4:
        warnflag = 0
    elif nfev > maxfun or n_iterations >= maxiter:
        warnflag = 1
    else:
        warnflag = 2
    msg = lb.status_messages[task[0]] + ": " + lb.task_messages[task[1]]
    return OptimizeResult(fun=f, jac=g, nfev=nfev, njev=nfev, nit=n_iterations, status=warnflag, message=msg, x=x,
                          success=(warnflag == 0))


_LBFGSB_OK = None


def lbfgsb_coroutine_ok():
    """True when :func:`lbfgsb_coroutine` reproduces ``scipy.optimize.minimize(method='L-BFGS-B')`` on this scipy
    -- every evaluation point, the result and the termination status, bit for bit -- on two bounded test problems
    (one ends on a bound, one needs many line searches).  Evaluated once per process."""
    global _LBFGSB_OK
    if _LBFGSB_OK is not None:
        return _LBFGSB_OK
    try:
        ok = True
        problems = [
            (lambda x: (float(np.sum((x - np.array([0.3, -2., 5.])) ** 2) + np.prod(np.cosh(x / 3.))),
                        2. * (x - np.array([0.3, -2., 5.])) + np.prod(np.cosh(x / 3.)) * np.tanh(x / 3.) / 3.),
             np.array([1., 1., 1.]), [(1e-8, 10.)] * 3),
            (lambda x: (float(100. * (x[1] - x[0] ** 2) ** 2 + (1. - x[0]) ** 2 + (x[2] - 0.5) ** 4),
                        np.array([-400. * x[0] * (x[1] - x[0] ** 2) - 2. * (1. - x[0]), 200. * (x[1] - x[0] ** 2),
                                  4. * (x[2] - 0.5) ** 3])),
             np.array([0.1, 3., 9.]), [(1e-8, 10.)] * 3),
        ]
        for fg, x0, bounds in problems:
            trace = []

            def fun(x):
                trace.append(np.array(x))
                return fg(x)[0]

            ref = scipy.optimize.minimize(fun, x0, method='L-BFGS-B', jac=lambda x: fg(x)[1], bounds=bounds)
            gen = lbfgsb_coroutine(x0, bounds)
            mine = []
            try:
                x = next(gen)
                while True:
                    mine.append(np.array(x))
                    x = gen.send(fg(x))
            except StopIteration as stop:
                res = stop.value
            ok = ok and len(mine) == len(trace) and all(np.array_equal(a, b) for a, b in zip(mine, trace))
            ok = ok and np.array_equal(res.x, ref.x) and res.fun == ref.fun and res.success == ref.success
            ok = ok and res.message == ref.message and res.nit == ref.nit and res.nfev == ref.nfev
        _LBFGSB_OK = bool(ok)
    except Exception:
        _LBFGSB_OK = False
    return _LBFGSB_OK


# ---------------------------------------------------------------------------------------------------------------
# backends
# ---------------------------------------------------------------------------------------------------------------
class BatchBackend(object):
    """The CUDA batch (``bpmodel.RemixtBatch``) as the evaluator of the lock-step M-step."""

    def __init__(self, batch):
        self.batch = batch
        self.G = batch.num_fits
        self.num_segments = batch.num_segments
        self.num_clones = batch.num_clones
        self._cache = {}
        self._param_names = ['negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1']
        if not getattr(batch, 'normal_contamination', True):
            self._param_names += ['negbin_hdel_mu', 'negbin_hdel_r_0', 'negbin_hdel_r_1', 'betabin_loh_p',
                                  'betabin_loh_M_0', 'betabin_loh_M_1']

    def new_iteration(self):
        self._cache.clear()

    def _arr(self, name):
        # the q-factors do not change during an M-step: one read per EM iteration serves every parameter
        if name not in self._cache:
            self._cache[name] = getattr(self.batch, name)
        return self._cache[name]

    def prefetch(self, plan):
        if any(p.startswith('negbin') for p in plan):
            self._arr('p_outlier_total')
        if any(p.startswith('betabin') for p in plan):
            self._arr('p_outlier_allele')

    def state_keys(self):
        """Per fit, the bytes of (h, every likelihood parameter): what the full-data E[LL] depends on."""
        cols = [np.array(self.batch.h)] + [np.array(getattr(self.batch, n)).reshape(self.G, 1) for n in self._param_names]
        block = np.ascontiguousarray(np.concatenate(cols, axis=1))
        return [block[g].tobytes() for g in range(self.G)]

    def get_h(self):
        return np.array(self.batch.h)

    def set_h(self, h, active=None):
        if active is None or all(active):
            self.batch.h = h
        else:
            cur = np.array(self.batch.h)
            cur[np.asarray(active, dtype=bool)] = np.asarray(h)[np.asarray(active, dtype=bool)]
            self.batch.h = cur

    def get_param(self, name):
        return np.array(getattr(self.batch, name))

    def set_param(self, name, values, active=None):
        self.batch.set_scalar_array(name, values, active)

    def outlier_weights(self, name, g):
        which = 'p_outlier_total' if name.startswith('negbin') else 'p_outlier_allele'
        return self._arr(which)[g][:, int(name[-1])]

    def weights_block(self, name, fits, out):
        """``get_param_sample_weight(name)`` (``cn_model.py:323-352``) of the fits ``fits`` for an outlier-weighted
        parameter, written to ``out (len(fits), N)``: the outlier posterior column over its own sum, exactly the
        per-fit ``weights / weights.sum()`` (``tests/test_mstep_cpu.py`` checks the row sums of the strided block
        against the per-fit sums).  Returns the flags (0 where the sum is not positive: uniform sample)."""
        which = 'p_outlier_total' if name.startswith('negbin') else 'p_outlier_allele'
        col = self._arr(which)[:, :, int(name[-1])]
        if len(fits) != col.shape[0]:
            col = col[np.asarray(fits)]
        norm = col.sum(axis=1)
        ok = norm > 0.
        np.divide(col, np.where(ok, norm, 1.)[:, np.newaxis], out=out)
        return ok.astype(np.uint8)

    def state_weights(self, which, g):
        return (self.batch.fit(g).posterior_marginals * (np.asarray(getattr(self.batch, which)) == 1)).sum(axis=-1)

    def stage(self, rows=None, counts=None):
        if rows is None:
            self.batch.mstep_stage_sample(None)
        else:
            self.batch.mstep_stage_rows(rows, counts)

    def evaluate(self, active=None, h=None, name=None, values=None, grad=False):
        return self.batch.mstep_evaluate(active=active, h=h, name=name, values=values, grad=grad)

    def param_search(self, name, lo, hi, grid, active, trace=0):
        return self.batch.mstep_param_search(name, lo, hi, grid, active=active, trace=trace)


class ModelListBackend(object):
    """The same interface over G independent per-fit model objects (the compiled reference's ``RemixtModel``, the
    oracle port, or single-fit CUDA models); evaluations are a loop.  Tests only."""

    def __init__(self, models):
        self.models = list(models)
        self.G = len(self.models)
        self.num_segments = self.models[0].num_segments
        self.num_clones = self.models[0].num_clones
        self.samples = None
        self.log = [[] for _ in self.models]        # per fit: (kind, point, value) of every evaluation, in order

    def new_iteration(self):
        pass

    def prefetch(self, plan):
        pass

    def get_h(self):
        return np.array([np.asarray(m.h) for m in self.models])

    def set_h(self, h, active=None):
        for g, m in enumerate(self.models):
            if active is None or active[g]:
                m.h = np.array(h[g], dtype=float)

    def get_param(self, name):
        return np.array([getattr(m, name) for m in self.models])

    def set_param(self, name, values, active=None):
        for g, m in enumerate(self.models):
            if active is None or active[g]:
                setattr(m, name, float(values[g]))

    def outlier_weights(self, name, g):
        which = 'p_outlier_total' if name.startswith('negbin') else 'p_outlier_allele'
        return np.asarray(getattr(self.models[g], which))[:, int(name[-1])]

    def state_weights(self, which, g):
        m = self.models[g]
        return (np.asarray(m.posterior_marginals) * (np.asarray(getattr(m, which)) == 1)).sum(axis=-1)

    def stage(self, rows=None, counts=None):
        if rows is None:
            self.samples = None
        else:
            self.samples = np.zeros((self.G, self.num_segments), dtype=np.int64)
            for g in range(self.G):
                self.samples[g, rows[g, :counts[g]]] = 1

    def evaluate(self, active=None, h=None, name=None, values=None, grad=False):
        n = 1 if name is None else np.asarray(values).reshape(self.G, -1).shape[1]
        ell = np.zeros((self.G, n))
        gr = np.zeros((self.G, self.num_clones))
        for g, m in enumerate(self.models):
            if active is not None and not active[g]:
                continue
            sample = (np.ones(self.num_segments, dtype=int) if self.samples is None
                      else np.ascontiguousarray(self.samples[g], dtype=int))
            if h is not None:
                m.h = np.array(h[g], dtype=float)
            for v in range(n):
                if name is not None:
                    setattr(m, name, float(np.asarray(values).reshape(self.G, -1)[g, v]))
                ell[g, v] = m.calculate_expected_log_likelihood(sample)
                self.log[g].append((name or ('h' if h is not None else 'all'),
                                    None if name is None else float(np.asarray(values).reshape(self.G, -1)[g, v]),
                                    ell[g, v]))
            if grad:
                out = np.zeros(self.num_clones)
                m.calculate_expected_log_likelihood_partial_h(sample, out)
                gr[g] = out
        return (ell, gr) if grad else ell


# ---------------------------------------------------------------------------------------------------------------
# lock-step rendezvous of G blocking optimiser threads
# ---------------------------------------------------------------------------------------------------------------
class _Rendezvous(object):
    """Worker thread g calls ``request(g, x)`` and blocks; once every running worker has a request pending the
    coordinator evaluates them all at once and releases the workers with their results."""

    def __init__(self, workers):
        self.lock = threading.Lock()
        self.wake = dict((g, threading.Semaphore(0)) for g in workers)
        self.coord = threading.Semaphore(0)
        self.pending = {}
        self.results = {}
        self.running = len(workers)

    def request(self, g, x):
        with self.lock:
            self.pending[g] = x
            ready = len(self.pending) >= self.running
        if ready:
            self.coord.release()
        self.wake[g].acquire()
        r = self.results.pop(g)
        if isinstance(r, BaseException):
            raise r
        return r

    def finish(self, g):
        with self.lock:
            self.running -= 1
            ready = self.running == 0 or len(self.pending) >= self.running
        if ready:
            self.coord.release()

    def serve(self, evaluate):
        while True:
            self.coord.acquire()
            with self.lock:
                if self.running == 0 and not self.pending:
                    return
                if len(self.pending) < self.running:
                    continue
                batch, self.pending = self.pending, {}
            try:
                res = evaluate(batch)
            except BaseException as e:      # surfaces in every waiting optimiser
                res = dict((g, e) for g in batch)
            for g in batch:
                self.results[g] = res[g]
                self.wake[g].release()


_CHOICE_ERRORS = {1: 'probabilities do not sum to 1', 2: 'probabilities are not non-negative',
                  3: 'probabilities contain NaN', 4: 'Fewer non-zero entries in p than size',
                  5: "Cannot take a larger sample than population when 'replace=False'"}


def native_choice(random_states, n, size, weights, num_threads=None, buffer=None, block=None):
    """``rs.choice(n, size=size, replace=False, p=weights[d][g])`` for every stream ``random_states[g]`` (numpy
    ``RandomState`` objects over MT19937) and draw ``d`` in order, by the library's host threads
    (``rmx_host_sample_segments``: numpy's legacy algorithm restated, bit for bit).  ``weights[d][g]`` is an ``(n,)``
    array or ``None`` (uniform); ``block = (w (D, G, n), flags (D, G))`` passes them in place instead.  The streams are advanced in place.  Returns ``chosen (D, G, size)`` and
    ``status (D, G)`` (see ``_CHOICE_ERRORS``; -1 = skipped after an earlier failure of the fit)."""
    from . import _lib
    G, D = len(random_states), len(weights)
    states = [rs.get_state(legacy=True) for rs in random_states]
    if any(st[0] != 'MT19937' for st in states):
        raise TypeError('native_choice needs MT19937 streams')
    key = np.ascontiguousarray(np.array([st[1] for st in states], dtype=np.uint32).reshape(G, 624))
    pos = np.ascontiguousarray(np.array([st[2] for st in states], dtype=np.int32))
    flags = np.zeros((D, G), dtype=np.uint8)
    w = None
    if block is not None:       # (D, G, n) weights already in place, with their (D, G) flags
        w, flags = block
    elif any(weights[d][g] is not None for d in range(D) for g in range(G)):
        if buffer is not None and buffer.get('shape') == (D, G, n):
            w = buffer['w']
        else:
            w = np.empty((D, G, n), dtype=np.float64)
            if buffer is not None:
                buffer['shape'], buffer['w'] = (D, G, n), w
        for d in range(D):
            for g in range(G):
                if weights[d][g] is not None:
                    w[d, g] = weights[d][g]
                    flags[d, g] = 1
    chosen = np.zeros((D, G, size), dtype=np.int64)
    status = np.zeros((D, G), dtype=np.int32)
    ptr = lambda a: a.ctypes.data
    _lib.check(_lib.lib.rmx_host_sample_segments(G, D, int(n), int(size), ptr(key), ptr(pos), None if w is None else ptr(w),
                                                 ptr(flags), ptr(chosen), ptr(status),
                                                 int(num_threads or _host_threads())), None)
    for g, (rs, st) in enumerate(zip(random_states, states)):
        rs.set_state((st[0], key[g], int(pos[g])) + tuple(st[3:]))
    return chosen, status


def _host_threads():
    try:
        return min(16, len(os.sched_getaffinity(0)))
    except AttributeError:
        return min(16, os.cpu_count() or 1)


# ---------------------------------------------------------------------------------------------------------------
# the driver
# ---------------------------------------------------------------------------------------------------------------
class GridMStep(object):
    """M-steps of the G fits of a batch in lock step.  ``drivers`` are the per-fit ``cn_model.BreakpointModel``
    objects (they supply the random streams, ``likelihood_params``, bounds, ``do_h_update`` and collect the
    ``decisions``); ``errors`` maps a fit index to the exception that took that fit out of the grid."""

    def __init__(self, backend, drivers, errors=None, lbfgsb='auto'):
        self.lbfgsb = lbfgsb    # 'auto' | 'coroutine' | 'threads'
        self.native_sampling = True
        self.overlap_sampling = True
        self.native_search = True   # update_param's brute + fmin inside the library (BatchBackend.param_search)
        self._full_cache = {}
        self._wbuf = {}         # reusable weight block of the native sampler (fresh pages cost more than the draws)
        self.be = backend
        self.drivers = drivers
        self.G = backend.G
        self.errors = errors if errors is not None else {}
        self.rounds = 0         # lock-step evaluation rounds (device calls) so far
        self.evaluations = 0    # objective evaluations they carried
        self.seconds = collections.defaultdict(float)      # wall clock by phase

    def _alive(self):
        return np.array([g not in self.errors for g in range(self.G)], dtype=bool)

    def _choose(self, g, weights=None):
        # cn_model.py:475-480 (_create_sample): the segments of the sample
        d = self.drivers[g]
        N = self.be.num_segments
        size = int(min(200, N / 10))
        rs = d.random_state if d.random_state is not None else np.random
        return rs.choice(N, size=size, replace=False, p=weights)

    def _draw_samples(self, plan, alive=None, timed=True):
        """The segment samples of one EM iteration, ``{step: (rows (G, size), counts (G))}``.  Every sample is a function of the
        E-step's q-factors and of the fit's random stream alone (nothing the M-step changes), so they are all drawn
        up front -- per fit in the reference's order h, then each parameter, which is what keeps every stream
        identical to the per-fit driver's -- with the fits spread over host threads (numpy releases the GIL inside
        the cumulative sums and searches that dominate ``RandomState.choice``).  A fit whose draw raises takes the
        error at the step that would have raised it."""
        t0 = time.perf_counter()
        be = self.be
        alive = np.flatnonzero(self._alive() if alive is None else alive)
        shared_stream = any(self.drivers[g].random_state is None for g in alive)

        def draw(g):
            out = []
            for step in plan:
                try:
                    out.append(self._choose(g, None if step == 'h' else self._weights(step, g)))
                except Exception as e:      # e.g. fewer non-zero weights than the sample size
                    out.append(e)
                    break
            return out

        native = (self.native_sampling and not shared_stream and len(alive) > 0 and
                  all(type(self.drivers[g].random_state) is np.random.RandomState for g in alive))
        if native:
            # weights on the calling thread (they read the backend), the draws on the library's host threads
            size = int(min(200, be.num_segments / 10))
            D, A, N = len(plan), len(alive), be.num_segments
            weighted = any(step != 'h' for step in plan)
            if weighted and (D, A, N) not in self._wbuf:
                self._wbuf[(D, A, N)] = np.empty((D, A, N), dtype=np.float64)
            w, flags = self._wbuf.get((D, A, N)) if weighted else None, np.zeros((D, A), dtype=np.uint8)
            def fill(d):
                step = plan[d]
                if step == 'h':
                    return
                if hasattr(be, 'weights_block') and step in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1'):
                    flags[d] = be.weights_block(step, alive, w[d])
                else:
                    for k, g in enumerate(alive):
                        wg = self._weights(step, g)
                        if wg is not None:
                            w[d, k], flags[d, k] = wg, 1

            # one strided sum and one division over (fits, N) per parameter: numpy releases the GIL inside both
            if hasattr(be, 'weights_block') and D > 2:
                import concurrent.futures
                with concurrent.futures.ThreadPoolExecutor(min(D, _host_threads())) as pool:
                    list(pool.map(fill, range(D)))
            else:
                for d in range(D):
                    fill(d)
            chosen, status = native_choice([self.drivers[g].random_state for g in alive], N, size,
                                           [None] * D if weighted else [[None] * A] * D,
                                           block=(w, flags) if weighted else None)
            drawn = []
            for k in range(len(alive)):
                out = []
                for d in range(len(plan)):
                    if status[d, k] == 0:
                        out.append(chosen[d, k])
                    else:
                        if status[d, k] > 0:
                            out.append(ValueError(_CHOICE_ERRORS[int(status[d, k])]))
                        break
                drawn.append(out)
        else:
            drawn = [draw(g) for g in alive]
        size = int(min(200, be.num_segments / 10))
        samples = dict((step, (np.zeros((self.G, size), dtype=np.int32), np.zeros(self.G, dtype=np.int32)))
                       for step in plan)
        failed = {}
        for g, out in zip(alive, drawn):
            for step, chosen in zip(plan, out):
                if isinstance(chosen, Exception):
                    failed[(step, int(g))] = chosen
                else:
                    # ascending, as the rows of the reference's 0/1 sample mask are visited (bpmodel.pyx:1137)
                    samples[step][0][g] = np.sort(chosen)
                    samples[step][1][g] = size
        if timed:
            self.seconds['sample'] += time.perf_counter() - t0
        return samples, failed

    def _evaluate(self, **kw):
        self.rounds += 1
        act = kw.get('active')
        self.evaluations += int(np.sum(act)) if act is not None else self.G
        return self.be.evaluate(**kw)

    def _full_ell(self, alive):
        """E[LL] over every segment for the fits ``alive``.  The value is a function of the fit's (h, dispersions)
        alone while the E-step's factors stand, and the device evaluation is deterministic, so a fit whose parameters
        are back at a state already evaluated in this M-step (a rejected step restores them) takes the stored value:
        bit for bit what a fresh evaluation would return."""
        t0 = time.perf_counter()
        keys = self.be.state_keys() if hasattr(self.be, 'state_keys') else None
        out = np.zeros(self.G)
        need = np.array(alive, dtype=np.uint8)
        if keys is not None:
            for g in np.flatnonzero(alive):
                v = self._full_cache.get((int(g), keys[g]))
                if v is not None:
                    out[g], need[g] = v, 0
        if need.any():
            self.be.stage(None)
            got = self._evaluate(active=need)[:, 0]
            for g in np.flatnonzero(need):
                out[g] = got[g]
                if keys is not None:
                    self._full_cache[(int(g), keys[g])] = got[g]
        self.seconds['full_data_ell'] += time.perf_counter() - t0
        return out

    # ------------------------------------------------------------------------------------------ update_h
    def _take_sample(self, step, samples):
        """The pre-drawn sample of ``step`` as ``(rows, counts)`` (drawn here when the step runs on its own)."""
        if samples is None:
            samples, failed = self._draw_samples([step])
        else:
            samples, failed = samples
        for (st, g), e in failed.items():
            if st == step:
                self.errors[g] = e
        return samples[step]

    def update_h(self, samples=None):
        """``update_h`` (``cn_model.py:482-531``) of every fit."""
        be, M = self.be, self.be.num_clones
        alive = self._alive()
        if not alive.any():
            return
        h_before = be.get_h()
        ell_before = self._full_ell(alive)
        rows, counts = self._take_sample('h', samples)
        alive = self._alive()
        if not alive.any():
            return
        t1 = time.perf_counter()
        be.stage(rows, counts)
        t2 = time.perf_counter()
        self.seconds['stage'] += t2 - t1

        workers = [int(g) for g in np.flatnonzero(alive)]
        results = {}

        def evaluate(batch):
            active = np.zeros(self.G, dtype=np.uint8)
            hv = np.array(h_before)
            for g, x in batch.items():
                active[g] = 1
                hv[g] = x
            ell, grad = self._evaluate(active=active, h=hv, grad=True)
            return dict((g, (float(ell[g, 0]), grad[g].copy())) for g in batch)

        bounds = [(1e-8, 10.)] * M
        if self.lbfgsb == 'coroutine' or (self.lbfgsb == 'auto' and lbfgsb_coroutine_ok()):
            gens, pending = {}, {}
            for g in workers:
                gens[g] = lbfgsb_coroutine(h_before[g], bounds)
                pending[g] = next(gens[g])
            while pending:
                answers = evaluate(pending)
                nxt = {}
                for g, (ell, grad) in answers.items():
                    try:
                        nxt[g] = gens[g].send((-ell, -grad))
                    except StopIteration as stop:
                        results[g] = stop.value
                    except BaseException as e:
                        results[g] = e
                pending = nxt
        else:
            self._update_h_threads(workers, evaluate, h_before, bounds, results)
        self.seconds['lbfgsb'] += time.perf_counter() - t2

        h_new = np.array(h_before)
        for g in workers:
            res = results[g]
            if isinstance(res, BaseException):
                self.errors[g] = res
                continue
            if not res.success:
                # cn_model.py:510-520: with this scipy the message never equals the bare task name, so a failed
                # line search is an error of the fit, exactly as in the reference
                self.errors[g] = ValueError('optimization failed\n{}'.format(res))
                continue
            h_new[g] = res.x
        alive = self._alive()
        if not alive.any():
            return
        be.set_h(h_new, alive)
        ell_after = self._full_ell(alive)
        final = np.array(h_new)
        for g in np.flatnonzero(alive):
            rejected = ell_after[g] < ell_before[g]
            self.drivers[g].decisions.append(('h', not rejected, float(ell_before[g]), float(ell_after[g])))
            if rejected:
                final[g] = h_before[g]
        be.set_h(final, alive)

    def _update_h_threads(self, workers, evaluate, h_before, bounds, results):
        """L-BFGS-B of every fit as the unmodified ``scipy.optimize.minimize`` on its own host thread, the objective
        blocking until the round has been evaluated (used when the coroutine driver is not available)."""
        rv = _Rendezvous(workers)

        def optimise(g):
            cache = {}

            def both(h):
                key = h.tobytes()
                if key not in cache:
                    cache.clear()
                    cache[key] = rv.request(g, np.array(h, dtype=float))
                return cache[key]

            try:
                results[g] = scipy.optimize.minimize(lambda h: -both(h)[0], h_before[g], method='L-BFGS-B',
                                                     jac=lambda h: -both(h)[1], bounds=bounds)
            except BaseException as e:
                results[g] = e
            finally:
                rv.finish(g)

        threads = [threading.Thread(target=optimise, args=(g,)) for g in workers]
        for t in threads:
            t.start()
        rv.serve(evaluate)
        for t in threads:
            t.join()

    # ------------------------------------------------------------------------------------------ update_param
    def _weights(self, name, g):
        be = self.be
        if name in ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1'):
            w = be.outlier_weights(name, g)
        elif name == 'negbin_hdel_mu':
            w = be.state_weights('is_hdel', g)
        elif name in ('negbin_hdel_r_0', 'negbin_hdel_r_1'):
            w = be.state_weights('is_hdel', g) * be.outlier_weights('negbin_r_' + name[-1], g)
        elif name == 'betabin_loh_p':
            w = be.state_weights('is_loh', g)
        elif name in ('betabin_loh_M_0', 'betabin_loh_M_1'):
            w = be.state_weights('is_loh', g) * be.outlier_weights('betabin_M_' + name[-1], g)
        else:
            raise KeyError(name)
        norm = w.sum()
        return w / norm if norm > 0. else None

    def update_param(self, name, samples=None):
        """``update_param(name)`` (``cn_model.py:533-569``) of every fit."""
        be = self.be
        alive = self._alive()
        if not alive.any():
            return
        lo, hi = self.drivers[0].likelihood_param_bounds[name]
        value_before = be.get_param(name)
        ell_before = self._full_ell(alive)
        rows, counts = self._take_sample(name, samples)
        alive = self._alive()
        if not alive.any():
            return
        t2 = time.perf_counter()
        be.stage(rows, counts)
        t3 = time.perf_counter()
        self.seconds['stage'] += t3 - t2

        grid = np.mgrid[slice(lo, hi, complex(20))]
        if self.native_search and hasattr(be, 'param_search'):
            # the same sequence of evaluations, driven by the library's host loop (no interpreter between rounds)
            result, (rounds, evaluations) = be.param_search(name, lo, hi, grid, alive.astype(np.uint8))
            self.rounds += rounds
            self.evaluations += evaluations
            self.seconds['param_search'] += time.perf_counter() - t3
            self._accept_param(name, alive, value_before, ell_before, result)
            return
        # scipy.optimize.brute: the grid, ascending, one call for all fits (values outside the bounds are inf and
        # are not assigned -- cn_model.py:541-543)
        inside = [float(v) for v in grid if not (v < lo or v > hi)]
        J = np.full((self.G, grid.shape[0]), np.inf)
        if inside:
            ell = self._evaluate(active=alive.astype(np.uint8), name=name,
                                 values=np.broadcast_to(np.array(inside), (self.G, len(inside))))
            self.rounds += len(inside) - 1
            self.evaluations += (len(inside) - 1) * int(alive.sum())
            J[:, [i for i, v in enumerate(grid) if not (v < lo or v > hi)]] = -ell
        t4 = time.perf_counter()
        self.seconds['brute_grid'] += t4 - t3
        # fmin from the grid minimum, every fit's simplex advancing one evaluation per round
        gens, pending, result = {}, {}, {}
        for g in np.flatnonzero(alive):
            gens[g] = neldermead_1d(float(grid[int(np.argmin(J[g]))]))
            pending[g] = next(gens[g])
        vals = np.zeros((self.G, 1))
        while pending:
            active = np.zeros(self.G, dtype=np.uint8)
            answers = {}
            for g, x in pending.items():
                if x < lo or x > hi:
                    answers[g] = np.inf
                else:
                    active[g] = 1
                    vals[g, 0] = x
            if active.any():
                ell = self._evaluate(active=active, name=name, values=vals)
                for g in np.flatnonzero(active):
                    answers[g] = -float(ell[g, 0])
            nxt = {}
            for g, f in answers.items():
                try:
                    nxt[g] = gens[g].send(f)
                except StopIteration as stop:
                    result[g] = stop.value[0]
            pending = nxt
        self.seconds['nelder_mead'] += time.perf_counter() - t4

        self._accept_param(name, alive, value_before, ell_before, result)

    def _accept_param(self, name, alive, value_before, ell_before, result):
        """The acceptance test of ``update_param`` (``cn_model.py:561-569``): full-data E[LL] with the parameter as the
        optimiser left it (its last evaluated value) against the value before."""
        ell_after = self._full_ell(alive)
        final = np.array(value_before)
        for g in np.flatnonzero(alive):
            rejected = ell_after[g] < ell_before[g]
            self.drivers[g].decisions.append((name, not rejected, float(ell_before[g]), float(ell_after[g])))
            final[g] = value_before[g] if rejected else result[g]
        self.be.set_param(name, final, alive)

    # ------------------------------------------------------------------------------------------ one EM iteration
    def run(self):
        """``em_update_h`` + ``em_update_params`` (``cn_model.py:411-416``) for the whole grid."""
        self.be.new_iteration()
        self._full_cache = {}
        d0 = self.drivers[0]
        plan = (['h'] if d0.do_h_update else []) + list(d0.likelihood_params)
        t0 = time.perf_counter()
        self.be.prefetch(plan)
        self.seconds['weights'] += time.perf_counter() - t0
        # the parameters' samples depend on the E-step's q-factors (read above) and on the fits' random streams only:
        # they are drawn on a host thread while update_h runs, after h's own draw (each stream keeps its order)
        standard = ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')
        if (self.overlap_sampling and self.native_sampling and d0.do_h_update and len(plan) > 1 and
                hasattr(self.be, 'weights_block') and all(p in standard for p in plan[1:])):
            import concurrent.futures
            samples = self._draw_samples(['h'])
            alive = self._alive()
            with concurrent.futures.ThreadPoolExecutor(1) as pool:
                rest = pool.submit(self._draw_samples, plan[1:], alive, False)
                self.update_h(samples)
                t1 = time.perf_counter()
                samples = rest.result()
                self.seconds['sample'] += time.perf_counter() - t1
        else:
            samples = self._draw_samples(plan)
            if d0.do_h_update:
                self.update_h(samples)
        for name in d0.likelihood_params:
            self.update_param(name, samples)
