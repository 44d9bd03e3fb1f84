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
import threading

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

    def new_iteration(self):
        self._cache.clear()

    def _arr(self, name):
        # the q-factors do not change during an M-step: one read per EM iteration serves every parameter
        if name not in self._cache:
            self._cache[name] = getattr(self.batch, name)
        return self._cache[name]

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

    def state_weights(self, which, g):
        return (self.batch.fit(g).posterior_marginals * (np.asarray(getattr(self.batch, which)) == 1)).sum(axis=-1)

    def stage(self, samples):
        self.batch.mstep_stage_sample(samples)

    def evaluate(self, active=None, h=None, name=None, values=None, grad=False):
        return self.batch.mstep_evaluate(active=active, h=h, name=name, values=values, grad=grad)


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

    def stage(self, samples):
        self.samples = samples

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


# ---------------------------------------------------------------------------------------------------------------
# the driver
# ---------------------------------------------------------------------------------------------------------------
class GridMStep(object):
    """M-steps of the G fits of a batch in lock step.  ``drivers`` are the per-fit ``cn_model.BreakpointModel``
    objects (they supply the random streams, ``likelihood_params``, bounds, ``do_h_update`` and collect the
    ``decisions``); ``errors`` maps a fit index to the exception that took that fit out of the grid."""

    def __init__(self, backend, drivers, errors=None):
        self.be = backend
        self.drivers = drivers
        self.G = backend.G
        self.errors = errors if errors is not None else {}
        self.rounds = 0         # lock-step evaluation rounds (device calls) so far
        self.evaluations = 0    # objective evaluations they carried

    def _alive(self):
        return np.array([g not in self.errors for g in range(self.G)], dtype=bool)

    def _sample(self, g, weights=None):
        d = self.drivers[g]
        N = self.be.num_segments
        size = int(min(200, N / 10))
        rs = d.random_state if d.random_state is not None else np.random
        chosen = rs.choice(N, size=size, replace=False, p=weights)
        sample = np.zeros((N,), dtype=np.int64)
        sample[chosen] = 1
        return sample

    def _evaluate(self, **kw):
        self.rounds += 1
        act = kw.get('active')
        self.evaluations += int(np.sum(act)) if act is not None else self.G
        return self.be.evaluate(**kw)

    def _full_ell(self, alive):
        self.be.stage(None)
        return self._evaluate(active=alive)[:, 0]

    # ------------------------------------------------------------------------------------------ update_h
    def update_h(self):
        """``update_h`` (``cn_model.py:482-531``) of every fit."""
        be, M = self.be, self.be.num_clones
        alive = self._alive()
        if not alive.any():
            return
        h_before = be.get_h()
        ell_before = self._full_ell(alive)
        samples = np.zeros((self.G, be.num_segments), dtype=np.int64)
        for g in np.flatnonzero(alive):
            samples[g] = self._sample(g)
        be.stage(samples)

        workers = [int(g) for g in np.flatnonzero(alive)]
        rv = _Rendezvous(workers)
        results = {}

        def evaluate(batch):
            active = np.zeros(self.G, dtype=np.uint8)
            hv = np.array(h_before)
            for g, x in batch.items():
                active[g] = 1
                hv[g] = x
            ell, grad = self._evaluate(active=active, h=hv, grad=True)
            return dict((g, (float(ell[g, 0]), grad[g].copy())) for g in batch)

        def optimise(g):
            cache = {}

            def both(h):
                key = h.tobytes()
                if key not in cache:
                    cache.clear()
                    cache[key] = rv.request(g, np.array(h, dtype=float))
                return cache[key]

            def nll(h):
                return -both(h)[0]

            def nll_grad(h):
                return -both(h)[1]

            try:
                results[g] = scipy.optimize.minimize(nll, h_before[g], method='L-BFGS-B', jac=nll_grad,
                                                     bounds=[(1e-8, 10.)] * M)
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
            self.drivers[g].decisions.append(('h', not rejected))
            if rejected:
                final[g] = h_before[g]
        be.set_h(final, alive)

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

    def update_param(self, name):
        """``update_param(name)`` (``cn_model.py:533-569``) of every fit."""
        be = self.be
        alive = self._alive()
        if not alive.any():
            return
        lo, hi = self.drivers[0].likelihood_param_bounds[name]
        value_before = be.get_param(name)
        ell_before = self._full_ell(alive)
        samples = np.zeros((self.G, be.num_segments), dtype=np.int64)
        for g in np.flatnonzero(alive):
            try:
                samples[g] = self._sample(g, self._weights(name, g))
            except Exception as e:      # e.g. fewer non-zero weights than the sample size
                self.errors[g] = e
        alive = self._alive()
        if not alive.any():
            return
        be.stage(samples)

        # scipy.optimize.brute: the grid, ascending, one call for all fits (values outside the bounds are inf and
        # are not assigned -- cn_model.py:541-543)
        grid = np.mgrid[slice(lo, hi, complex(20))]
        inside = [float(v) for v in grid if not (v < lo or v > hi)]
        J = np.full((self.G, grid.shape[0]), np.inf)
        if inside:
            ell = self._evaluate(active=alive.astype(np.uint8), name=name,
                                 values=np.broadcast_to(np.array(inside), (self.G, len(inside))))
            self.rounds += len(inside) - 1
            self.evaluations += (len(inside) - 1) * int(alive.sum())
            J[:, [i for i, v in enumerate(grid) if not (v < lo or v > hi)]] = -ell
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

        ell_after = self._full_ell(alive)
        final = np.array(value_before)
        for g in np.flatnonzero(alive):
            rejected = ell_after[g] < ell_before[g]
            self.drivers[g].decisions.append((name, not rejected))
            final[g] = value_before[g] if rejected else result[g]
        be.set_param(name, final, alive)

    # ------------------------------------------------------------------------------------------ one EM iteration
    def run(self):
        """``em_update_h`` + ``em_update_params`` (``cn_model.py:411-416``) for the whole grid."""
        self.be.new_iteration()
        d0 = self.drivers[0]
        if d0.do_h_update:
            self.update_h()
        for name in d0.likelihood_params:
            self.update_param(name)
