"""Lock-step M-step (remixt_b200/mstep.py) on the CPU: the Nelder-Mead coroutine against scipy's own fmin, and the
lock-step grid driver against the per-fit driver on the compiled reference model (cn_model.py:482-569)."""
import numpy as np
import pytest
import scipy.optimize

from remixt_b200 import mstep
from tests import util


def _drive(gen_fn, func):
    trace = []
    gen = gen_fn()
    x = next(gen)
    while True:
        f = func(x)
        trace.append((x, f))
        try:
            x = gen.send(f)
        except StopIteration as stop:
            return stop.value, trace


FUNCS = [
    lambda x: (x - 3.) ** 2,
    lambda x: abs(x - 700.) ** 1.5 + 3.,
    lambda x: np.cos(x / 50.) * 10. + (x / 400.) ** 2,
    lambda x: np.inf if (x < 10. or x > 2000.) else -1e3 * np.log(x) + x,       # bounded like the dispersions
    lambda x: np.inf if (x < 10. or x > 2000.) else (2000. - x) * 1e-3,         # minimum on the upper bound
    lambda x: 5.,                                                                # flat
    lambda x: np.floor(x / 7.) * 1e-5,                                           # plateaus: ties and shrinks
]


@pytest.mark.parametrize('k', range(len(FUNCS)))
@pytest.mark.parametrize('x0', [0., 1., 10., 1476.3157894736842, 2000., -4.5])
def test_neldermead_coroutine_equals_scipy_fmin(k, x0):
    func = FUNCS[k]
    ref_trace = []

    def logged(x):
        assert x.shape == (1,)
        f = func(float(x[0]))
        ref_trace.append((float(x[0]), f))
        return f

    xopt, fopt, it, calls, warn = scipy.optimize.fmin(logged, x0, full_output=1, disp=0)
    (x, fval, iterations, funcalls, warnflag), trace = _drive(lambda: mstep.neldermead_1d(x0), func)
    assert [t[0] for t in trace] == [t[0] for t in ref_trace]      # same points, same order, bit for bit
    assert (x, iterations, funcalls, warnflag) == (float(xopt[0]), it, calls, warn)
    assert fval == fopt or (np.isnan(fval) and np.isnan(fopt))


def test_native_choice_equals_numpy_randomstate_choice():
    """rmx_host_sample_segments == RandomState.choice(n, size, replace=False, p) draw by draw, and leaves every stream
    in the state numpy leaves it in (so all later draws agree too)."""
    n, size, G = 3011, 200, 5
    rng = np.random.RandomState(3)
    plans = []
    for g in range(G):
        w1 = rng.rand(n) ** 8          # spiky: many collisions, several rounds of the rejection loop
        w2 = rng.rand(n)
        w2[rng.rand(n) < 0.9] = 0.     # barely more non-zero entries than the sample size
        plans.append([None, w1 / w1.sum(), w2 / w2.sum(), None, w1 / w1.sum()])
    weights = [[plans[g][d] for g in range(G)] for d in range(5)]
    mine = [np.random.RandomState(100 + g) for g in range(G)]
    ref = [np.random.RandomState(100 + g) for g in range(G)]
    for rs in mine + ref:
        rs.rand(1000 + 7)              # somewhere inside the stream, position not a multiple of 624
    chosen, status = mstep.native_choice(mine, n, size, weights, num_threads=3)
    assert not status.any()
    for g in range(G):
        for d in range(5):
            want = ref[g].choice(n, size=size, replace=False, p=plans[g][d])
            assert np.array_equal(chosen[d, g], want), (g, d)
        assert np.array_equal(mine[g].get_state()[1], ref[g].get_state()[1]) and mine[g].get_state()[2] == ref[g].get_state()[2]
        assert mine[g].rand() == ref[g].rand()
    # numpy's errors
    bad = np.zeros(n)
    bad[:50] = 1. / 50
    _, st = mstep.native_choice([np.random.RandomState(1)], n, size, [[bad], [None]])
    assert st[0, 0] == 4 and st[1, 0] == -1
    with pytest.raises(ValueError, match='Fewer non-zero'):
        np.random.RandomState(1).choice(n, size=size, replace=False, p=bad)
    _, st = mstep.native_choice([np.random.RandomState(1)], n, size, [[np.full(n, 2. / n)]])
    assert st[0, 0] == 1


def test_weights_block_equals_per_fit_weights():
    """BatchBackend.weights_block (one strided (G, N) block) == the per-fit ``w / w.sum()`` of get_param_sample_weight,
    bit for bit (numpy's pairwise sum runs over the same strided elements in both forms)."""
    class FakeBatch(object):
        num_fits, num_segments, num_clones = 6, 30105, 3
    be = mstep.BatchBackend(FakeBatch())
    rng = np.random.RandomState(2)
    be._cache['p_outlier_total'] = rng.rand(6, 30105, 2) ** 3
    be._cache['p_outlier_total'][4, :, 1] = 0.          # nothing to weight by: uniform sample
    out = np.empty((6, 30105))
    flags = be.weights_block('negbin_r_1', list(range(6)), out)
    assert list(flags) == [1, 1, 1, 1, 0, 1]
    for g in (0, 1, 2, 3, 5):
        w = be.outlier_weights('negbin_r_1', g)
        assert np.array_equal(out[g], w / w.sum())
    sub = np.empty((2, 30105))
    be.weights_block('negbin_r_1', [1, 5], sub)
    assert np.array_equal(sub[0], out[1]) and np.array_equal(sub[1], out[5])


def test_rendezvous_batches_every_round():
    import threading
    workers = list(range(7))
    rv = mstep._Rendezvous(workers)
    seen = []
    out = {}

    def evaluate(batch):
        seen.append(sorted(batch))
        return dict((g, x * 2) for g, x in batch.items())

    def work(g):
        acc = 0
        for i in range(g + 1):          # worker g makes g + 1 requests
            acc += rv.request(g, i)
        out[g] = acc
        rv.finish(g)

    ts = [threading.Thread(target=work, args=(g,)) for g in workers]
    for t in ts:
        t.start()
    rv.serve(evaluate)
    for t in ts:
        t.join()
    assert out == dict((g, sum(2 * i for i in range(g + 1))) for g in workers)
    assert seen == [list(range(r, 7)) for r in range(7)]       # round r holds exactly the workers still running


def test_lbfgsb_coroutine_reproduces_scipy_minimize():
    """The generator over scipy's compiled setulb visits the points scipy.optimize.minimize(method='L-BFGS-B') visits
    and returns its result (checked on this scipy; GridMStep falls back to blocking scipy calls on threads if not)."""
    assert mstep.lbfgsb_coroutine_ok()


@pytest.mark.slow
@pytest.mark.parametrize('lbfgsb', ['coroutine', 'threads'])
def test_lockstep_equals_per_fit_driver_on_reference_model(ref_bpmodel, lbfgsb):
    """G reference models advanced by GridMStep == the same models advanced one by one by BreakpointModel's own
    update_h / update_param: h, dispersions, decisions bit for bit, and the same objective evaluations in order."""
    exp = util.load_experiment('small')
    G = 3

    class Logged(object):
        """RemixtModel proxy that records calculate_expected_log_likelihood calls."""
        def __init__(self, model, log):
            object.__setattr__(self, '_m', model)
            object.__setattr__(self, '_log', log)

        def __getattr__(self, k):
            return getattr(self._m, k)

        def __setattr__(self, k, v):
            setattr(self._m, k, v)

        def calculate_expected_log_likelihood(self, sample):
            v = self._m.calculate_expected_log_likelihood(sample)
            self._log.append((int(np.sum(sample)), v))
            return v

    def prepared(seed_offset):
        ds = []
        for g in range(G):
            d = util.make_driver(exp, util.init_params('small', 5 * g + 1), model_class=ref_bpmodel.RemixtModel,
                                 random_state=np.random.RandomState(77))
            for _ in range(2):
                d._variational_update_stepwise()
            ds.append(d)
        return ds

    a, b = prepared(0), prepared(0)
    logs_a = [[] for _ in range(G)]
    for g, d in enumerate(a):
        d.model = Logged(d.model, logs_a[g])
        d.em_update_h()
        d.em_update_params()
    be = mstep.ModelListBackend([d.model for d in b])
    ms = mstep.GridMStep(be, b, lbfgsb=lbfgsb)
    ms.run()
    assert not ms.errors
    for g in range(G):
        assert np.array_equal(np.asarray(a[g].model.h), np.asarray(b[g].model.h))
        for name in a[g].likelihood_params:
            assert getattr(a[g].model, name) == getattr(b[g].model, name), name
        assert a[g].decisions == b[g].decisions
        assert [v for _, v in logs_a[g]] == [v for _, _, v in be.log[g]]
    # the lock-step driver needed far fewer evaluation rounds than evaluations
    assert ms.rounds < ms.evaluations
