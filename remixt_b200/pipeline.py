"""Fit orchestration: the mirror of ``remixt/analysis/pipeline.py`` ``fit`` (``:127-228``) and of
the solution selection of ``store_optimal_solution`` (``:253-264``), plus an in-process grid
driver that replaces the reference's pypeliner process fan-out over ``init_id``
(``workflow.py:329-340``) with one host thread per GPU.

An *experiment* is anything with attributes ``x (N,3)``, ``l (N,)``, ``adjacencies``,
``breakpoints`` (dict id -> frozenset of two (segment, side)) and ``segment_chromosome_id``;
the reference's ``remixt.analysis.experiment.Experiment`` qualifies, and so does
:class:`ArrayExperiment`, which is built from the plain arrays of the ``tests/golden`` fixtures.
"""
import threading

import numpy as np

from . import cn_model

# the fit-relevant subset of remixt/defaults.py:114-163
DEFAULTS = {
    'h_normal': None,
    'h_tumour': None,
    'max_copy_number': 12,
    'tumour_mix_fractions': [0.45, 0.3, 0.2, 0.1],
    'min_ploidy': 1.5,
    'max_ploidy': 6.0,
    'max_prop_diverge': 0.5,
    'normal_contamination': True,
    'likelihood_min_segment_length': 10000,
    'likelihood_min_proportion_genotyped': 0.01,
    'divergence_weights': [1e-6, 1e-7, 1e-8],
    'num_em_iter': 5,
    'num_update_iter': 5,
    'disable_breakpoints': False,
    'do_h_update': True,
    'is_female': True,
}


def get_param(config, name):
    return config[name] if name in config else DEFAULTS[name]


class ArrayExperiment(object):
    """Experiment container over plain arrays (see ``oracle/make_inputs.py`` for the layout)."""

    def __init__(self, x, l, adjacencies, breakpoints, breakpoint_ids=None, chromosome=None, start=None, end=None,
                 major_is_allele_a=None):
        self.x = np.asarray(x, dtype=float)
        self.l = np.asarray(l, dtype=float)
        n = self.l.shape[0]
        # coordinates feed only the result tables and the (unused) high_quality column of the depth table
        self.segment_start = np.zeros(n, dtype=int) if start is None else np.asarray(start)
        self.segment_end = np.maximum(self.l.astype(int) - 1, 0) if end is None else np.asarray(end)
        self.segment_major_is_allele_a = (np.ones(n, dtype=int) if major_is_allele_a is None
                                          else np.asarray(major_is_allele_a))
        self.adjacencies = set((int(a), int(b)) for a, b in np.asarray(adjacencies).reshape(-1, 2))
        bps = np.asarray(breakpoints).reshape(-1, 4)
        if breakpoint_ids is None:
            breakpoint_ids = [str(i) for i in range(bps.shape[0])]
        self.breakpoints = dict()
        for bp_id, (n1, s1, n2, s2) in zip(breakpoint_ids, bps):
            self.breakpoints[str(bp_id)] = frozenset([(int(n1), int(s1)), (int(n2), int(s2))])
        self.segment_chromosome_id = (np.asarray(chromosome) if chromosome is not None
                                      else np.array(['1'] * self.x.shape[0]))

    @classmethod
    def from_npz(cls, path):
        d = np.load(path, allow_pickle=False)
        return cls(d['x'], d['l'], d['adjacencies'], d['breakpoints'],
                   breakpoint_ids=d['breakpoint_ids'] if 'breakpoint_ids' in d else None,
                   chromosome=d['chromosome'] if 'chromosome' in d else None)


def init(experiment, config, store=None):
    """The initialisation grid of one experiment (``analysis/pipeline.py:12-109``): candidate haploid depths from
    the modes of the minor-allele depth x tumour mix fractions, filtered by the ploidy they imply, x divergence
    weights, all sharing one ``max_depth``.  Returns ``{init_id: init_params}``.

    ``experiment`` is the object itself (the reference unpickles it from a file name); ``store`` (optional,
    ``results.open_store``) receives ``read_depth`` and ``minor_modes`` as the reference's HDF5 store does."""
    import itertools
    from . import readdepth
    min_ploidy = get_param(config, 'min_ploidy')
    max_ploidy = get_param(config, 'max_ploidy')
    h_normal = get_param(config, 'h_normal')
    h_tumour = get_param(config, 'h_tumour')
    tumour_mix_fractions = get_param(config, 'tumour_mix_fractions')
    divergence_weights = get_param(config, 'divergence_weights')
    max_copy_number = get_param(config, 'max_copy_number')
    np.random.seed(config.get('random_seed', 1234))

    read_depth = readdepth.calculate_depth(experiment)
    minor_modes = readdepth.calculate_minor_modes(read_depth)
    init_h_mono = readdepth.calculate_candidate_h_monoclonal(minor_modes, h_normal=h_normal, h_tumour=h_tumour)

    candidates = []     # (params, ploidy, max_depth)
    for mode_idx, h_mono in enumerate(init_h_mono):
        ploidy = readdepth.estimate_ploidy(h_mono, experiment, read_depth=read_depth)
        assert not np.isinf(ploidy) and not np.isnan(ploidy)
        max_depth = 2. * h_mono[0] + (max_copy_number + 0.25) * h_mono[1]
        for mix_frac in tumour_mix_fractions:
            candidates.append(({'mode_idx': mode_idx, 'h_normal': h_mono[0], 'h_tumour': h_mono[1],
                                'mix_frac': mix_frac}, ploidy, max_depth))

    def distance(ploidy):       # how far outside [min_ploidy, max_ploidy]
        if min_ploidy is not None and ploidy < min_ploidy:
            return min_ploidy - ploidy
        if max_ploidy is not None and ploidy > max_ploidy:
            return ploidy - max_ploidy
        return 0.

    keep = [distance(p) == 0. for _, p, _ in candidates]
    if not any(keep):           # bounds too strict: the candidates closest to them
        dists = [distance(p) for _, p, _ in candidates]
        keep = [a == min(dists) for a in dists]
    candidates = [c for c, k in zip(candidates, keep) if k]

    # one depth threshold for all initialisations, so that their objectives are comparable
    max_depth = min(d for _, _, d in candidates)
    depth = experiment.x[:, 2] / experiment.l
    below = np.sum((depth <= max_depth) * experiment.l) / np.sum(experiment.l)
    if below < 0.75:
        raise ValueError('Unable to model {} of the genome, consider reducing max ploidy or increasing max copy number'
                         .format(1. - below))

    init_params = []
    for (h_p, _, _), w in itertools.product(candidates, divergence_weights):
        params = h_p.copy()
        params['divergence_weight'] = w
        params['max_depth'] = max_depth
        init_params.append(params)
    if store is not None:
        import pandas as pd
        store['read_depth'] = read_depth
        store['minor_modes'] = pd.Series(minor_modes, index=range(len(minor_modes)))
    return dict(enumerate(init_params))


def build_model(experiment, init_params, config, **model_kwargs):
    """The ``BreakpointModel`` construction of ``analysis/pipeline.py:164-189``."""
    is_female = get_param(config, 'is_female')
    normal_copies = np.ones((experiment.l.shape[0], 2), dtype=int)
    if not is_female:
        on_x = experiment.segment_chromosome_id == 'X'
        normal_copies[on_x, :] = np.array([1, 0])
        if np.any(experiment.x[on_x, 0:2] > 0):
            raise Exception('inconsistent allele read counts for chromosome X')

    # the dummy-segment layout depends on the experiment only: built once, adopted by every later fit
    donor = getattr(experiment, '_rmx_segmentation', None)
    model = cn_model.BreakpointModel(
        experiment.x, experiment.l, experiment.adjacencies, experiment.breakpoints,
        segmentation=donor,
        max_copy_number=get_param(config, 'max_copy_number'),
        normal_contamination=get_param(config, 'normal_contamination'),
        divergence_weight=init_params['divergence_weight'],
        min_segment_length=get_param(config, 'likelihood_min_segment_length'),
        min_proportion_genotyped=get_param(config, 'likelihood_min_proportion_genotyped'),
        max_depth=init_params['max_depth'],
        normal_copies=normal_copies,
        disable_breakpoints=get_param(config, 'disable_breakpoints'),
        breakpoint_init=None,
        do_h_update=get_param(config, 'do_h_update'),
        **model_kwargs)
    model.num_em_iter = get_param(config, 'num_em_iter')
    model.num_update_iter = get_param(config, 'num_update_iter')
    if donor is None:
        try:
            experiment._rmx_segmentation = model
        except AttributeError:      # experiment types with __slots__
            pass
    return model


def init_h(init_params):
    # analysis/pipeline.py:128-132
    return np.array([
        init_params['h_normal'],
        init_params['h_tumour'] * init_params['mix_frac'],
        init_params['h_tumour'] * (1. - init_params['mix_frac']),
    ])


def fit(experiment, init_params, config, **model_kwargs):
    """One fit from one initialisation; returns the reference's ``fit_results`` dict
    (``analysis/pipeline.py:127-228``)."""
    model = build_model(experiment, init_params, config, **model_kwargs)
    model.fit(init_h(init_params))

    cn = np.zeros((model.model.num_segments, model.model.num_clones, model.model.num_alleles), dtype=int)
    model.model.infer_cn(cn)
    return _fit_results(model, experiment, init_params, config, cn)


def select_solution(stats_by_init, config=None):
    """``init_id`` the reference's collate step would pick (``analysis/pipeline.py:253-259``):
    highest ELBO among solutions with ``proportion_divergent < max_prop_diverge`` (all solutions
    if none qualifies); ties keep the earlier ``init_id`` (stable sort)."""
    max_prop_diverge = get_param(config or {}, 'max_prop_diverge')
    ids = list(stats_by_init.keys())
    ok = [i for i in ids if stats_by_init[i]['proportion_divergent'] < max_prop_diverge]
    pool = ok if ok else ids
    best = pool[0]
    for i in pool[1:]:
        if stats_by_init[i]['elbo'] > stats_by_init[best]['elbo']:
            best = i
    return best


def fit_grid(experiment, init_params_by_id, config, device=None, random_seed=1234, num_threads=4,
             init_ids=None):
    """Fit the initialisations ``init_ids`` (default: all) of the grid on ONE GPU, ``num_threads``
    fits in flight at a time (one host thread and one CUDA stream each), and pick the solution.

    The reference runs each ``init_id`` as its own process (``workflow.py:329-340``).  Here every
    fit draws its M-step subsamples from its own ``RandomState(random_seed)``, which is the stream
    a process that called ``numpy.random.seed(random_seed)`` before ``fit`` would see, so results
    do not depend on how fits interleave.  Across GPUs the grid is sharded by rank
    (``init_ids = ids[rank::world_size]``) with no collective; rank 0 gathers the stats.
    """
    ids = list(init_params_by_id.keys()) if init_ids is None else list(init_ids)
    results = {}
    errors = {}
    model_kwargs = {} if device is None else {'device': device}

    def worker(slot):
        for pos in range(slot, len(ids), num_threads):
            init_id = ids[pos]
            try:
                results[init_id] = fit(experiment, init_params_by_id[init_id], config, model_kwargs=model_kwargs,
                                       random_state=np.random.RandomState(random_seed))
            except Exception as e:  # a failed initialisation must not take the grid down
                errors[init_id] = e

    num_threads = max(1, min(num_threads, len(ids)))
    if num_threads == 1:
        worker(0)
    else:
        threads = [threading.Thread(target=worker, args=(s,)) for s in range(num_threads)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    if errors and not results:
        raise next(iter(errors.values()))
    stats = dict((i, results[i]['stats']) for i in ids if i in results)
    best = select_solution(stats, config) if stats else None
    return results, best, errors


def shard_ids(ids, rank, world_size):
    """Initialisations of the grid that rank ``rank`` fits: a strided slice, so every rank gets a
    mix of (mode, mix_frac, divergence_weight) combinations of similar cost."""
    return list(ids)[rank::world_size]


def fit_grid_distributed(experiment, init_params_by_id, config, dist=None, device=None, random_seed=1234,
                         grid_fit=None, **fit_kwargs):
    """The init grid over the ranks of a ``torch.distributed`` process group (one process per GPU).

    The grid is the unit that shards: fits are independent, so there is no data-path
    collective; every rank fits ``shard_ids(ids, rank, world)`` on its own device as ONE batch
    (:func:`fit_grid_batched`: shared data, every kernel launch serves all of the rank's fits, lock-step
    M-step) and the per-fit ``stats`` dicts (about ten scalars each) are exchanged once at the end with
    ``all_gather_object``.  Every rank then applies the reference's selection rule
    (``analysis/pipeline.py:253-259``) to the full table, so all ranks return the same ``best`` id.
    The reference fans ``init_id`` out one process per initialisation (``workflow.py:329-340``).

    ``grid_fit`` (tests) replaces :func:`fit_grid_batched` as the per-rank grid driver.
    Returns ``(local_results, all_stats, best, all_errors)``.
    """
    ids = list(init_params_by_id.keys())
    rank = dist.get_rank() if dist is not None else 0
    world = dist.get_world_size() if dist is not None else 1
    mine = shard_ids(ids, rank, world)
    results, errors = {}, {}
    if mine:
        try:
            results, _, errors = (grid_fit or fit_grid_batched)(
                experiment, init_params_by_id, config, device=device, random_seed=random_seed, init_ids=mine,
                **fit_kwargs)
        except Exception as e:          # every fit of this rank failed: the other ranks' fits still count
            errors = dict((i, e) for i in mine)
    local_stats = dict((i, results[i]['stats']) for i in mine if i in results)
    local_errors = dict((i, repr(e)) for i, e in errors.items())
    if dist is not None and world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, (local_stats, local_errors))
        all_stats, all_errors = {}, {}
        for st, er in gathered:
            all_stats.update(st)
            all_errors.update(er)
    else:
        all_stats, all_errors = local_stats, local_errors
    ordered = dict((i, all_stats[i]) for i in ids if i in all_stats)
    best = select_solution(ordered, config) if ordered else None
    return results, ordered, best, all_errors


def _fit_results(model, experiment, init_params, config, cn, pre=None):
    """The ``fit_results`` dict of ``analysis/pipeline.py:195-228`` from a fitted driver and its MAP copy number.
    ``pre`` (grid drivers) carries quantities already read for the whole grid in one device call."""
    pre = pre or {}
    brk_cn = pre['brk_cn'] if 'brk_cn' in pre else model.decode_breakpoints(cn)
    fwd = model.seg_fwd_remap
    cn = cn[fwd]
    if get_param(config, 'disable_breakpoints'):
        brk_cn = cn_model.decode_breakpoints_naive(cn, experiment.adjacencies, experiment.breakpoints)
    h = pre['h'] if 'h' in pre else model.h
    results = dict()
    results['h'] = h
    results['cn'] = cn
    results['brk_cn'] = brk_cn
    results['p_outlier_total'] = pre['p_outlier_total'][fwd] if 'p_outlier_total' in pre else model.p_outlier_total
    results['p_outlier_allele'] = pre['p_outlier_allele'][fwd] if 'p_outlier_allele' in pre else model.p_outlier_allele
    results['total_likelihood_mask'] = (pre['total_likelihood_mask'][fwd] if 'total_likelihood_mask' in pre
                                        else model.total_likelihood_mask)
    results['allele_likelihood_mask'] = (pre['allele_likelihood_mask'][fwd] if 'allele_likelihood_mask' in pre
                                         else model.allele_likelihood_mask)
    stats = dict()
    stats['elbo'] = model.prev_elbo
    stats['elbo_diff'] = model.prev_elbo_diff
    stats['error_message'] = ''
    stats.update(pre['params'] if 'params' in pre else model.get_likelihood_param_values())
    l = experiment.l
    stats['num_clones'] = len(h)
    stats['num_segments'] = len(experiment.x)
    tumour = cn[:, 1:, :]
    stats['ploidy'] = (tumour.mean(axis=1).T * l).sum() / l.sum()
    divergent = (tumour.max(axis=1) != tumour.min(axis=1)) * 1.
    stats['proportion_divergent'] = (divergent.T * l).sum() / (2. * l.sum())
    stats['mode_idx'] = init_params.get('mode_idx', 0)
    stats['divergence_weight'] = init_params['divergence_weight']
    results['stats'] = stats
    # extension (not a reference key): the outcome of every M-step acceptance test, in call order
    results['decisions'] = list(getattr(model, 'decisions', []))
    return results


def decode_breakpoints_grid(driver, cn_all, brk_states, transition_penalty):
    """``BreakpointModel.decode_breakpoints`` (``cn_model.py:578-596``) for all G fits of a grid at once: the
    breakend rows, orientations and breakpoint states are shared, so the reference's accumulation (rows ascending,
    clones ascending -- kept, it fixes the rounding of every partial sum) runs once over ``(G, B)`` slabs instead
    of G times over ``(B,)`` vectors.  Returns one ``{breakpoint_id: brk_states row}`` dict per fit."""
    G = cn_all.shape[0]
    bidx, borient = driver.breakpoint_idx, driver.breakpoint_orient
    K, B, M = driver.num_breakpoints, brk_states.shape[0], brk_states.shape[1]
    rows = np.flatnonzero(bidx[:cn_all.shape[1] - 1] >= 0)
    log_p = np.zeros((G, K, B))
    if rows.size:
        tot_a = cn_all[:, rows].sum(axis=-1)            # (G, R, M)
        tot_b = cn_all[:, rows + 1].sum(axis=-1)
        d = tot_a - tot_b
        for r, n in enumerate(rows):
            for m in range(M):
                log_p[:, bidx[n], :] += -transition_penalty * np.abs(d[:, r, m][:, np.newaxis] - borient[n] * brk_states[:, m])
    best = log_p.argmax(axis=-1)                        # (G, K), first maximum as the reference's argmax
    return [dict((driver.breakpoint_ids[k], brk_states[best[g, k]]) for k in range(K)) for g in range(G)]


class _ModelListGrid(object):
    """G independent single-fit models behind the part of the ``RemixtBatch`` surface the grid driver uses."""

    def __init__(self, models):
        self.models = models
        self.num_segments = models[0].num_segments
        self.num_clones = models[0].num_clones

    def variational_update(self, n_iter):
        for m in self.models:
            for _ in range(n_iter):
                for step in ('update_p_allele_swap', 'update_p_cn', 'update_p_breakpoint', 'update_p_outlier_total',
                             'update_p_outlier_allele'):
                    getattr(m, step)()

    def calculate_elbo(self):
        return np.array([m.calculate_elbo() for m in self.models])

    def infer_cn(self, cn):
        for g, m in enumerate(self.models):
            out = np.zeros(cn.shape[1:], dtype=int)
            m.infer_cn(out)
            cn[g] = out

    def close(self):
        pass


#: fit_grid_batched fits a grid as interleaved sub-batches of at least this many fits, three at most, each on its own
#: host thread and stream
SUB_BATCH_MIN_FITS = 16
SUB_BATCH_MAX = 3


def fit_grid_batched(experiment, init_params_by_id, config, device=None, random_seed=1234, init_ids=None,
                     mstep='lockstep', num_threads=8, profile=None, model_class=None, trace=None, sub_batches=None):
    """The initialisation grid on one GPU.  ``sub_batches`` = K > 1 splits the grid into K interleaved groups
    (``ids[k::K]``), each ONE batch on its own stream driven by its own host thread (:func:`_fit_grid_one_batch`): the
    lock-step M-step is half host work (the optimisers' bookkeeping in Python / scipy) and half short device calls, so
    one group's host work runs under the other's device work (the C-ABI calls release the GIL), and one group's
    E-step kernels fill the SMs its neighbour's M-step leaves idle.  A fit's trajectory does not depend on the batch
    it rides in (its own random stream, every kernel per fit), so results, statistics and the selected solution are
    those of one batch.  Default for lock-step grids on the device model: one group per ``SUB_BATCH_MIN_FITS`` fits,
    ``SUB_BATCH_MAX`` at most (48 fits: 3 x 16; measured on config 2: 0.637 / 0.573 / 0.560 / 0.584 s for K = 1 / 2 / 3 /
    4); ``config['sub_batches']`` overrides.  With K > 1 ``profile`` holds the largest per-phase time of a
    group (the phases of different groups overlap) and ``sub_batches``.  Returns ``(results, best, errors)``."""
    ids = list(init_params_by_id.keys()) if init_ids is None else list(init_ids)
    K = sub_batches if sub_batches is not None else config.get('sub_batches')
    if K is None:
        K = (min(SUB_BATCH_MAX, len(ids) // SUB_BATCH_MIN_FITS)
             if (model_class is None and trace is None and mstep == 'lockstep') else 1)
    K = max(1, min(int(K), len(ids)))
    if K == 1 or trace is not None:
        return _fit_grid_one_batch(experiment, init_params_by_id, config, device=device, random_seed=random_seed,
                                   init_ids=ids, mstep=mstep, num_threads=num_threads, profile=profile,
                                   model_class=model_class, trace=trace)
    import concurrent.futures
    groups = [ids[k::K] for k in range(K)]
    profs = [dict() for _ in groups]

    def one(k):
        try:
            return _fit_grid_one_batch(experiment, init_params_by_id, config, device=device, random_seed=random_seed,
                                       init_ids=groups[k], mstep=mstep, num_threads=num_threads, profile=profs[k],
                                       model_class=model_class, select=False)
        except Exception as e:          # every fit of the group failed (or its batch could not be built)
            return {}, None, dict((i, e) for i in groups[k])

    with concurrent.futures.ThreadPoolExecutor(K) as pool:
        parts = list(pool.map(one, range(K)))
    results, errors = {}, {}
    for r, _, e in parts:
        results.update(r)
        errors.update(e)
    if not results:
        raise next(iter(errors.values()))
    import time
    t_res = time.perf_counter()
    pending = [i for i in results if callable(results[i])]
    from . import mstep as _mstep
    with concurrent.futures.ThreadPoolExecutor(max(1, min(len(pending), _mstep._host_threads()))) as pool:
        for i, r in zip(pending, pool.map(lambda i: results[i](), pending)):
            results[i] = r
    t_res = time.perf_counter() - t_res
    results = dict((i, results[i]) for i in ids if i in results)      # the caller's order
    stats = dict((i, results[i]['stats']) for i in results)
    best = select_solution(stats, config) if stats else None
    if profile is not None:
        for key in set().union(*[set(p) for p in profs]):
            vals = [p[key] for p in profs if key in p]
            if isinstance(vals[0], dict):
                profile[key] = dict((kk, max(v.get(kk, 0.) for v in vals)) for kk in vals[0])
            elif key in ('mstep_rounds', 'mstep_evaluations'):
                profile[key] = sum(vals)
            else:
                profile[key] = max(vals)
        profile['results'] = profile.get('results', 0.) + t_res
        profile['sub_batches'] = K
    return results, best, errors


def _fit_grid_one_batch(experiment, init_params_by_id, config, device=None, random_seed=1234, init_ids=None,
                        mstep='lockstep', num_threads=8, profile=None, model_class=None, trace=None, select=True):
    """The given initialisations as ONE batch on one GPU: every E-step kernel launch serves all fits
    (``bpmodel.RemixtBatch``) and the M-steps of all fits advance in lock step (``mstep.GridMStep``: every round of
    objective evaluations of ``cn_model.py:482-569`` is one device call for the whole grid).

    Per fit this performs exactly the calls ``fit`` performs, in the same order, with the M-step
    subsamples drawn from ``RandomState(random_seed)`` per fit (the stream a process that called
    ``numpy.random.seed(random_seed)`` before ``pipeline.fit`` would see).  ``mstep='threads'`` keeps the per-fit
    scipy drivers of ``cn_model.BreakpointModel`` on ``num_threads`` host threads (A/B measurements).
    ``profile`` (a dict) receives wall-clock seconds per phase.  ``model_class`` (tests) runs the same driver over G
    independent models of another ``RemixtModel`` implementation.  ``trace`` (a list) receives ``(h, dispersions,
    elbo)`` of all fits after every EM iteration.  Returns ``(results, best, errors)`` like
    :func:`fit_grid`; a fit that fails is reported in ``errors`` and leaves the grid.
    """
    import time
    from . import bpmodel
    from . import mstep as _mstep
    prof = profile if profile is not None else {}
    for k in ('build', 'e_step', 'm_step', 'elbo', 'decode', 'results'):
        prof.setdefault(k, 0.)
    t0 = time.perf_counter()
    ids = list(init_params_by_id.keys()) if init_ids is None else list(init_ids)
    drivers = [build_model(experiment, init_params_by_id[i], config, random_state=np.random.RandomState(random_seed))
               for i in ids]
    h = np.array([init_h(init_params_by_id[i]) for i in ids])
    dw = np.array([init_params_by_id[i]['divergence_weight'] for i in ids], dtype=float)
    kwargs = {} if device is None else {'device': device}
    d0 = drivers[0]
    if model_class is not None:
        # tests: G independent models of another RemixtModel implementation (the compiled reference, the oracle)
        # behind the same grid driver
        for d, hg in zip(drivers, h):
            d.model_class = model_class
            d.attach(d._make_model(hg))
        batch = _ModelListGrid([d.model for d in drivers])
        backend = _mstep.ModelListBackend(batch.models)
    else:
        batch = bpmodel.RemixtBatch(*(d0.model_args(h, compact=True) + (dw,)), **kwargs)
        backend = _mstep.BatchBackend(batch)
        shared = all(np.array_equal(d._total_likelihood_mask, d0._total_likelihood_mask) and
                     np.array_equal(d._allele_likelihood_mask, d0._allele_likelihood_mask) and
                     d.transition_model == d0.transition_model and d.breakpoint_init is None for d in drivers)
        if shared:      # the grid shares max_depth (analysis/pipeline.py:85-101): one upload serves every fit
            batch.total_likelihood_mask = d0._total_likelihood_mask.astype(int)
            batch.allele_likelihood_mask = d0._allele_likelihood_mask.astype(int)
            batch.transition_model = d0.transition_model
            for g, d in enumerate(drivers):
                d.model = batch.fit(g)
        else:
            for g, d in enumerate(drivers):
                d.attach(batch.fit(g))
    num_em_iter = get_param(config, 'num_em_iter')
    num_update_iter = get_param(config, 'num_update_iter')
    # The ELBO of the starting state (cn_model.py:407) only ever enters the FIRST iteration's `elbo_diff`; with two or
    # more EM iterations nothing `fit` returns depends on it.  On the GPU it is the one expensive ELBO of a fit (the
    # uniform pair posteriors of the constructor state take the general pair sums over every row: 42 ms for the
    # 48-fit config-2 grid, 6 % of the whole fit), so it is evaluated only when it is observable.
    if model_class is not None or num_em_iter < 2 or trace is not None or any(d.verbose for d in drivers):
        elbo = batch.calculate_elbo()
        for g, d in enumerate(drivers):
            d.prev_elbo = float(elbo[g])
    else:
        for d in drivers:
            d.prev_elbo = float('nan')
    errors = {}         # fit index -> exception
    lock = _mstep.GridMStep(backend, drivers, errors)
    prof['build'] += time.perf_counter() - t0

    def m_step_threads(slot):
        for g in range(slot, len(drivers), num_threads):
            if g in errors:
                continue
            try:
                if drivers[g].do_h_update:
                    drivers[g].em_update_h()
                drivers[g].em_update_params()
            except Exception as e:   # a failed initialisation must not take the grid down
                errors[g] = e

    for i in range(num_em_iter):
        t0 = time.perf_counter()
        batch.variational_update(num_update_iter)
        t1 = time.perf_counter()
        if mstep == 'lockstep':
            lock.run()
        else:
            nt = max(1, min(num_threads, len(drivers)))
            threads = [threading.Thread(target=m_step_threads, args=(s,)) for s in range(nt)]
            for t in threads:
                t.start()
            for t in threads:
                t.join()
        t2 = time.perf_counter()
        elbo = batch.calculate_elbo()
        for g, d in enumerate(drivers):
            d.record_iteration(i, float(elbo[g]))
        if trace is not None:       # per EM iteration: h (G,M), the dispersions (G,P), ELBO (G)
            trace.append((backend.get_h(), np.array([backend.get_param(n) for n in d0.likelihood_params]).T.copy(),
                          np.array(elbo, dtype=float)))
        t3 = time.perf_counter()
        prof['e_step'] += t1 - t0
        prof['m_step'] += t2 - t1
        prof['elbo'] += t3 - t2
    t0 = time.perf_counter()
    cn = np.zeros((len(drivers), batch.num_segments, batch.num_clones, 2), dtype=np.int64)
    batch.infer_cn(cn)
    t1 = time.perf_counter()
    results = {}
    pre = [dict() for _ in drivers]
    if model_class is None:
        # one device read per quantity for the whole grid instead of one per fit
        brk = decode_breakpoints_grid(d0, cn, np.asarray(batch.brk_states), abs(d0.transition_log_prob))
        h_all, pot, poa = np.array(batch.h), batch.p_outlier_total, batch.p_outlier_allele
        par = dict((name, np.array(getattr(batch, name))) for name in d0.likelihood_params)
        as_int = {}          # the masks are usually the same for the whole grid: convert each distinct one once

        def mask_int(a):
            key = a.tobytes()
            if key not in as_int:
                as_int[key] = a.astype(int)
            return as_int[key]

        for g, d in enumerate(drivers):
            pre[g] = dict(brk_cn=brk[g], h=h_all[g], p_outlier_total=pot[g], p_outlier_allele=poa[g],
                          params=dict((name, float(par[name][g])) for name in d0.likelihood_params),
                          total_likelihood_mask=mask_int(d._total_likelihood_mask),
                          allele_likelihood_mask=mask_int(d._allele_likelihood_mask))
    todo = [g for g in range(len(drivers)) if g not in errors]

    def assemble(g):
        return _fit_results(drivers[g], experiment, init_params_by_id[ids[g]], config, cn[g], pre[g])

    # index remaps and length-weighted statistics over N segments per fit: numpy releases the GIL inside them
    nt = min(len(todo), _mstep._host_threads()) if model_class is None else 1
    if not select:
        # a sub-batch of a grid: the caller assembles the results of all sub-batches in ONE pool (every group doing it
        # on its own threads at the same moment oversubscribes the host), from host arrays only
        done = [(lambda g=g: assemble(g)) for g in todo]
    elif nt > 1:
        import concurrent.futures
        with concurrent.futures.ThreadPoolExecutor(nt) as pool:
            done = list(pool.map(assemble, todo))
    else:
        done = [assemble(g) for g in todo]
    for g, r in zip(todo, done):
        results[ids[g]] = r
    prof['mstep_rounds'] = lock.rounds
    prof['mstep_evaluations'] = lock.evaluations
    prof['mstep_seconds'] = dict(lock.seconds)
    batch.close()       # results hold host arrays only; the device blocks go back to the pool now
    errors_by_id = dict((ids[g], e) for g, e in errors.items())
    if errors_by_id and not results:
        raise next(iter(errors_by_id.values()))
    best = None
    if select:
        stats = dict((i, results[i]['stats']) for i in ids if i in results)
        best = select_solution(stats, config) if stats else None
    prof['decode'] += t1 - t0
    prof['results'] += time.perf_counter() - t1
    return results, best, errors_by_id
