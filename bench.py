"""Benchmark of the ReMixT variational-inference hot path on B200 (contract: see DESIGN.md, Measurement).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME] [--fits G]

Workload: a simulated whole genome (tests/golden/exp_<workload>.npz, produced by the reference's own
simulator, oracle/make_inputs.py) fitted from every initialisation of the reference's (h, ploidy) grid
(stored in the same fixture by oracle/make_golden.py init).  One *step* is one EM iteration of the whole grid
(cn_model.py:409-419): `num_update_iter` (5, remixt/defaults.py:157) rounds of the five coordinate updates
(cn_model.py:444-460) on the likelihood table of the current (h, dispersions), the M-step -- update_h (L-BFGS-B) and
update_param for the four dispersions (brute grid + Nelder-Mead), cn_model.py:482-569 -- and the ELBO, for all G fits.

    value = N1 * S * num_update_iter * G_total / step_time      [segment-state evals / s]

`value` has the batch resident in HBM (built once, outside the timed region); `e2e` is the COMPLETE fit of the grid
through the public API from HOST arrays: batch construction (H2D), num_em_iter (5) EM iterations, MAP decode, result
assembly on the host (D2H) and the selection of the best initialisation; `full_fit` carries its wall time by phase.

N > 1 (torchrun): the init grid is SPLIT over the ranks (`ids[rank::world]`, one RemixtBatch per rank, no data-path
collective; the per-fit stats are gathered once for the selection): scaling = strong, G_total fixed at the grid size.
`replicas` repeats the resident measurement with every rank holding the whole grid (weak, round 1's number).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')
NUM_UPDATE_ITER = 5
NUM_EM_ITER = 5
METRIC = 'segment_state_evals_per_s'
UNIT = 'segment-state evals/s'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default=None)
    ap.add_argument('--fits', type=int, default=0, help='fits per GPU (0 = the whole init grid)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--cpu-sample-segments', type=int, default=400)
    ap.add_argument('--no-replicas', action='store_true', help='N > 1: skip the secondary replica (weak) measurement')
    return ap.parse_args()


def pick_workload(name):
    if name:
        return name
    for cand in ('config2', 'genome5k'):
        if os.path.exists(os.path.join(GOLDEN, 'exp_%s.npz' % cand)):
            return cand
    raise SystemExit('no workload fixture under tests/golden')


class PlainExperiment(object):
    """The attributes pipeline.init / pipeline.fit read, over the fixture's plain arrays.  Defined here (and not
    taken from remixt_b200.pipeline) so that the reference arm never imports the package under test."""

    def __init__(self, x, l, adjacencies, breakpoints, breakpoint_ids, chromosome):
        self.x = np.asarray(x, dtype=float)
        self.l = np.asarray(l, dtype=float)
        self.adjacencies = set((int(a), int(b)) for a, b in np.asarray(adjacencies).reshape(-1, 2))
        self.breakpoints = dict()
        for bp_id, (n1, s1, n2, s2) in zip(breakpoint_ids, np.asarray(breakpoints).reshape(-1, 4)):
            self.breakpoints[str(bp_id)] = frozenset([(int(n1), int(s1)), (int(n2), int(s2))])
        self.segment_chromosome_id = np.asarray(chromosome)
        n = self.l.shape[0]
        self.segment_start = np.zeros(n, dtype=int)
        self.segment_end = np.maximum(self.l.astype(int) - 1, 0)
        self.segment_major_is_allele_a = np.ones(n, dtype=int)


# workloads that reuse another workload's genome with a different model configuration
DERIVED = {
    # BASELINE config 3: the config-2 genome with max copy number 8 (S = 165 states for 2 tumour clones + normal);
    # its initialisation grid comes from this repo's pipeline.init (the reference's grid builder, tests/test_init_results.py)
    'config3': ('config2', {'max_copy_number': 8, 'max_ploidy': 4.0}),
    # the reference's DEFAULT max_copy_number (analysis/pipeline.py defaults: 12 -> S = 355 states): above every
    # tensor-path chain kernel and above k_hist2 -- the L2-streamed chain kernel and the first-generation histograms
    'cn12': ('config2', {'max_copy_number': 12, 'max_ploidy': 4.0}),
}


def load_workload(name):
    if name in DERIVED:
        base, cfg = DERIVED[name]
        exp, _ = load_workload(base)
        exp._name = name
        CONFIG.update(cfg)
        from remixt_b200 import pipeline
        grid = pipeline.init(exp, dict(cfg))
        return exp, [grid[i] for i in sorted(grid)]
    path = os.path.join(GOLDEN, 'exp_%s.npz' % name)
    d = np.load(path, allow_pickle=False)
    exp = PlainExperiment(d['x'], d['l'], d['adjacencies'], d['breakpoints'], d['breakpoint_ids'], d['chromosome'])
    exp._name = name
    inits = []
    for i in range(len(d['init_h_normal'])):
        inits.append(dict(mode_idx=int(d['init_mode_idx'][i]), h_normal=float(d['init_h_normal'][i]),
                          h_tumour=float(d['init_h_tumour'][i]), mix_frac=float(d['init_mix_frac'][i]),
                          divergence_weight=float(d['init_divergence_weight'][i]),
                          max_depth=float(d['init_max_depth'][i])))
    return exp, inits


def subsample_experiment(exp, nseg, start=None):
    """A window of `nseg` consecutive segments of the workload with the adjacencies / breakpoints that lie wholly inside,
    renumbered from 0.  start=None picks, among the windows starting at multiples of nseg, the one holding the most
    breakpoints (the reference's BreakpointModel cannot be built without any, cn_model.py:59)."""
    n_all = exp.x.shape[0]
    nseg = min(nseg, n_all)

    def inside(s0):
        out = []
        for bp_id, bp in exp.breakpoints.items():
            (n1, s1), (n2, s2) = tuple(bp)
            if (s0 <= n1 < s0 + nseg and s0 <= n2 < s0 + nseg and not (n1 == s0 and s1 == 0) and
                    not (n2 == s0 and s2 == 0)):
                out.append((bp_id, (n1 - s0, s1, n2 - s0, s2)))
        return out

    if start is None:
        start = max(range(0, n_all - nseg + 1, nseg), key=lambda s0: (len(inside(s0)), -s0))
    keep = inside(start)
    adj = np.array([(a - start, b - start) for a, b in sorted(exp.adjacencies)
                    if start <= a < start + nseg and start <= b < start + nseg], dtype=np.int64).reshape(-1, 2)
    sub = PlainExperiment(exp.x[start:start + nseg], exp.l[start:start + nseg], adj,
                          np.array([b for _, b in keep], dtype=np.int64).reshape(-1, 4), [i for i, _ in keep],
                          exp.segment_chromosome_id[start:start + nseg])
    sub._window = (int(start), int(nseg), len(keep))
    return sub


CONFIG = {'max_copy_number': 6, 'num_em_iter': NUM_EM_ITER, 'num_update_iter': NUM_UPDATE_ITER}


class ClockSampler(object):
    """nvidia-smi clock / throttle-reason sampling during the timed region (B200_PROFILING.md)."""
    QUERY = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.QUERY,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.lines:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def measured_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), 'measured'
    return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0}, 'fallback'


def dist_setup(n_gpus):
    """torch.distributed over NCCL when launched by torchrun; returns (rank, world, dist or None)."""
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        # NCCL announces its version on stdout when the first communicator comes up; stdout carries the one JSON
        # line only, so the start-up (init + first collective) runs with fd 1 pointed at stderr
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group(backend='nccl', device_id=torch.device('cuda', local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
        return rank, world, local, dist
    return 0, 1, 0, None


def grid_batch(exp, inits, device, random_seed=1234):
    """One RemixtBatch for the given initialisations with its host drivers and lock-step M-step, built from HOST arrays
    through the public constructor (the H2D copies happen inside) exactly as pipeline.fit_grid_batched builds it."""
    from remixt_b200 import bpmodel, mstep, pipeline
    drivers = [pipeline.build_model(exp, ip, CONFIG, random_state=np.random.RandomState(random_seed)) for ip in inits]
    h = np.array([pipeline.init_h(ip) for ip in inits])
    dw = np.array([ip['divergence_weight'] for ip in inits], dtype=float)
    args = drivers[0].model_args(h, compact=True) + (dw,)
    batch = bpmodel.RemixtBatch(*args, device=device)
    batch.total_likelihood_mask = drivers[0]._total_likelihood_mask.astype(int)      # the grid shares max_depth
    batch.allele_likelihood_mask = drivers[0]._allele_likelihood_mask.astype(int)
    batch.transition_model = drivers[0].transition_model
    for g, d in enumerate(drivers):
        d.model = batch.fit(g)
    lock = mstep.GridMStep(mstep.BatchBackend(batch), drivers, {})
    h2d = sum(int(np.asarray(a).nbytes) for a in args if isinstance(a, np.ndarray)) + 2 * 8 * batch.num_segments
    h2d += sum(int(a.table.nbytes + a.normal_copies.nbytes) for a in args if isinstance(a, bpmodel.CompactCnStates))
    return batch, drivers, lock, h2d


# algorithmic work per launch of each kernel class: (bound, bytes, fp64 flop)  -- DESIGN.md, Kernels
def kernel_models(N1, S, NC, G, R):
    ns = float(N1) * S * G
    return {
        'chain': ('fp64', 32. * ns, 2. * 2. * S * S * (N1 - NC) * G),
        # executed FP64-pipe instructions per (n,s) of the likelihood kernel (ncu: dadd + dmul + dfma thread
        # instructions, profiles/), each counted as one FMA = 2 flop
        'lltable': ('fp64', 48. * ns, 2. * LLTABLE_FP64_INSTR * ns),
        'frame': ('hbm', 64. * ns, 14. * ns),
        'qsums': ('hbm', 56. * ns, 12. * ns),
        'post': ('hbm', 24. * ns, 3. * ns),
        'hist': ('fp64', 0., 2. * 8. * S * S * R * G),
        'elbo_rows': ('hbm', 16. * ns, 4. * ns),
        'round': ('hbm', 88. * ns, 29. * ns),
    }


LLTABLE_FP64_INSTR = 200.
NCU_DRAM_BYTES_PER_UNIT = {'chain': 29.0, 'lltable': 48.8, 'round': 89.1}


def result_bytes(results):
    n = 0
    for r in results.values():
        for k in ('h', 'cn', 'p_outlier_total', 'p_outlier_allele', 'total_likelihood_mask', 'allele_likelihood_mask'):
            n += int(np.asarray(r[k]).nbytes)
        n += sum(int(np.asarray(v).nbytes) for v in r['brk_cn'].values())
    return n


def gpu_arm(args):
    rank, world, local, dist = dist_setup(args.gpus)
    from remixt_b200 import _lib, pipeline
    if _lib.device_count() <= 0:
        raise SystemExit('bench.py: no CUDA device; remixt_b200 has no CPU fallback')
    name = pick_workload(args.workload)
    exp, inits = load_workload(name)
    if args.fits > 0:
        inits = inits[:args.fits]
    G_total = len(inits)
    ids = list(range(G_total))
    mine = pipeline.shard_ids(ids, rank, world)         # strong scaling: the grid is split over the ranks
    my_inits = [inits[i] for i in mine]
    G = len(my_inits)
    peaks, peak_src = measured_peaks()

    def barrier(batch=None):
        if batch is not None:
            batch.synchronize()
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- resident arm: the rank's batch built once, every step = one EM iteration of its fits
    def resident(fit_inits, steps, warmup, sample_clocks):
        batch, drivers, lock, h2d = grid_batch(exp, fit_inits, local)

        def em_iteration():
            batch.variational_update(NUM_UPDATE_ITER)
            t0 = time.perf_counter()
            lock.run()
            em_iteration.m_s += time.perf_counter() - t0
            return batch.calculate_elbo()
        em_iteration.m_s = 0.

        for _ in range(warmup):
            em_iteration()
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        barrier(batch)
        batch.kernel_timing(True)
        em_iteration.m_s = 0.
        r0, e0, l0 = lock.rounds, lock.evaluations, _lib.global_launch_count()
        batch.event_record(0)
        for _ in range(steps):
            elbo = em_iteration()
        batch.event_record(1)
        barrier(batch)
        ms_timed = batch.event_elapsed(0, batch, 1)
        launches_timed = _lib.global_launch_count() - l0
        ktime = batch.kernel_timing(True)
        clocks = sampler.stop() if sampler else None
        # the likelihood table of ALL fits for a new h, on its own: inside the step the M-step also rebuilds single
        # planes of it for subsets of fits, so the step's launches of this class are not uniform
        h_now = np.array(batch.h)
        for _ in range(3):
            batch.h = h_now
            batch.update_framelogprob()
        ktime['lltable'] = batch.kernel_timing(False)['lltable']
        out = dict(ms=ms_timed, launches=launches_timed,
                   ktime=ktime, clocks=clocks,
                   m_step_ms=em_iteration.m_s * 1e3, mstep_rounds=lock.rounds - r0, mstep_evals=lock.evaluations - e0,
                   N1=batch.num_segments, S=batch.num_cn_states, K=batch.num_breakpoints,
                   NC=int(np.sum(np.asarray(batch.is_telomere)[:-1] > 0)) + 1,
                   R=int(np.sum(np.asarray(batch.breakpoint_idx)[:-1] >= 0)), h2d=h2d,
                   elbo_checksum=float(np.sum(elbo)), errors=len(lock.errors))
        batch.close()
        return out

    res = resident(my_inits, args.steps, args.warmup, True)
    N1, S, NC, R = res['N1'], res['S'], res['NC'], res['R']

    # ---------------- e2e: the complete fit of the rank's share of the grid from host arrays, through the public API
    full = []
    n_full = max(1, min(args.steps, 3))
    for rep in range(n_full + 1):            # first run untimed (pool warm-up)
        prof = {}
        barrier()
        t0 = time.perf_counter()
        results, stats, best, errors = pipeline.fit_grid_distributed(exp, dict(enumerate(inits)), CONFIG, dist=dist,
                                                                     device=local, profile=prof)
        wall = time.perf_counter() - t0
        if rep > 0:
            full.append((wall, prof, best, len(errors)))
    d2h = result_bytes(results)
    full.sort(key=lambda t: t[0])
    wall, prof, best, n_err = full[len(full) // 2]

    rep_res = None
    if world > 1 and not args.no_replicas:
        rep_res = resident(inits, max(1, min(args.steps, 3)), 1, False)

    ms, launches, m_ms = res['ms'], res['launches'], res['m_step_ms']
    if dist is not None:
        import torch
        vals = [ms, wall, m_ms] + ([rep_res['ms']] if rep_res else [])
        t = torch.tensor(vals, dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, wall, m_ms = float(t[0]), float(t[1]), float(t[2])
        if rep_res:
            rep_res['ms'] = float(t[3])
        tl = torch.tensor([launches, d2h, res['h2d']], dtype=torch.int64, device='cuda')
        dist.all_reduce(tl)
        launches, d2h, h2d = int(tl[0]), int(tl[1]), int(tl[2])
    else:
        h2d = res['h2d']
    ms_per_step = ms / args.steps
    evals_step = float(N1) * S * NUM_UPDATE_ITER * G_total
    value = evals_step / (ms_per_step * 1e-3)

    # ---------------- rooflines from the CUDA-event durations of the timed region (this rank's launches)
    try:
        fp64_peak = _lib.fp64_peak(local)
    except Exception:
        fp64_peak = None
    hbm_peak = peaks.get('hbm_gbs')
    models = kernel_models(N1, S, NC, G, R)
    ktime = res['ktime']
    total_kernel_ms = sum(v[0] for k, v in ktime.items() if k != 'lltable') + ktime['lltable'][0] / max(ktime['lltable'][1], 1)
    kernels = {}
    for k, (tot, n) in ktime.items():
        if n == 0:
            continue
        ent = {'launches': n, 'ms_per_launch': tot / n, 'share_of_kernel_time': (tot / n if k == 'lltable' else tot) / total_kernel_ms}
        if k == 'lltable':
            ent['note'] = 'full rebuild for all fits, timed on its own after the timed region (one per EM iteration in the E-step)'
        # launches of the M-step's full-data evaluations rebuild the table for the fits whose parameters changed
        # only, so per-launch work varies: rates are quoted for the E-step classes, whose launches are uniform
        if k in models and k not in ('qsums',):
            bound, nbytes, flops = models[k]
            t_s = tot / n * 1e-3
            ent['bound'] = bound
            ent['gbs'] = nbytes / t_s / 1e9
            ent['tflops'] = flops / t_s / 1e12
            ent['hbm_frac'] = ent['gbs'] / hbm_peak if hbm_peak else None
            ent['fp64_frac'] = ent['tflops'] / fp64_peak if fp64_peak else None
        kernels[k] = ent
    rated = [k for k in kernels if 'bound' in kernels[k] and k != 'lltable']
    dom = max(rated, key=lambda k: kernels[k]['share_of_kernel_time'])
    de = kernels[dom]
    if de['bound'] == 'hbm':
        roofline = {'kernel': dom, 'bound': 'hbm', 'achieved': de['gbs'], 'peak': hbm_peak, 'unit': 'GB/s',
                    'frac': de['hbm_frac'], 'traffic': None, 'peak_source': 'MEASURED_PEAKS.json (%s)' % peak_src}
    else:
        roofline = {'kernel': dom, 'bound': 'fp64', 'achieved': de['tflops'], 'peak': fp64_peak, 'unit': 'TFLOP/s',
                    'frac': de['fp64_frac'], 'traffic': None,
                    'peak_source': 'FP64 FMA micro-benchmark run in this process (rmx_fp64_peak); not in MEASURED_PEAKS.json',
                    'hbm_achieved_gbs': de['gbs'], 'hbm_peak_gbs': hbm_peak, 'hbm_frac': de['hbm_frac']}
    if dom in NCU_DRAM_BYTES_PER_UNIT:
        roofline['traffic'] = NCU_DRAM_BYTES_PER_UNIT[dom] * float(N1) * S * G
        roofline['traffic_source'] = 'ncu --set full capture, %.1f B per (segment, state)' % NCU_DRAM_BYTES_PER_UNIT[dom]
    roofline['launch_ms'] = de['ms_per_launch']
    roofline['share_of_step'] = de['ms_per_launch'] * de['launches'] / res['ms']
    roofline['fits_per_launch'] = G

    e_ms = (res['ms'] - res['m_step_ms']) / args.steps
    evals_fit = evals_step * NUM_EM_ITER
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None,
        'dtype': 'f64', 'data': 'synthetic (remixt.simulations genome, fixture exp_%s.npz)' % name,
        'config': {'workload': name, 'segments': int(exp.x.shape[0]), 'model_segments': int(N1), 'cn_states': int(S),
                   'breakpoints': int(res['K']), 'chains': NC, 'fits_total': G_total, 'fits_per_gpu': G,
                   'step': 'one EM iteration of every fit: %d VI rounds + M-step (h, 4 dispersions) + ELBO' % NUM_UPDATE_ITER,
                   'vi_updates_per_step': NUM_UPDATE_ITER,
                   'l2': 'inputs larger than L2: per-step working set %.1f GB vs 126 MB' % (G * N1 * S * 8 * 11 / 1e9)},
        'em_iterations_per_s': G_total / (ms_per_step * 1e-3),
        'e_step_ms_per_step': e_ms, 'm_step_ms_per_step': m_ms / args.steps,
        'e_steps_per_s': G_total / (e_ms * 1e-3),
        'mstep': {'rounds_per_step': res['mstep_rounds'] / args.steps, 'evaluations_per_step': res['mstep_evals'] / args.steps},
        'clocks': res['clocks'], 'gpu_launches': int(launches),
        'e2e': {'value': evals_fit / wall, 'unit': UNIT, 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
                'ms_per_step': wall * 1e3,
                'step': 'the complete fit of the grid from host arrays: construction, %d EM iterations, MAP decode, '
                        'results on the host, selection' % NUM_EM_ITER,
                'statistic': 'median of %d fits' % n_full, 'wall_s_all': [round(f[0], 4) for f in full]},
        'full_fit': {'wall_s': wall, 'em_iterations_per_s': G_total * NUM_EM_ITER / wall,
                     'e_step_share': prof['e_step'] / wall, 'm_step_share': prof['m_step'] / wall,
                     'phases_s': dict((k, v) for k, v in prof.items() if isinstance(v, float)),
                     'mstep_seconds': prof.get('mstep_seconds'), 'selected_init': best, 'failed_inits': n_err,
                     'inits': G_total, 'num_em_iter': NUM_EM_ITER, 'num_update_iter': NUM_UPDATE_ITER,
                     'sub_batches': int(prof.get('sub_batches', 1)),
                     'note': 'sub_batches > 1: the grid runs as interleaved sub-batches on their own host threads and '
                             'streams, whose phases overlap; phases_s and the shares are the largest of a sub-batch'},
        'roofline': roofline, 'kernels': kernels,
        'elbo_checksum': res['elbo_checksum'],
    }
    if rep_res:
        line['replicas'] = {'scaling': 'weak', 'fits_per_gpu': G_total, 'ms_per_step': rep_res['ms'] / max(1, min(args.steps, 3)),
                            'value': evals_step * world / (rep_res['ms'] / max(1, min(args.steps, 3)) * 1e-3)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line['cpu_baseline'] = cpu_baseline(name, args.cpu_sample_segments)
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# reference arm: the UNMODIFIED reference (its own cn_model.py driving its own compiled bpmodel.pyx, both under
# oracle/_ref) on the host cores.  Nothing of remixt_b200 is imported here.
# ---------------------------------------------------------------------------------------------
_W = {}


def _ref_worker_init(name, nseg):
    import contextlib
    import io
    import warnings
    from oracle import refenv
    kind = 'reference'
    rcn = None
    import glob
    if refenv.available() and glob.glob(os.path.join(refenv.BIN_PKG, 'bpmodel*.so')):
        refenv.install()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            import remixt.cn_model as rcn
    cn_max = 6
    if name in DERIVED:         # the base genome and grid with the derived model configuration (no remixt_b200 import here)
        base, cfg = DERIVED[name]
        exp, inits = load_workload(base)
        cn_max = cfg.get('max_copy_number', 6)
        depth = min(2. * ip['h_normal'] + (cn_max + 0.25) * ip['h_tumour'] for ip in inits)      # analysis/pipeline.py:47,85
        inits = [dict(ip, max_depth=depth) for ip in inits]
    else:
        exp, inits = load_workload(name)
    _W.update(kind=kind, rcn=rcn, sub=subsample_experiment(exp, nseg), inits=inits, cn_max=cn_max,
              quiet=lambda: contextlib.redirect_stdout(io.StringIO()))


def _ref_worker_build(idx):
    ip = _W['inits'][idx % len(_W['inits'])]
    sub = _W['sub']
    h = np.array([ip['h_normal'], ip['h_tumour'] * ip['mix_frac'], ip['h_tumour'] * (1. - ip['mix_frac'])])
    kw = dict(max_copy_number=_W['cn_max'], normal_contamination=True, divergence_weight=ip['divergence_weight'],
              min_segment_length=10000, min_proportion_genotyped=0.01, max_depth=ip['max_depth'],
              normal_copies=np.array([[1, 1]] * sub.l.shape[0]))
    if _W['rcn'] is not None:
        bm = _W['rcn'].BreakpointModel(sub.x, sub.l, sub.adjacencies, sub.breakpoints, **kw)
    else:           # no compiled reference on this box: the oracle port under this repo's host mirror
        from oracle import bpmodel_np
        from remixt_b200 import cn_model
        _W['kind'] = 'port'
        bm = cn_model.BreakpointModel(sub.x, sub.l, sub.adjacencies, sub.breakpoints, model_class=bpmodel_np.RemixtModel, **kw)
    bm.num_em_iter = 0
    np.random.seed(1234)
    with _W['quiet']():
        bm.fit(h)           # builds the RemixtModel, masks, first ELBO (cn_model.py:354-407)
    _W['bm'] = bm
    return (bm.model.num_segments, bm.model.num_cn_states, _W['kind'])


def _ref_worker_step(idx):
    """One EM iteration, the loop body of BreakpointModel.fit (cn_model.py:409-419), through the reference's own methods."""
    if 'bm' not in _W:          # pool.map does not pin one build task to every worker
        _ref_worker_build(idx)
    bm = _W['bm']
    t0 = time.perf_counter()
    with _W['quiet']():
        for _ in range(NUM_UPDATE_ITER):
            bm.variational_update()
        t1 = time.perf_counter()
        bm.em_update_h()
        bm.em_update_params()
        elbo = bm.model.calculate_elbo()
    return (elbo, t1 - t0, time.perf_counter() - t1)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_run(name, nseg, steps, warmup, cores=None):
    import multiprocessing as mp
    cores = cores or host_cores()
    # the reference holds 3 (N-1,S,S) arrays per model: ~0.23 MB per segment at S=97
    ctx = mp.get_context('spawn')
    with ctx.Pool(cores, initializer=_ref_worker_init, initargs=(name, nseg)) as pool:
        info = pool.map(_ref_worker_build, range(cores), chunksize=1)
        N1, S, kind = info[0]
        for _ in range(warmup):
            pool.map(_ref_worker_step, range(cores), chunksize=1)
        t0 = time.perf_counter()
        parts = []
        for _ in range(steps):
            parts += pool.map(_ref_worker_step, range(cores), chunksize=1)
        dt = (time.perf_counter() - t0) / max(steps, 1)
    evals = float(N1) * S * NUM_UPDATE_ITER * cores
    e_s = float(np.mean([p[1] for p in parts]))
    m_s = float(np.mean([p[2] for p in parts]))
    return {'value': evals / dt, 'unit': UNIT, 'cores': cores, 'kind': kind, 'ms_per_step': dt * 1e3,
            'e_step_s_per_fit': e_s, 'm_step_s_per_fit': m_s,
            # one fit on one core: the reference is single threaded (its workflow fans inits out over processes)
            'single_core': {'value': float(N1) * S * NUM_UPDATE_ITER / (e_s + m_s), 'unit': UNIT,
                            's_per_em_iteration': e_s + m_s},
            'sample': '%d consecutive segments of %s holding the most breakpoints (%d model segments, S=%d), one fit per '
                      'core on %d cores, step = one EM iteration (%d VI rounds + M-step + ELBO) by the reference\'s own '
                      'cn_model.py on its compiled bpmodel' % (nseg, name, N1, S, cores, NUM_UPDATE_ITER)}


def cpu_baseline(name, nseg):
    """Bounded CPU sample inside the default run: ONE step of the reference on every host core."""
    try:
        return reference_run(name, nseg, steps=1, warmup=0)
    except Exception as e:
        return {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'unavailable', 'sample': repr(e)}


def reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    name = pick_workload(args.workload)
    r = reference_run(name, args.cpu_sample_segments, args.steps, args.warmup)
    exp, inits = load_workload(name)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': r['ms_per_step'], 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic (remixt.simulations genome, fixture exp_%s.npz)' % name,
        'config': {'workload': name, 'vi_updates_per_step': NUM_UPDATE_ITER, 'sample_segments': args.cpu_sample_segments,
                   'step': 'one EM iteration: %d VI rounds + M-step (h, 4 dispersions) + ELBO' % NUM_UPDATE_ITER},
        'cpu_baseline': dict((k, r[k]) for k in ('value', 'unit', 'cores', 'kind', 'sample', 'single_core',
                                                 'e_step_s_per_fit', 'm_step_s_per_fit')),
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        # the full config-2 grid at this rate (stated extrapolation: per-(segment,state) cost of the reference is
        # flat in N; the M-step's 200-row samples are 42 rows on this sample, which flatters the CPU)
        'full_fit_extrapolated': {'wall_s': float(exp.x.shape[0]) * 1.0035 * 97 * NUM_UPDATE_ITER * NUM_EM_ITER * len(inits) / r['value'],
                                  'inits': len(inits), 'cores': r['cores']},
        'gpu_launches': 0,
    }
    print(json.dumps(line))


def main():
    args = parse_args()
    if args.workload is None:
        args.workload = pick_workload(None)
    if args.impl == 'reference':
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == '__main__':
    main()
