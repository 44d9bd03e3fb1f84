"""Benchmark of the ReMixT variational-inference hot path on B200 (contract: see DESIGN.md, Measurement).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME] [--fits G]

Workload: a simulated whole genome (tests/golden/exp_<workload>.npz, produced by the reference's own
simulator, oracle/make_inputs.py) fitted from every initialisation of the reference's (h, ploidy) grid
(stored in the same fixture by oracle/make_golden.py init).  One *step* is one E-step of the EM loop for
the whole grid: the likelihood table for the current (h, dispersions), `num_update_iter` (5,
remixt/defaults.py:157) rounds of the five coordinate updates (cn_model.py:444-460) and the ELBO
(cn_model.py:418), for all G fits.

    value = N1 * S * num_update_iter * G_total / step_time      [segment-state evals / s]

`value` has all inputs resident in HBM; `e2e` builds every model from HOST arrays through the public
RemixtModel constructor (H2D inside the timed region), runs the same E-step and reads ELBO + outlier
posteriors + posterior marginals back to the host (D2H).

N > 1 (torchrun): the init grid is sharded over ranks (`ids[rank::world]`, no data-path collective;
scaling = weak: every rank fits a full G-sized grid slice of its own).
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
    return ap.parse_args()


def pick_workload(name):
    if name:
        return name
    for cand in ('config2', 'genome5k'):
        if os.path.exists(os.path.join(GOLDEN, 'exp_%s.npz' % cand)):
            return cand
    raise SystemExit('no workload fixture under tests/golden')


def load_workload(name):
    from remixt_b200 import pipeline
    path = os.path.join(GOLDEN, 'exp_%s.npz' % name)
    d = np.load(path, allow_pickle=False)
    exp = pipeline.ArrayExperiment.from_npz(path)
    exp._name = name
    inits = []
    for i in range(len(d['init_h_normal'])):
        inits.append(dict(mode_idx=int(d['init_mode_idx'][i]), h_normal=float(d['init_h_normal'][i]),
                          h_tumour=float(d['init_h_tumour'][i]), mix_frac=float(d['init_mix_frac'][i]),
                          divergence_weight=float(d['init_divergence_weight'][i]),
                          max_depth=float(d['init_max_depth'][i])))
    return exp, inits


def subsample_experiment(exp, nseg):
    """The first `nseg` segments of the workload with the adjacencies / breakpoints that lie wholly inside."""
    from remixt_b200 import pipeline
    nseg = min(nseg, exp.x.shape[0])
    adj = np.array([(a, b) for a, b in sorted(exp.adjacencies) if a < nseg and b < nseg], dtype=np.int64).reshape(-1, 2)
    ids, bps = [], []
    for bp_id, bp in exp.breakpoints.items():
        (n1, s1), (n2, s2) = tuple(bp)
        if n1 < nseg and n2 < nseg and not (n1 == 0 and s1 == 0) and not (n2 == 0 and s2 == 0):
            ids.append(bp_id)
            bps.append((n1, s1, n2, s2))
    return pipeline.ArrayExperiment(exp.x[:nseg], exp.l[:nseg], adj, np.array(bps, dtype=np.int64).reshape(-1, 4),
                                    breakpoint_ids=ids, chromosome=exp.segment_chromosome_id[:nseg])


CONFIG = {'max_copy_number': 6, 'num_em_iter': 1, 'num_update_iter': NUM_UPDATE_ITER}


def make_driver(exp, ip, model_class=None, device=None):
    from remixt_b200 import pipeline
    kw = {}
    if model_class is not None:
        kw['model_class'] = model_class
    if device is not None:
        kw['model_kwargs'] = {'device': device}
    bm = pipeline.build_model(exp, ip, CONFIG, **kw)
    return bm


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


def run_threads(fns):
    if len(fns) == 1:
        fns[0]()
        return
    errs = []

    def wrap(f):
        try:
            f()
        except Exception as e:   # surfaced after join
            errs.append(e)
    ts = [threading.Thread(target=wrap, args=(f,)) for f in fns]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]


def build_batch(exp, inits, device):
    """One RemixtBatch for the given initialisations, built from HOST arrays through the public constructor
    (the H2D copies happen inside), with the per-fit likelihood masks and transition model applied."""
    from remixt_b200 import bpmodel, pipeline
    drivers = [pipeline.build_model(exp, ip, CONFIG) for ip in inits]
    h = np.array([pipeline.init_h(ip) for ip in inits])
    dw = np.array([ip['divergence_weight'] for ip in inits], dtype=float)
    args = drivers[0].model_args(h, compact=True) + (dw,)
    batch = bpmodel.RemixtBatch(*args, device=device)
    same_masks = all(np.array_equal(d._total_likelihood_mask, drivers[0]._total_likelihood_mask) and
                     np.array_equal(d._allele_likelihood_mask, drivers[0]._allele_likelihood_mask) for d in drivers)
    if same_masks:       # max_depth is shared by the grid (analysis/pipeline.py:90-96): one upload for all fits
        batch.total_likelihood_mask = drivers[0]._total_likelihood_mask.astype(int)
        batch.allele_likelihood_mask = drivers[0]._allele_likelihood_mask.astype(int)
        batch.transition_model = drivers[0].transition_model
        for g, d in enumerate(drivers):
            d.model = batch.fit(g)
    else:
        for g, d in enumerate(drivers):
            d.attach(batch.fit(g))
    h2d = sum(int(np.asarray(a).nbytes) for a in args if isinstance(a, np.ndarray)) + 2 * 8 * batch.num_segments * len(inits)
    h2d += sum(int(a.table.nbytes + a.normal_copies.nbytes) for a in args if isinstance(a, bpmodel.CompactCnStates))
    return batch, drivers, h2d


LLTABLE_CONVENTION_OPS = 900.
NCU_DRAM_BYTES_PER_UNIT = {'chain': 29.0, 'lltable': 48.8, 'round': 89.1}


# algorithmic work per (segment, state) of each kernel class: (bound, bytes, fp64 ops)  -- DESIGN.md, Kernels
def kernel_models(N1, S, NC, G, R):
    ns = float(N1) * S * G
    return {
        'chain': ('fp64', 32. * ns, 2. * 2. * S * S * (N1 - NC) * G),
        # 200 executed FP64-pipe instructions per (n,s) (ncu: dadd + dmul + dfma thread instructions of k_lltable2,
        # profiles/r01_summary.md), each counted as one FMA = 2 flop; SURVEY.md 8d's convention for the reference-minimal
        # evaluation (12 lgamma + 4 log = 900 op-equivalents) is reported next to it (LLTABLE_CONVENTION_OPS)
        'lltable': ('fp64', 48. * ns, 400. * ns),
        'frame': ('hbm', 64. * ns, 14. * ns),
        'qsums': ('hbm', 56. * ns, 12. * ns),
        'post': ('hbm', 24. * ns, 3. * ns),
        'hist': ('fp64', 0., 2. * 8. * S * S * R * G),
        'elbo_rows': ('hbm', 16. * ns, 4. * ns),
        'round': ('hbm', 88. * ns, 29. * ns),
    }


def gpu_arm(args):
    rank, world, local, dist = dist_setup(args.gpus)
    from remixt_b200 import _lib
    if _lib.device_count() <= 0:
        raise SystemExit('bench.py: no CUDA device; remixt_b200 has no CPU fallback')
    name = pick_workload(args.workload)
    exp, inits = load_workload(name)
    G = args.fits if args.fits > 0 else len(inits)
    # weak scaling: every rank fits G initialisations (its own rotation of the grid)
    my_inits = [inits[(rank * G + g) % len(inits)] for g in range(G)]
    peaks, peak_src = measured_peaks()

    # ---------------- resident arm: batch built once, inputs live in HBM
    batch, drivers, h2d = build_batch(exp, my_inits, local)
    N1, S = batch.num_segments, batch.num_cn_states
    NC = int(np.sum(np.asarray(batch.is_telomere)[:-1] > 0)) + 1
    R = int(np.sum(np.asarray(batch.breakpoint_idx)[:-1] >= 0))
    state = {}

    h_cur = np.array(batch.h)

    def estep():
        # every EM iteration of the reference starts from a new h / dispersions (cn_model.py:411-416), so the
        # likelihood table of every fit is rebuilt once per step: writing h marks it stale
        batch.h = h_cur
        batch.variational_update(NUM_UPDATE_ITER)
        state['elbo'] = batch.calculate_elbo()

    def barrier():
        batch.synchronize()
        if dist is not None:
            dist.barrier()
            import torch
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        estep()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    batch.kernel_timing(True)
    l0 = _lib.global_launch_count()
    batch.event_record(0)
    for _ in range(args.steps):
        estep()
    batch.event_record(1)
    barrier()
    ms = batch.event_elapsed(0, batch, 1)
    launches = _lib.global_launch_count() - l0
    ktime = batch.kernel_timing(False)
    clocks = sampler.stop()

    # ---------------- e2e arm: the batch rebuilt from host arrays every step, results read back to the host
    def e2e_step():
        b2, d2, _ = build_batch(exp, my_inits, local)
        b2.variational_update(NUM_UPDATE_ITER)
        elbo = b2.calculate_elbo()
        cn = np.zeros((G, b2.num_segments, b2.num_clones, 2), dtype=np.int64)
        b2.infer_cn(cn)
        out = (elbo, cn, b2.p_outlier_total, b2.p_outlier_allele, b2.h)
        b2.close()      # device blocks back to the pool now: the next step's batch reuses them
        return sum(int(np.asarray(o).nbytes) for o in out)

    del state['elbo']
    d2h = e2e_step()
    e2e_steps = max(1, min(args.steps, 5))
    barrier()
    e2e_times = []
    for _ in range(e2e_steps):
        t0 = time.perf_counter()
        e2e_step()
        e2e_times.append(time.perf_counter() - t0)
    # the median step: a host hiccup (page faults of a fresh box, another tenant on the host cores) lands on one step
    # of a handful and would otherwise own the mean; all step times are reported next to it
    e2e_s = float(np.median(e2e_times))

    if dist is not None:
        import torch
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])
        tl = torch.tensor([launches], dtype=torch.int64, device='cuda')
        dist.all_reduce(tl)
        launches = int(tl[0])
    ms_per_step = ms / args.steps
    evals = float(N1) * S * NUM_UPDATE_ITER * G * world
    value = evals / (ms_per_step * 1e-3)

    # ---------------- rooflines from the CUDA-event durations of the timed region
    try:
        fp64_peak = _lib.fp64_peak(local)
    except Exception:
        fp64_peak = None
    hbm_peak = peaks.get('hbm_gbs')
    models = kernel_models(N1, S, NC, G, R)
    total_kernel_ms = sum(v[0] for v in ktime.values())
    kernels = {}
    for k, (tot, n) in ktime.items():
        if n == 0:
            continue
        ent = {'launches': n, 'ms_per_launch': tot / n, 'share_of_kernel_time': tot / total_kernel_ms}
        if k in models:
            bound, nbytes, flops = models[k]
            t_s = tot / n * 1e-3
            ent['bound'] = bound
            ent['gbs'] = nbytes / t_s / 1e9
            ent['tflops'] = flops / t_s / 1e12
            ent['hbm_frac'] = ent['gbs'] / hbm_peak if hbm_peak else None
            ent['fp64_frac'] = ent['tflops'] / fp64_peak if fp64_peak else None
            if k == 'lltable' and fp64_peak:
                ent['fp64_frac_survey_convention_900'] = ent['tflops'] * (LLTABLE_CONVENTION_OPS / 400.) / fp64_peak
        kernels[k] = ent
    dom = max(kernels, key=lambda k: kernels[k]['share_of_kernel_time'] if k in models else -1.)
    de = kernels[dom]
    if de['bound'] == 'hbm':
        roofline = {'kernel': dom, 'bound': 'hbm', 'achieved': de['gbs'], 'peak': hbm_peak, 'unit': 'GB/s',
                    'frac': de['hbm_frac'], 'traffic': None, 'peak_source': 'MEASURED_PEAKS.json (%s)' % peak_src}
    else:
        roofline = {'kernel': dom, 'bound': 'fp64', 'achieved': de['tflops'], 'peak': fp64_peak, 'unit': 'TFLOP/s',
                    'frac': de['fp64_frac'], 'traffic': None,
                    'peak_source': 'FP64 FMA micro-benchmark run in this process (rmx_fp64_peak); not in MEASURED_PEAKS.json',
                    'hbm_achieved_gbs': de['gbs'], 'hbm_peak_gbs': hbm_peak, 'hbm_frac': de['hbm_frac']}
    # DRAM bytes per launch of the dominant kernel from its `ncu --set full` capture (dram__bytes_read.sum +
    # dram__bytes_write.sum of an 8-fit launch, profiles/r01_summary.md), carried per (segment, state) unit
    if dom in NCU_DRAM_BYTES_PER_UNIT:
        roofline['traffic'] = NCU_DRAM_BYTES_PER_UNIT[dom] * float(N1) * S * G
        roofline['traffic_source'] = 'ncu capture of an 8-fit launch, %.1f B per (segment, state)' % NCU_DRAM_BYTES_PER_UNIT[dom]
    roofline['launch_ms'] = de['ms_per_launch']
    roofline['share_of_step'] = de['ms_per_launch'] * de['launches'] / ms
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f64', 'data': 'synthetic (remixt.simulations genome, fixture exp_%s.npz)' % name,
        'config': {'workload': name, 'segments': int(exp.x.shape[0]), 'model_segments': int(N1), 'cn_states': int(S),
                   'breakpoints': int(batch.num_breakpoints), 'chains': NC, 'fits_per_gpu': G,
                   'vi_updates_per_step': NUM_UPDATE_ITER,
                   'l2': 'inputs larger than L2: per-step working set %.1f GB vs 126 MB' % (G * N1 * S * 8 * 11 / 1e9)},
        'em_iterations_per_s': G * world / (ms_per_step * 1e-3),
        'clocks': clocks, 'gpu_launches': int(launches),
        'e2e': {'value': evals / e2e_s, 'unit': UNIT, 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
                'ms_per_step': e2e_s * 1e3, 'statistic': 'median of %d steps' % e2e_steps,
                'ms_all_steps': [round(t * 1e3, 1) for t in e2e_times]},
        'roofline': roofline, 'kernels': kernels,
        'elbo_checksum': float(np.sum(batch.calculate_elbo())),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line['cpu_baseline'] = cpu_baseline(exp, inits, args.cpu_sample_segments, budget_s=20.)
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# reference arm: the compiled UNMODIFIED reference extension (oracle/_ref) on the host cores
# ---------------------------------------------------------------------------------------------
_W = {}


def _ref_worker_init(name, nseg, n_workers):
    from oracle import refenv
    from remixt_b200 import pipeline   # host driver only; the model class below is the reference's
    ref = refenv.load_ref_bpmodel()
    kind = 'reference'
    if ref is None:
        from oracle import bpmodel_np as ref   # the oracle port
        kind = 'port'
    exp, inits = load_workload(name)
    sub = subsample_experiment(exp, nseg)
    _W.update(ref=ref, kind=kind, sub=sub, inits=inits, pipeline=pipeline)


def _ref_worker_build(idx):
    ip = _W['inits'][idx % len(_W['inits'])]
    d = make_driver(_W['sub'], ip, model_class=_W['ref'].RemixtModel)
    d.prepare(_W['pipeline'].init_h(ip))
    _W['driver'] = d
    return (d.model.num_segments, d.model.num_cn_states, _W['kind'])


def _ref_worker_step(idx):
    if 'driver' not in _W:          # pool.map does not pin one build task to every worker
        _ref_worker_build(idx)
    m = _W['driver'].model
    for _ in range(NUM_UPDATE_ITER):
        m.update_p_allele_swap()
        m.update_p_cn()
        m.update_p_breakpoint()
        m.update_p_outlier_total()
        m.update_p_outlier_allele()
    return m.calculate_elbo()


def reference_run(name, nseg, steps, warmup, cores=None):
    import multiprocessing as mp
    if cores is None:
        try:
            cores = len(os.sched_getaffinity(0))
        except AttributeError:
            cores = os.cpu_count() or 1
    # the reference holds 3 (N-1,S,S) arrays per model: ~0.23 MB per segment at S=97
    ctx = mp.get_context('spawn')
    with ctx.Pool(cores, initializer=_ref_worker_init, initargs=(name, nseg, cores)) as pool:
        info = pool.map(_ref_worker_build, range(cores), chunksize=1)
        N1, S, kind = info[0]
        for _ in range(warmup):
            pool.map(_ref_worker_step, range(cores), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_ref_worker_step, range(cores), chunksize=1)
        dt = (time.perf_counter() - t0) / max(steps, 1)
    evals = float(N1) * S * NUM_UPDATE_ITER * cores
    return {'value': evals / dt, 'unit': UNIT, 'cores': cores, 'kind': kind, 'ms_per_step': dt * 1e3,
            'sample': 'first %d segments of %s (%d model segments, S=%d), one fit per core, %d VI rounds + ELBO per step'
                      % (nseg, name, N1, S, NUM_UPDATE_ITER)}


def cpu_baseline(exp, inits, nseg, budget_s=20.):
    """Bounded CPU sample inside the default run: ONE step of the reference on every host core."""
    try:
        return reference_run(exp._name, nseg, steps=1, warmup=0)
    except Exception as e:
        return {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'unavailable', 'sample': repr(e)}


def reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    name = pick_workload(args.workload)
    r = reference_run(name, args.cpu_sample_segments, args.steps, args.warmup)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': r['ms_per_step'], 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic (remixt.simulations genome, fixture exp_%s.npz)' % name,
        'config': {'workload': name, 'vi_updates_per_step': NUM_UPDATE_ITER, 'sample_segments': args.cpu_sample_segments},
        'cpu_baseline': {'value': r['value'], 'unit': UNIT, 'cores': r['cores'], 'kind': r['kind'], 'sample': r['sample']},
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
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
