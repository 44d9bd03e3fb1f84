"""BASELINE config 5: a batch of independent simulated genomes (benchmark/sim_defs.yaml-style parameters, seeds
999 ...), every genome fitted completely -- init grid from this repo's pipeline.init, all initialisations as one
batch, 5 EM x 5 VI, MAP decode, selection -- and the genomes dealt round-robin to the ranks (no data-path collective:
SURVEY.md section 8e row 2, "replicas").

    python scripts/config5_batch.py [--genomes 8] [--segments 5000]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/config5_batch.py --genomes 64

Rank 0 prints one JSON line: genomes / fits / wall time (max over ranks) / aggregate segment-state evals/s."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    args = sys.argv[1:]
    opt = lambda f, d: int(args[args.index(f) + 1]) if f in args else d
    genomes, nseg = opt('--genomes', 8), opt('--segments', 5000)
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group('nccl')
    from remixt_b200 import pipeline, simulate
    cfg = {'max_copy_number': 6, 'num_em_iter': 5, 'num_update_iter': 5}
    mine = list(range(rank, genomes, world))
    # inputs first (host work of the reference's own upstream stages: simulation, minor modes, init grid)
    t0 = time.perf_counter()
    jobs = []
    for i in mine:
        exp = simulate.simulate_experiment(dict(N=nseg, random_seed=999 + i))
        jobs.append((i, exp, pipeline.init(exp, cfg)))
    t_prep = time.perf_counter() - t0
    if jobs:      # warm the library and the memory pools on this rank's device
        pipeline.fit_grid_batched(jobs[0][1], dict(list(jobs[0][2].items())[:2]), cfg, device=local)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    evals, fits, picked = 0., 0, {}
    for i, exp, grid in jobs:
        res, best, errors = pipeline.fit_grid_batched(exp, grid, cfg, device=local)
        n1 = next(iter(res.values()))['cn'].shape[0] if res else 0
        s_states = 97
        evals += float(n1) * s_states * cfg['num_em_iter'] * cfg['num_update_iter'] * len(res)
        fits += len(res)
        picked[i] = best
    wall = time.perf_counter() - t0
    rec = dict(rank=rank, genomes=len(jobs), fits=fits, wall_s=wall, prep_s=t_prep, evals=evals, picked=picked)
    if dist is not None:
        out = [None] * world
        dist.all_gather_object(out, rec)
    else:
        out = [rec]
    if rank == 0:
        wall_max = max(r['wall_s'] for r in out)
        print(json.dumps({
            'workload': 'config5', 'genomes': sum(r['genomes'] for r in out), 'segments_per_genome': nseg, 'n_gpus': world,
            'fits': sum(r['fits'] for r in out), 'wall_s': wall_max, 'genomes_per_s': sum(r['genomes'] for r in out) / wall_max,
            'segment_state_evals_per_s': sum(r['evals'] for r in out) / wall_max,
            'host_prep_s_per_rank': max(r['prep_s'] for r in out),
            'what': 'complete fits (init grid as one batch, 5 EM x 5 VI, M-steps, MAP decode, selection) of independent '
                    'simulated genomes dealt round-robin to the ranks; wall time = slowest rank',
            'selected_init_per_genome': {str(k): v for r in out for k, v in r['picked'].items()}}))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
