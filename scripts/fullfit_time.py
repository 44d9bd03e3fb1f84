"""Wall time of a complete grid fit (all inits, EM loop, MAP decode, selection) by phase.
    python scripts/fullfit_time.py [workload] [--mstep lockstep|threads] [--fits G] [--em E]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    args = sys.argv[1:]
    name = args[0] if args and not args[0].startswith('--') else 'config2'
    mstep = args[args.index('--mstep') + 1] if '--mstep' in args else 'lockstep'
    fits = int(args[args.index('--fits') + 1]) if '--fits' in args else 0
    em = int(args[args.index('--em') + 1]) if '--em' in args else 5
    from remixt_b200 import pipeline
    exp, inits = bench.load_workload(name)
    if fits:
        inits = inits[:fits]
    cfg = {'max_copy_number': 6, 'num_em_iter': em, 'num_update_iter': 5}
    if '--sub-batches' in args:
        cfg['sub_batches'] = int(args[args.index('--sub-batches') + 1])
    for rep in range(3):
        prof = {}
        t0 = time.perf_counter()
        res, best, errors = pipeline.fit_grid_batched(exp, dict(enumerate(inits)), cfg, mstep=mstep, profile=prof)
        wall = time.perf_counter() - t0
        print(json.dumps({'workload': name, 'mstep': mstep, 'rep': rep, 'fits': len(inits), 'wall_s': wall, 'best': best,
                          'errors': len(errors), 'profile': prof}))


if __name__ == '__main__':
    main()
