"""A short E-step on the bench workload, for ncu captures:  python scripts/prof_estep.py [fits] [rounds] [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

fits = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 2
name = sys.argv[3] if len(sys.argv) > 3 else 'config2'
exp, inits = bench.load_workload(name)
batch, drivers, lock, _ = bench.grid_batch(exp, inits[:fits], 0)
batch.variational_update(rounds)
print('elbo', float(batch.calculate_elbo().sum()))
