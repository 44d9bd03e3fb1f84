"""Device time of every E-step kernel class per launch:  python scripts/time_estep.py [workload] [fits]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'config2'
fits = int(sys.argv[2]) if len(sys.argv) > 2 else 8
exp, inits = bench.load_workload(name)
print(name, 'inits', len(inits))
batch, drivers, lock, _ = bench.grid_batch(exp, inits[:fits], 0)
print('N1', batch.num_segments, 'S', batch.num_cn_states, 'K', batch.num_breakpoints)
batch.variational_update(2)
batch.kernel_timing(True)
batch.event_record(0)
batch.variational_update(5)
elbo = batch.calculate_elbo()
batch.event_record(1)
batch.synchronize()
print('5 rounds + elbo: %.2f ms' % batch.event_elapsed(0, batch, 1))
for k, (ms, n) in batch.kernel_timing(False).items():
    if n:
        print('  %-10s %3d launches  %.3f ms each' % (k, n, ms / n))
print('elbo', float(elbo.sum()))
