"""Cost of one lock-step objective round (48 fits, 200-row samples) by parameter: wall per call and device time of
the objective kernel."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

exp, inits = bench.load_workload('config2')
batch, drivers, lock, _ = bench.grid_batch(exp, inits, 0)
batch.variational_update(2)
G, N = batch.num_fits, batch.num_segments
rng = np.random.RandomState(0)
rows = np.sort(np.array([rng.choice(N, 200, replace=False) for _ in range(G)]), axis=1).astype(np.int32)
batch.mstep_stage_rows(rows, np.full(G, 200))
for name, c in (('negbin_r_0', 700.), ('betabin_M_0', 900.), (None, None)):
    vals = np.full((G, 1), c) if name else None
    for _ in range(5):
        batch.mstep_evaluate(name=name, values=vals)
    batch.kernel_timing(True)
    t0 = time.perf_counter()
    n = 200
    for k in range(n):
        if name:
            vals[:, 0] = c * (1 + 1e-5 * k)
        batch.mstep_evaluate(name=name, values=vals)
    dt = (time.perf_counter() - t0) / n
    kt = batch.kernel_timing(False)
    print('%-12s wall %.1f us per round; objective kernel %.1f us per launch (%d launches)' % (
        name or 'all terms', dt * 1e6, kt['ell'][0] / max(kt['ell'][1], 1) * 1e3, kt['ell'][1]))
h = np.array(batch.h)
for _ in range(3):
    batch.mstep_evaluate(h=h, grad=True)
t0 = time.perf_counter()
for k in range(100):
    batch.mstep_evaluate(h=h * (1 + 1e-6 * k), grad=True)
print('h + gradient wall %.1f us per round' % ((time.perf_counter() - t0) / 100 * 1e6))
# every segment (the acceptance tests of the M-step): table planes of the changed parameter + their sums + row values
batch.mstep_stage_sample(None)
for name, c in (('negbin_r_0', 700.), ('betabin_M_0', 900.), ('h', None)):
    for rep in range(2):
        batch.kernel_timing(True)
        t0 = time.perf_counter()
        n = 20
        for k in range(n):
            if name == 'h':
                batch.mstep_evaluate(h=h * (1 + 1e-6 * (k + 1)))
            else:
                batch.mstep_evaluate(name=name, values=np.full((G, 1), c * (1 + 1e-5 * (k + 1 + 100 * rep))))
        dt = (time.perf_counter() - t0) / n
        kt = batch.kernel_timing(False)
    print('full data %-12s wall %.2f ms per evaluation; table %.2f ms, sums %.2f ms per launch' % (
        name, dt * 1e3, kt['lltable'][0] / max(kt['lltable'][1], 1), kt['qsums'][0] / max(kt['qsums'][1], 1)))
