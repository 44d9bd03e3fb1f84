"""Device time of the chain-message kernel per launch at several fit counts, with and without breakend rows:
    python scripts/time_chain.py [workload]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'config2'
exp, inits = bench.load_workload(name)
for nobrk in (False, True):
    bench.CONFIG['disable_breakpoints'] = nobrk
    if hasattr(exp, '_rmx_segmentation'):
        del exp._rmx_segmentation
    for fits in (8, 48):
        batch, drivers, lock, _ = bench.grid_batch(exp, inits[:fits], 0)
        tel = np.flatnonzero(np.asarray(batch.is_telomere) > 0)
        longest = int(np.max(np.diff(np.concatenate([[-1], tel]))))
        nbrk = int(np.sum(np.asarray(batch.breakpoint_idx) >= 0))
        batch.variational_update(2)
        batch.chain_timing(True)
        batch.variational_update(5)
        ms, n = batch.chain_timing(False)
        print('breakends %4d fits %2d: chain %.3f ms per launch; longest chain %d segments -> %.0f cycles per segment of it; '
              'elbo sum %.10f' % (nbrk, fits, ms / n, longest, ms / n * 1.965e6 / longest, float(batch.calculate_elbo().sum())))
        batch.close()
