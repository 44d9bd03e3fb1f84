"""Per-phase cycle counts of the chain kernel (a build with -DC2_PROFILE prints them): one 8-fit update on config 2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
exp, inits = bench.load_workload(sys.argv[1] if len(sys.argv) > 1 else 'config2')
batch, drivers, lock, _ = bench.grid_batch(exp, inits[:8], 0)
batch.variational_update(1)
batch.chain_timing(True)
batch.variational_update(1)
ms, n = batch.chain_timing(False)
print('chain %.3f ms per launch' % (ms / n))
batch.close()
