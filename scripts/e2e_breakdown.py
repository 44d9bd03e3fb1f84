"""Where the end-to-end (host arrays in, results out) step of bench.py spends its time."""
import sys
import time

import numpy as np

sys.path.insert(0, '.')
sys.argv = [sys.argv[0]]
import bench  # noqa: E402
from remixt_b200 import bpmodel, pipeline  # noqa: E402

exp, inits = bench.load_workload(bench.pick_workload(None))
G = len(inits)


def tick(label, t0):
    t1 = time.perf_counter()
    print('%-28s %8.1f ms' % (label, (t1 - t0) * 1e3))
    return t1


for rep in range(2):
    print('--- rep', rep)
    t = time.perf_counter()
    drivers = [pipeline.build_model(exp, ip, bench.CONFIG) for ip in inits]
    t = tick('build_model x%d' % G, t)
    h = np.array([pipeline.init_h(ip) for ip in inits])
    dw = np.array([ip['divergence_weight'] for ip in inits], dtype=float)
    args = drivers[0].model_args(h, compact=True) + (dw,)
    t = tick('model_args', t)
    batch = bpmodel.RemixtBatch(*args, device=0)
    t = tick('RemixtBatch()', t)
    batch.total_likelihood_mask = drivers[0]._total_likelihood_mask.astype(int)
    batch.allele_likelihood_mask = drivers[0]._allele_likelihood_mask.astype(int)
    batch.transition_model = drivers[0].transition_model
    t = tick('masks', t)
    batch.variational_update(5)
    t = tick('variational_update(5)', t)
    elbo = batch.calculate_elbo()
    t = tick('elbo', t)
    cn = np.zeros((G, batch.num_segments, batch.num_clones, 2), dtype=np.int64)
    batch.infer_cn(cn)
    t = tick('infer_cn', t)
    batch.infer_cn(cn)
    t = tick('infer_cn (output pages resident)', t)
    out = (batch.p_outlier_total, batch.p_outlier_allele, batch.h)
    t = tick('read q', t)
    del batch
    t = tick('destroy', t)
