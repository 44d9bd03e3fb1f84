"""N > 1 path on CPU: the init grid sharded over a world_size-2 gloo process group
(remixt_b200.pipeline.fit_grid_distributed).  Every rank runs the real per-rank grid driver (fit_grid_batched: one
grid object per rank, lock-step M-step); the model class is the compiled reference where built, else the numpy
oracle -- the distribution logic under test is host code and does not depend on which RemixtModel implementation
runs."""
import os
import socket
import sys

import numpy as np
import pytest

from tests import util


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _model_class():
    from oracle import refenv
    ref = refenv.load_ref_bpmodel()
    if ref is not None:
        return ref.RemixtModel
    from oracle import bpmodel_np
    return bpmodel_np.RemixtModel


def _experiment():
    sys.path.insert(0, util.ROOT)
    import bench
    exp = util.load_experiment('small')
    return bench.subsample_experiment(exp, 60)


CFG = {'max_copy_number': 6, 'num_em_iter': 1, 'num_update_iter': 1}
IDS = [0, 5, 11, 17]


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from remixt_b200 import pipeline
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    exp = _experiment()
    inits = dict((i, util.init_params('small', i)) for i in IDS)
    results, stats, best, errors = pipeline.fit_grid_distributed(exp, inits, CFG, dist=dist,
                                                                 model_class=_model_class())
    assert not errors
    np.save(os.path.join(out_dir, 'r%d.npy' % rank),
            np.array([best] + [stats[i]['elbo'] for i in IDS] + sorted(results.keys())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.slow
def test_grid_sharded_over_two_gloo_ranks(tmp_path):
    import torch.multiprocessing as mp
    from remixt_b200 import pipeline
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = np.load(str(tmp_path / 'r0.npy'))
    r1 = np.load(str(tmp_path / 'r1.npy'))
    # both ranks hold the full stats table and agree on the selected initialisation
    assert np.array_equal(r0[:1 + len(IDS)], r1[:1 + len(IDS)])
    # each rank fitted its strided slice only
    assert list(r0[1 + len(IDS):]) == IDS[0::2] and list(r1[1 + len(IDS):]) == IDS[1::2]
    # and the result equals the single-process grid
    exp = _experiment()
    inits = dict((i, util.init_params('small', i)) for i in IDS)
    _, stats, best, _ = pipeline.fit_grid_distributed(exp, inits, CFG, dist=None, model_class=_model_class())
    assert best == int(r0[0])
    assert np.array_equal(np.array([stats[i]['elbo'] for i in IDS]), r0[1:1 + len(IDS)])
