"""Chromosome-sharded fit (remixt_b200/sharded.py) on CPU: the sharding plan, and the sharded model driven with
the numpy oracle as the per-shard model -- in one process (LocalComm, 3 shards) and over a world_size-2 gloo
group (TorchDistComm) -- against the same oracle unsharded.  The reductions under test are host code; the CUDA
shards are covered by tests/test_sharded_gpu.py."""
import os
import socket

import numpy as np
import pytest

from tests import util


def test_plan_shards():
    from remixt_b200.sharded import plan_shards
    tel = np.zeros(100, dtype=int)
    tel[[9, 29, 49, 79, 99]] = 1
    for k in range(1, 6):
        b = plan_shards(tel, k)
        assert len(b) == k and b[0][0] == 0 and b[-1][1] == 100
        assert all(b[i][1] == b[i + 1][0] for i in range(k - 1))
        assert all(tel[e - 1] == 1 for _, e in b)            # groups end at chromosome ends
    assert plan_shards(tel, 2) == [(0, 50), (50, 100)]
    with pytest.raises(ValueError):
        plan_shards(tel, 6)
    bidx = -np.ones(100, dtype=int)
    bidx[49] = 0
    with pytest.raises(ValueError):
        plan_shards(tel, 2, bidx)


def _drivers(num_shards, comm=None):
    from oracle import bpmodel_np
    from remixt_b200 import sharded
    exp = util.load_experiment('small')
    ip = util.init_params('small', 1)
    ref = util.make_driver(exp, ip, model_class=bpmodel_np.RemixtModel)
    shd = util.make_driver(exp, ip, model_class=sharded.ShardedRemixtModel,
                           model_kwargs=dict(comm=comm, num_shards=num_shards, model_class=bpmodel_np.RemixtModel))
    return ref, shd


def _run_and_compare(ref, shd, rtol=1e-10):
    a, b = ref.model, shd.model
    assert abs(a.calculate_elbo() - b.calculate_elbo()) <= rtol * abs(a.calculate_elbo())
    for _ in range(2):
        for step in util.STEPS:
            getattr(a, step)()
            getattr(b, step)()
        util.assert_close(b.p_breakpoint, np.asarray(a.p_breakpoint), rtol=rtol, atol=1e-280, what='p_breakpoint')
        util.assert_close(b.posterior_marginals, np.asarray(a.posterior_marginals), rtol=rtol, atol=1e-200, what='post')
        for f in ('calculate_variational_energy', 'calculate_variational_entropy', 'calculate_elbo'):
            util.assert_close(getattr(b, f)(), getattr(a, f)(), rtol=rtol, what=f)
        util.assert_close(b.hmm_log_norm_const, a.hmm_log_norm_const, rtol=rtol, what='logZ')
    for name in ('p_outlier_total', 'p_outlier_allele', 'p_allele_swap'):
        util.assert_close(getattr(b, name), np.asarray(getattr(a, name)), rtol=rtol, atol=1e-280, what=name)
    rs = np.random.RandomState(5)
    sample = (rs.rand(a.num_segments) < 0.2).astype(int)
    util.assert_close(b.calculate_expected_log_likelihood(sample), a.calculate_expected_log_likelihood(sample),
                      rtol=rtol, what='E[LL]')
    g1, g2 = np.zeros(3), np.zeros(3)
    a.calculate_expected_log_likelihood_partial_h(sample, g1)
    b.calculate_expected_log_likelihood_partial_h(sample, g2)
    util.assert_close(g2, g1, rtol=rtol, what='dE[LL]/dh')
    # parameter writes reach every shard
    for m in (a, b):
        m.betabin_M_0 = 321.
        m.h = np.asarray(m.h) * 1.01
    util.assert_close(b.calculate_expected_log_likelihood(sample), a.calculate_expected_log_likelihood(sample),
                      rtol=rtol, what='E[LL] after setattr')
    c1 = np.zeros((a.num_segments, 3, 2), dtype=np.int64)
    c2 = np.zeros_like(c1)
    a.infer_cn(c1)
    b.infer_cn(c2)
    assert np.array_equal(c1, c2)
    return b.calculate_elbo()


def test_three_shards_in_process():
    ref, shd = _drivers(3)
    assert len(shd.model.shards) == 3
    _run_and_compare(ref, shd)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from remixt_b200 import sharded
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    ref, shd = _drivers(2, comm=sharded.TorchDistComm())
    assert len(shd.model.shards) == 1
    elbo = _run_and_compare(ref, shd)
    np.save(os.path.join(out_dir, 'r%d.npy' % rank), np.array([elbo]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_shards_over_gloo(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert np.array_equal(np.load(str(tmp_path / 'r0.npy')), np.load(str(tmp_path / 'r1.npy')))
