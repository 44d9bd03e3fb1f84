"""Chromosome-sharded fit on the B200: CUDA shards (C-ABI, RMX_DESC_SHARD, rmx_breakpoint_logq /
rmx_breakpoint_apply_logq) against the unsharded CUDA model and the oracle.  Several shards live on one GPU here
(LocalComm); scripts/sharded_nccl_check.py runs the same comparison with one shard per GPU over NCCL."""
import numpy as np
import pytest

from tests import util
from tests.test_sharded_cpu import _run_and_compare

pytestmark = pytest.mark.gpu


def _pair(name, init_id, num_shards, ref_class=None):
    from remixt_b200 import sharded
    exp = util.load_experiment(name)
    ip = util.init_params(name, init_id)
    ref = util.make_driver(exp, ip, model_class=ref_class)
    shd = util.make_driver(exp, ip, model_class=sharded.ShardedRemixtModel, model_kwargs=dict(num_shards=num_shards))
    return ref, shd


def test_small_three_shards_vs_unsharded():
    ref, shd = _pair('small', 1, 3)
    assert len(shd.model.shards) == 3
    _run_and_compare(ref, shd, rtol=1e-9)


def test_small_three_shards_vs_oracle():
    from oracle import bpmodel_np
    ref, shd = _pair('small', 7, 3, ref_class=bpmodel_np.RemixtModel)
    _run_and_compare(ref, shd, rtol=1e-6)


def test_genome5k_eight_shards():
    """23 chromosomes over 8 shards; breakpoints whose breakends sit on different shards couple them."""
    ref, shd = _pair('genome5k', 1, 8)
    m = shd.model
    owner = np.zeros(m.num_segments, dtype=int)
    for k, (a, b) in enumerate(m.bounds):
        owner[a:b] = k
    split = 0
    for kk in range(m.num_breakpoints):
        rows = np.nonzero(m.breakpoint_idx == kk)[0]
        split += int(len(set(owner[rows])) > 1)
    assert split > 0, 'fixture has no breakpoint spanning two shards'
    _run_and_compare(ref, shd, rtol=1e-9)


def test_shard_flag_required():
    """Without shard=True a model that lacks the last breakpoint is the reference's ValueError."""
    from remixt_b200 import bpmodel
    exp = util.load_experiment('small')
    ip = util.init_params('small', 1)
    drv = util.make_driver(exp, ip)
    args = list(drv.model_args(np.asarray(drv.model.h))) + [ip['divergence_weight']]
    n = int(np.nonzero(np.asarray(args[11]) == args[2] - 1)[0][0])      # up to the last breakpoint's first breakend
    assert n >= 4
    bidx = np.ascontiguousarray(args[11][:n])
    sub = [args[0], n, args[2], args[3], np.ascontiguousarray(args[4][:n]), args[5], args[6],
           np.ascontiguousarray(args[7][:n]), np.ascontiguousarray(args[8][:n]), np.ascontiguousarray(args[9][:n]),
           np.ascontiguousarray(args[10][:n]), bidx, np.ascontiguousarray(args[12][:n]), args[13], args[14]]
    with pytest.raises(ValueError):
        bpmodel.RemixtModel(*sub)
    bpmodel.RemixtModel(*sub, shard=True)


def test_in_library_nccl_single_rank_equals_unsharded():
    """rmx_comm_init with one rank: every in-stream ncclAllReduce path of a shard (update_p_breakpoint, log
    normaliser, ELBO terms with the p_breakpoint entropy counted once, M-step objective / gradient / batched
    evaluation, empty samples) runs and returns what the plain model returns.  The multi-rank case is
    scripts/sharded_nccl_check.py under torchrun (profiles/r02_sharded_nccl_*.log)."""
    import torch  # noqa: F401  (brings libnccl.so.2 into the process)
    from remixt_b200 import bpmodel
    exp = util.load_experiment('small')
    ip = util.init_params('small', 4)
    plain = util.make_driver(exp, ip).model
    shard = util.make_driver(exp, ip, model_kwargs=dict(shard=True)).model
    shard.comm_init(bpmodel.nccl_unique_id(), 0, 1)
    plain.variational_update(2)
    shard.variational_update(2)
    assert shard.calculate_elbo() == plain.calculate_elbo()
    assert shard.hmm_log_norm_const == plain.hmm_log_norm_const
    assert np.array_equal(shard.p_breakpoint, plain.p_breakpoint)
    for step in util.STEPS:
        getattr(plain, step)()
        getattr(shard, step)()
    assert shard.calculate_variational_entropy() == plain.calculate_variational_entropy()
    sample = np.zeros(plain.num_segments, dtype=np.int64)
    sample[::9] = 1
    assert shard.calculate_expected_log_likelihood(sample) == plain.calculate_expected_log_likelihood(sample)
    g1, g2 = np.zeros(3), np.zeros(3)
    plain.calculate_expected_log_likelihood_partial_h(sample, g1)
    shard.calculate_expected_log_likelihood_partial_h(sample, g2)
    assert np.array_equal(g1, g2)
    none = np.zeros(plain.num_segments, dtype=np.int64)
    assert shard.calculate_expected_log_likelihood(none) == 0.
    shard.comm_destroy()
