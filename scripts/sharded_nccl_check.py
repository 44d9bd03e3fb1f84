"""One chromosome shard per GPU over NCCL (run under torchrun, N = 2, 4 or 8):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/sharded_nccl_check.py [workload]

Every rank builds the same host driver, owns one shard on its own GPU, runs two rounds of the five coordinate
updates with the (K, B) allreduce inside update_p_breakpoint, and compares ELBO / p_breakpoint / posteriors /
MAP copy number with an unsharded model on its GPU.  Rank 0 prints one JSON line with the timings."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from remixt_b200 import sharded
    from tests import util
    args = [a for a in sys.argv[1:] if not a.startswith('--')]
    in_lib = '--host-staged' not in sys.argv
    name = args[0] if args else 'genome5k'
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl')
    comm = (sharded.NcclLibComm if in_lib else sharded.TorchDistComm)(device=local)
    exp = util.load_experiment(name)
    ip = util.init_params(name, 1)
    ref = util.make_driver(exp, ip, model_kwargs=dict(device=local)).model
    shd = util.make_driver(exp, ip, model_class=sharded.ShardedRemixtModel,
                           model_kwargs=dict(comm=comm, local_kwargs=dict(shard=True, device=local))).model
    t_ref = t_shd = 0.
    for r in range(3):
        t0 = time.perf_counter()
        if in_lib:
            ref.variational_update(1)
        else:
            for step in util.STEPS:
                getattr(ref, step)()
        e_ref = ref.calculate_elbo()
        t1 = time.perf_counter()
        dist.barrier()
        t2 = time.perf_counter()
        if in_lib:
            shd.variational_update(1)
        else:
            for step in util.STEPS:
                getattr(shd, step)()
        e_shd = shd.calculate_elbo()
        t3 = time.perf_counter()
        if r > 0:
            t_ref += t1 - t0
            t_shd += t3 - t2
        assert abs(e_ref - e_shd) <= 1e-9 * abs(e_ref), (e_ref, e_shd)
    util.assert_close(shd.p_breakpoint, ref.p_breakpoint, rtol=1e-9, atol=1e-280, what='p_breakpoint')
    util.assert_close(shd.posterior_marginals, ref.posterior_marginals, rtol=1e-9, atol=1e-200, what='posterior')
    # the M-step objective and its gradient on a sample that leaves some shards without rows
    sample = np.zeros(ref.num_segments, dtype=np.int64)
    sample[:ref.num_segments // (2 * comm.size):7] = 1
    ell = [ref.calculate_expected_log_likelihood(sample), shd.calculate_expected_log_likelihood(sample)]
    assert abs(ell[0] - ell[1]) <= 1e-10 * abs(ell[0]), ell
    g1, g2 = np.zeros(3), np.zeros(3)
    ref.calculate_expected_log_likelihood_partial_h(sample, g1)
    shd.calculate_expected_log_likelihood_partial_h(sample, g2)
    util.assert_close(g2, g1, rtol=1e-9, what='partial_h')
    assert abs(ref.hmm_log_norm_const - shd.hmm_log_norm_const) <= 1e-10 * abs(ref.hmm_log_norm_const)
    c1 = np.zeros((ref.num_segments, 3, 2), dtype=np.int64)
    c2 = np.zeros_like(c1)
    ref.infer_cn(c1)
    shd.infer_cn(c2)
    assert np.array_equal(c1, c2)
    if comm.rank == 0:
        print(json.dumps({'check': 'sharded_nccl', 'workload': name, 'ranks': comm.size, 'segments': int(ref.num_segments),
                          'breakpoints': int(ref.num_breakpoints), 'elbo': e_shd, 'elbo_unsharded': e_ref,
                          'round_ms_unsharded_1gpu': 1e3 * t_ref / 2, 'round_ms_sharded': 1e3 * t_shd / 2,
                          'shard_segments': [b - a for a, b in shd.bounds], 'collectives': 'in-library ncclAllReduce on the model stream' if in_lib else 'host-staged torch.distributed',
                          'ell_check': ell, 'ok': True}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
