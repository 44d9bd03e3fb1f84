"""The simulator mirror (remixt_b200/simulate.py): count sampler bit-exact against the reference's ExperimentSampler on
the same genome mixture and seed; the fast genome generator produces experiments the model's assumptions hold on."""
import warnings

import numpy as np
import pytest


def test_sample_experiment_equals_reference_sampler():
    from oracle import refenv
    import os
    if not os.path.isdir(os.path.join(refenv.REF_ROOT, 'remixt', 'simulations')):
        pytest.skip('reference simulator not available')
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import remixt.simulations.experiment as rsim
    from remixt_b200 import simulate
    params = dict(N=400, chromosome_lengths={'20': 63025520, '21': 48129895}, num_ancestral_events=12,
                  num_descendent_events=5, frac_normal=0.4, frac_clone_1=0.35, h_total=0.1)
    mix = simulate.sample_genome_mixture_fast(params, rng=np.random.RandomState(5))

    class RefMix(object):       # the attributes ExperimentSampler.sample_experiment and Experiment.__init__ read
        N, l, cn, frac = mix.N, mix.l, mix.cn, mix.frac
    np.random.seed(77)
    ref = rsim.ExperimentSampler(params).sample_experiment(RefMix())
    np.random.seed(77)
    mine = simulate.sample_experiment(mix, params)
    assert np.array_equal(mine['x'], ref.x)
    assert np.array_equal(mine['phi'], ref.phi)
    assert np.array_equal(mine['h'], ref.h)
    assert np.array_equal(mine['major_is_allele_a'], ref.segment_major_is_allele_a)
    assert np.array_equal(mine['is_outlier_total'], ref.is_outlier_total)
    # both left the stream at the same place
    a = np.random.random_sample()
    np.random.seed(77)
    rsim.ExperimentSampler(params).sample_experiment(RefMix())
    assert a == np.random.random_sample()


def test_fast_generator_makes_model_consistent_experiments():
    from remixt_b200 import cn_model, simulate
    params = dict(N=600, num_ancestral_events=30, num_descendent_events=10, num_false_breakpoints=8, random_seed=3,
                  frac_normal=0.4, frac_clone_1=0.4)
    exp = simulate.simulate_experiment(params)
    again = simulate.simulate_experiment(params)
    assert np.array_equal(exp.x, again.x) and exp.breakpoints == again.breakpoints          # seeded
    assert exp.x.shape == (600, 3) and np.all(exp.x[:, 1] <= exp.x[:, 0]) and np.all(exp.l >= 1)
    cn = exp.true_cn
    assert np.all(cn[:, 0, :] == 1) and cn.min() >= 0 and cn[:, 1:, :].sum(axis=2).max() <= 6
    assert np.abs(cn[:, 1, :] - cn[:, 2, :]).max() <= 1          # within the model's state space
    # true breakpoints join segment ends across which some clone's copy number changes (later events may cancel a
    # change at a shared boundary, so not every single one)
    n_true = n_change = 0
    for bp_id, bp in exp.breakpoints.items():
        if exp.true_brk_cn[bp_id].sum() == 0:
            continue
        for n, side in bp:
            other = n + 1 if side == 1 else n - 1
            n_true += 1
            n_change += bool(np.any(cn[n] != cn[other]))
    assert n_true >= 60 and n_change >= 0.8 * n_true
    # the host driver accepts it: dummy segments, one breakend per model boundary
    bm = cn_model.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, max_depth=1.)
    assert bm.N1 >= 600 and np.all(np.bincount(bm.breakpoint_idx[bm.breakpoint_idx >= 0]) == 2)
