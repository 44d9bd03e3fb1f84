"""Other state spaces than configs 1/2's S = 97 (S = 165 at max_copy_number 8 is config 3; four clones
give S = 457 at max_copy_number 6): the CUDA paths that S <= 128 does not use -- transition matrix streamed from L2
in the chain kernel, the histogram kernel with shared-memory atomics, the unfused update sequence -- side by side
with the oracle on a subsample the oracle's (N, S, S) arrays allow."""
import sys

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

RTOL = 1e-6


def _sub(nseg):
    sys.path.insert(0, util.ROOT)
    import bench
    return bench.subsample_experiment(util.load_experiment('small'), nseg)


def _drive(exp, h, ref_class, **kw):
    from remixt_b200 import cn_model
    ip = util.init_params('small', 1)
    args = dict(max_copy_number=6, normal_contamination=True, divergence_weight=ip['divergence_weight'],
                min_segment_length=10000, min_proportion_genotyped=0.01, max_depth=ip['max_depth'],
                normal_copies=np.array([[1, 1]] * exp.l.shape[0]))
    args.update(kw)
    out = []
    for mc in (None, ref_class):
        bm = cn_model.BreakpointModel(exp.x, exp.l, exp.adjacencies, exp.breakpoints, model_class=mc, **args)
        bm.num_em_iter = 0
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            bm.fit(np.asarray(h, dtype=float))
        out.append(bm.model)
    return out


def _compare(cu, ref, rounds=2):
    util.assert_close(cu.calculate_elbo(), ref.calculate_elbo(), rtol=RTOL, what='ctor elbo')
    for r in range(rounds):
        for step in util.STEPS:
            getattr(cu, step)()
            getattr(ref, step)()
            util.assert_close(cu.calculate_elbo(), ref.calculate_elbo(), rtol=RTOL, what='round %d %s elbo' % (r, step))
        util.assert_close(cu.p_breakpoint, np.asarray(ref.p_breakpoint), rtol=RTOL, atol=1e-280, what='p_breakpoint')
        util.assert_close(cu.posterior_marginals, np.asarray(ref.posterior_marginals), rtol=RTOL, atol=1e-200,
                          what='posterior')
        util.assert_close(cu.hmm_log_norm_const, ref.hmm_log_norm_const, rtol=RTOL, what='logZ')
    for a in ('p_allele_swap', 'p_outlier_total', 'p_outlier_allele'):
        util.assert_close(getattr(cu, a), np.asarray(getattr(ref, a)), rtol=RTOL, atol=1e-280, what=a)
    M = cu.num_clones
    c1 = np.zeros((cu.num_segments, M, 2), dtype=np.int64)
    c2 = np.zeros_like(c1)
    cu.infer_cn(c1)
    ref.infer_cn(c2)
    assert np.array_equal(c1, c2)
    g1, g2 = np.zeros(M), np.zeros(M)
    ones = np.ones(cu.num_segments, dtype=int)
    cu.calculate_expected_log_likelihood_partial_h(ones, g1)
    ref.calculate_expected_log_likelihood_partial_h(ones, g2)
    util.assert_close(g1, g2, rtol=RTOL, what='dE[LL]/dh')


def test_max_copy_number_8():
    """Config 3's state space: M = 3, max_copy_number = 8 -> S = 165, B = 25."""
    from oracle import bpmodel_np
    ip = util.init_params('small', 1)
    from remixt_b200 import pipeline
    cu, ref = _drive(_sub(80), pipeline.init_h(ip), bpmodel_np.RemixtModel, max_copy_number=8)
    assert cu.num_cn_states == 165
    _compare(cu, ref)
    # the fused loop entry point takes the unfused sequence here and matches the step-by-step state
    cu2, _ = _drive(_sub(80), pipeline.init_h(ip), bpmodel_np.RemixtModel, max_copy_number=8)
    cu2.variational_update(2)
    assert np.array_equal(cu2.posterior_marginals, cu.posterior_marginals)


def test_max_copy_number_12(ref_bpmodel):
    """The reference's DEFAULT max_copy_number (12): M = 3 -> S = 355.  The chain messages take k_chain_mma_s (eight fits
    per read of the L2-resident class matrix, B fragments loaded under the DMMAs, scalar breakend rows); against the
    COMPILED reference on 120 segments with their breakpoints (three (N-1, S, S) arrays of 0.12 GB each)."""
    ip = util.init_params('small', 1)
    from remixt_b200 import pipeline
    cu, ref = _drive(_sub(120), pipeline.init_h(ip), ref_bpmodel.RemixtModel, max_copy_number=12,
                     max_depth=2. * ip['h_normal'] + 12.25 * ip['h_tumour'])
    assert cu.num_cn_states == 355 and cu.num_breakpoints > 0
    _compare(cu, ref)


def test_four_clones():
    """Three tumour clones + normal: M = 4, max_copy_number = 6 -> S = 457."""
    from oracle import bpmodel_np
    ip = util.init_params('small', 1)
    h = np.array([ip['h_normal'], 0.5 * ip['h_tumour'], 0.3 * ip['h_tumour'], 0.2 * ip['h_tumour']])
    cu, ref = _drive(_sub(30), h, bpmodel_np.RemixtModel)
    assert cu.num_clones == 4 and cu.num_cn_states == 457
    _compare(cu, ref, rounds=1)


def test_four_clones_max_copy_number_8():
    """SURVEY.md section 8's stress state space: three tumour clones + normal at max_copy_number 8 -> S = 805, B = 57
    (the literal reading of BASELINE config 3).  24 segments against the numpy oracle (three (N-1, S, S) arrays of
    0.12 GB): the vector-pipe chain kernel, the first-generation histograms, the unfused round."""
    from oracle import bpmodel_np
    ip = util.init_params('small', 1)
    h = np.array([ip['h_normal'], 0.5 * ip['h_tumour'], 0.3 * ip['h_tumour'], 0.2 * ip['h_tumour']])
    cu, ref = _drive(_sub(24), h, bpmodel_np.RemixtModel, max_copy_number=8,
                     max_depth=2. * ip['h_normal'] + 8.25 * ip['h_tumour'])
    assert cu.num_clones == 4 and cu.num_cn_states == 805
    _compare(cu, ref, rounds=1)


def _chromosome(name, chrom):
    """One whole chromosome of a fixture as its own experiment (breakpoints with both ends on it)."""
    sys.path.insert(0, util.ROOT)
    import bench
    exp = util.load_experiment(name)
    on = np.flatnonzero(np.asarray(exp.segment_chromosome_id).astype(str) == str(chrom))
    assert on.size and np.all(np.diff(on) == 1)
    return bench.subsample_experiment(exp, int(on.size), start=int(on[0]))


def test_config3_state_space_on_config1_vs_compiled_reference(ref_bpmodel):
    """BASELINE config 3's state space (max copy number 8, S = 165) on the 1000 segments / 64 breakpoints of config 1,
    against the COMPILED reference (three (N-1, S, S) arrays of 0.23 GB each): every update of two rounds, MAP path,
    gradient."""
    exp = util.load_experiment('config1')
    ip = util.init_params('config1', 1)
    from remixt_b200 import pipeline
    args = dict(max_copy_number=8, max_depth=2. * ip['h_normal'] + 8.25 * ip['h_tumour'],
                normal_copies=np.array([[1, 1]] * exp.l.shape[0]))
    cu = util.make_driver(exp, ip, **args).model
    ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel, **args).model
    assert cu.num_cn_states == 165 and cu.num_segments > 1000
    _compare(cu, ref, rounds=2)


def test_whole_chromosome_of_config2_vs_compiled_reference(ref_bpmodel):
    """The benched workload at the scale the reference can hold: chromosome 1 of config 2 (about 2.4 k segments,
    3 x 0.18 GB of transition arrays), GPU against the compiled reference through two rounds and the MAP decode."""
    exp = _chromosome('config2', '1')
    assert exp.l.shape[0] > 2000
    ip = util.init_params('config2', 5)
    cu = util.make_driver(exp, ip).model
    ref = util.make_driver(exp, ip, model_class=ref_bpmodel.RemixtModel).model
    assert cu.num_cn_states == 97 and cu.num_breakpoints >= 5
    _compare(cu, ref, rounds=2)
