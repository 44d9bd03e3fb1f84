"""Host steps around the fit: initialisation candidates and grid (readdepth / pipeline.init), result
tables, collate, write_results -- against the reference's own functions run in this process (oracle/refenv: the
reference tree or the copies oracle/build_ref.py stages under oracle/_ref) and against the grids the reference produced
into the fixtures."""
import os
import warnings

import numpy as np
import pandas as pd
import pytest

from tests import util


@pytest.fixture(scope='module')
def ref():
    from oracle import refenv
    if not refenv.available():
        pytest.skip('reference modules not available')
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import remixt.analysis.experiment as rexp
        import remixt.analysis.pipeline as rpipe
        import remixt.analysis.readdepth as rdepth
    return dict(experiment=rexp, pipeline=rpipe, readdepth=rdepth)


def _ref_experiment(name):
    from oracle import make_golden
    return make_golden.load_experiment(name)[1]


@pytest.mark.parametrize('name', ['small', 'config1', 'config2'])
def test_init_grid_equals_fixture_grid(name):
    """pipeline.init on the fixture == the grid the reference's pipeline.init stored in it (oracle/make_golden.py init).
    The k-means of calculate_minor_modes is sklearn's on both sides; its threaded reductions are not bit-reproducible
    from process to process, hence 1e-12 and not equality here (equality in-process is the next test)."""
    from remixt_b200 import pipeline
    d = util.golden('exp_%s.npz' % name)
    grid = pipeline.init(util.load_experiment(name), {'max_copy_number': 6, 'max_ploidy': float(d['init_max_ploidy'])})
    assert len(grid) == len(d['init_h_normal'])
    for i, p in grid.items():
        assert p['mode_idx'] == int(d['init_mode_idx'][i])
        assert p['mix_frac'] == float(d['init_mix_frac'][i]) and p['divergence_weight'] == float(d['init_divergence_weight'][i])
        for k in ('h_normal', 'h_tumour', 'max_depth'):
            assert abs(p[k] - float(d['init_' + k][i])) <= 1e-12 * abs(p[k]), (i, k)


def test_readdepth_equals_reference_in_process(ref):
    """calculate_depth / candidates / estimate_ploidy == the reference's, bit for bit; calculate_minor_modes to the
    reproducibility of sklearn's own k-means."""
    from remixt_b200 import readdepth
    exp = util.load_experiment('config1')
    rex = _ref_experiment('config1')
    mine = readdepth.calculate_depth(exp)
    theirs = ref['readdepth'].calculate_depth(rex)
    assert list(mine.columns) == list(theirs.columns)
    for c in mine.columns:
        assert np.array_equal(mine[c].values, theirs[c].values), c
    np.random.seed(1234)
    m1 = readdepth.calculate_minor_modes(mine)
    np.random.seed(1234)
    m2 = ref['readdepth'].calculate_minor_modes(theirs)
    # sklearn's k-means reduces over threads: the same call on the same data is reproducible to an ulp or two only
    assert m1.shape == m2.shape and np.allclose(m1, m2, rtol=1e-12, atol=0.)
    c1 = readdepth.calculate_candidate_h_monoclonal(m1)
    c2 = ref['readdepth'].calculate_candidate_h_monoclonal(m1)
    assert len(c1) == len(c2) and all(np.array_equal(a, b) for a, b in zip(c1, c2))
    for h in c1[:3]:
        assert readdepth.estimate_ploidy(h, exp) == ref['readdepth'].estimate_ploidy(h, rex)
        assert readdepth.estimate_ploidy(h, exp, read_depth=mine) == ref['readdepth'].estimate_ploidy(h, rex)
    fixed = readdepth.calculate_candidate_h_monoclonal(m1, h_normal=0.03, h_tumour=0.05)
    assert np.array_equal(fixed, ref['readdepth'].calculate_candidate_h_monoclonal(m1, h_normal=0.03, h_tumour=0.05))


def test_init_raises_when_depth_cut_too_strict():
    """analysis/pipeline.py:90-94: more than a quarter of the genome above max_depth is an error (the simulated genomes
    at max_copy_number 6 need max_ploidy <= 4)."""
    from remixt_b200 import pipeline
    with pytest.raises(ValueError, match='Unable to model'):
        pipeline.init(util.load_experiment('config2'), {'max_copy_number': 6})


def _fake_results(exp, seed):
    rng = np.random.RandomState(seed)
    n = exp.l.shape[0]
    cn = rng.randint(0, 4, size=(n, 3, 2))
    cn[:, 0, :] = 1
    brk_cn = dict((k, rng.randint(0, 3, size=3)) for k in exp.breakpoints)
    h = np.array([0.04, 0.03, 0.02]) * (1 + seed)
    stats = dict(elbo=-1000. - 7 * ((seed * 5) % 3), elbo_diff=0.1, error_message='', negbin_r_0=500., negbin_r_1=10.,
                 betabin_M_0=500., betabin_M_1=10., num_clones=3, num_segments=n, ploidy=2. + 0.1 * seed,
                 proportion_divergent=0.1 + 0.25 * seed, mode_idx=seed, divergence_weight=1e-7)
    return dict(h=h, cn=cn, brk_cn=brk_cn, p_outlier_total=rng.rand(n, 2), p_outlier_allele=rng.rand(n, 2),
                total_likelihood_mask=rng.randint(0, 2, size=n), allele_likelihood_mask=rng.randint(0, 2, size=n),
                stats=stats)


def test_result_tables_equal_reference(ref):
    """create_cn_table / create_brk_cn_table column for column against the reference's."""
    from remixt_b200 import results
    exp = util.load_experiment('small')
    rex = _ref_experiment('small')
    r = _fake_results(exp, 1)
    mine = results.create_cn_table(exp, r['cn'], r['h'])
    theirs = ref['experiment'].create_cn_table(rex, r['cn'], r['h'])
    assert list(mine.columns) == list(theirs.columns)
    for c in mine.columns:
        assert np.array_equal(mine[c].values, theirs[c].values), c
    seg = results.breakpoint_segment_table(exp)
    assert set(seg['prediction_id']) == set(exp.breakpoints)
    mine_b = results.create_brk_cn_table(r['brk_cn'], seg)
    theirs_b = ref['experiment'].create_brk_cn_table(r['brk_cn'], seg)
    assert list(mine_b.columns) == list(theirs_b.columns) and np.array_equal(mine_b.values, theirs_b.values)
    assert list(results.create_brk_cn_table({}, seg).columns) == ['prediction_id']


def test_collate_store_layout_and_selection(ref, tmp_path):
    """collate: the reference's keys; the selected solution is the one the reference's store_optimal_solution picks;
    the store round-trips through disk; write_results_tables writes the three user-facing files."""
    from oracle import make_golden
    from remixt_b200 import results
    exp = util.load_experiment('small')
    fits = dict((i, _fake_results(exp, i)) for i in range(3))
    cfg = {}
    path = str(tmp_path / 'collated')
    with results.TableStore(path, 'w') as store:
        best = results.collate(store, exp, fits, cfg, init_store={'minor_modes': pd.Series([0.1, 0.2])})
    rows = [dict(init_id=i, elbo=f['stats']['elbo'], proportion_divergent=f['stats']['proportion_divergent'])
            for i, f in fits.items()]
    assert best == make_golden.reference_selection(rows, cfg)
    from remixt_b200 import pipeline
    assert best == pipeline.select_solution(dict((i, f['stats']) for i, f in fits.items()), cfg)
    with results.TableStore(path) as store:
        keys = set(store.keys())
        want = {'/stats', '/minor_modes', '/cn', '/mix', '/brk_cn'}
        for i in fits:
            want |= {'/solutions/solution_%d/%s' % (i, k) for k in ('h', 'cn', 'mix', 'brk_cn')}
        assert keys == want
        assert np.allclose(store['solutions/solution_%d/h' % best].values, fits[best]['h'])
        assert np.array_equal(store['/cn']['major_1'].values, fits[best]['cn'][:, 1, 0])
        assert np.array_equal(store['/cn']['prob_is_outlier_total'].values, fits[best]['p_outlier_total'][:, 1])
        assert np.allclose(store['/mix'].values, fits[best]['h'] / fits[best]['h'].sum())
        assert list(store['stats']['init_id']) == [0, 1, 2]
        sol = results.write_results_tables(store, str(tmp_path / 'cn.tsv'), str(tmp_path / 'brk_cn.tsv'),
                                           str(tmp_path / 'meta.yaml'))
    assert sol == best
    cn = pd.read_csv(str(tmp_path / 'cn.tsv'), sep='\t')
    assert len(cn) == exp.l.shape[0] and 'minor_2' in cn.columns
    import yaml
    meta = yaml.safe_load(open(str(tmp_path / 'meta.yaml')))
    assert meta['init_id'] == best and len(meta['h']) == 3
