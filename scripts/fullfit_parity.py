"""Free-running full-fit parity campaign (GPU): every initialisation of a fixture's grid fitted by
pipeline.fit_grid_batched (CUDA batch, lock-step M-step) against the reference's own free-running fits recorded in
tests/golden/grid_<name>.npz (oracle/make_golden.py grid), plus the reference's OWN sensitivity to a 1e-13
perturbation of h_normal (grid_<name>_perturbed.npz) as the yardstick.

    python scripts/fullfit_parity.py [names...] [--out profiles/r02_fullfit_parity.json] [--mstep lockstep|threads]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402

PARAMS4 = ('negbin_r_0', 'negbin_r_1', 'betabin_M_0', 'betabin_M_1')


def compare(ref, k, cand):
    """cand: dict(elbo, h, params, cn, brk_cn, accept, proportion_divergent) vs row k of the golden grid."""
    rel = lambda a, b: float(np.max(np.abs(np.asarray(a, float) - np.asarray(b, float)) / np.maximum(np.abs(np.asarray(b, float)), 1e-300)))
    cn_ref = ref['cn'][k]
    flips = np.any(cand['cn'].reshape(cn_ref.shape[0], -1) != cn_ref.reshape(cn_ref.shape[0], -1), axis=1)
    acc_ref = ref['accept'][k].ravel().astype(bool)
    acc = np.asarray(cand['accept']).ravel().astype(bool)
    return {
        'elbo_rel_diff': rel(cand['elbo'], ref['elbo'][k]),
        'h_rel_diff': rel(cand['h'], ref['h'][k]),
        'params_rel_diff': rel(cand['params'], ref['params'][k]),
        'map_flip_fraction': float(flips.mean()),
        'map_flips': int(flips.sum()),
        'brk_cn_mismatches': int(np.any(cand['brk_cn'] != ref['brk_cn'][k], axis=1).sum()),
        'decisions_match': bool(acc.shape == acc_ref.shape and np.array_equal(acc, acc_ref)),
        'first_decision_mismatch': (int(np.flatnonzero(acc != acc_ref)[0]) if acc.shape == acc_ref.shape and (acc != acc_ref).any() else None),
        'proportion_divergent_diff': float(abs(cand['proportion_divergent'] - ref['proportion_divergent'][k])),
        # E[LL] after - before of the acceptance test at the first mismatching decision (a margin at rounding level
        # means the reference's own decision there is decided by rounding)
        'margin_at_first_mismatch': (float(cand['margins'][int(np.flatnonzero(acc != acc_ref)[0])])
                                     if 'margins' in cand and acc.shape == acc_ref.shape and (acc != acc_ref).any() else None),
    }


def summarise(rows):
    if not rows:
        return {}
    f = lambda key: [r[key] for r in rows]
    return {
        'fits': len(rows),
        'fits_map_bit_exact': int(sum(r['map_flips'] == 0 for r in rows)),
        'fits_brk_cn_bit_exact': int(sum(r['brk_cn_mismatches'] == 0 for r in rows)),
        'fits_all_decisions_match': int(sum(r['decisions_match'] for r in rows)),
        'map_flip_fraction_mean': float(np.mean(f('map_flip_fraction'))),
        'map_flip_fraction_max': float(np.max(f('map_flip_fraction'))),
        'elbo_rel_diff_median': float(np.median(f('elbo_rel_diff'))),
        'elbo_rel_diff_max': float(np.max(f('elbo_rel_diff'))),
        'h_rel_diff_median': float(np.median(f('h_rel_diff'))),
        'h_rel_diff_max': float(np.max(f('h_rel_diff'))),
    }


def reference_self_sensitivity(name):
    path = os.path.join(util.GOLDEN, 'grid_%s_perturbed.npz' % name)
    if not os.path.exists(path):
        return None
    a, b = util.golden('grid_%s.npz' % name), np.load(path, allow_pickle=False)
    ids_b = list(b['init_ids'])
    rows = []
    for k, i in enumerate(a['init_ids']):
        if int(i) not in ids_b:
            continue
        j = ids_b.index(int(i))
        cand = dict(elbo=b['elbo'][j], h=b['h'][j], params=b['params'][j], cn=b['cn'][j], brk_cn=b['brk_cn'][j],
                    accept=b['accept'][j], proportion_divergent=b['proportion_divergent'][j])
        r = compare(a, k, cand)
        r['init_id'] = int(i)
        rows.append(r)
    return {'summary': summarise(rows), 'selected_init': [int(a['best_init_id']), int(b['best_init_id'])],
            'what': 'the reference against ITSELF with h_normal of every init scaled by 1 + 1e-13'}


def run(name, mstep='lockstep'):
    """`name` is a fixture (`small`, `config1`) or a fixture with the tag of a golden grid at another maximum copy
    number (`small_cn8`, `small_cn12`: tests/golden/grid_small_cn8.npz ...), fitted on the initialisations it holds."""
    from remixt_b200 import pipeline
    full = name
    g = util.golden('grid_%s.npz' % full)
    name = full.split('_cn')[0]
    exp = util.load_experiment(name)
    ids = sorted(int(i) for i in list(g['init_ids']) + list(g['failed_ids']))        # what the reference fitted
    inits = dict((i, util.init_params(name, i)) for i in ids)
    cfg = {'max_copy_number': int(g['max_copy_number']) if 'max_copy_number' in g.files else 6,
           'num_em_iter': int(g['num_em_iter']), 'num_update_iter': int(g['num_update_iter'])}
    prof = {}
    t0 = time.perf_counter()
    trace = []
    res, best, errors = pipeline.fit_grid_batched(exp, inits, cfg, random_seed=int(g['seed']), mstep=mstep, profile=prof,
                                                  trace=trace)
    wall = time.perf_counter() - t0
    gi = [int(i) for i in g['init_ids']]
    rows = []
    for i, r in res.items():
        if i not in gi:
            continue
        cand = dict(elbo=r['stats']['elbo'], h=r['h'], params=[r['stats'][k] for k in PARAMS4], cn=r['cn'],
                    brk_cn=np.array([r['brk_cn'][str(b)] for b in g['brk_ids']]),
                    accept=[d[1] for d in r['decisions']], margins=[d[3] - d[2] for d in r['decisions']], proportion_divergent=r['stats']['proportion_divergent'])
        row = compare(g, gi.index(i), cand)
        row['init_id'] = int(i)
        # trajectory: relative deviation of (h, dispersions, ELBO) from the reference after every EM iteration
        k = gi.index(i)
        rel = lambda a, b: float(np.max(np.abs(np.asarray(a, float) - np.asarray(b, float)) / np.maximum(np.abs(np.asarray(b, float)), 1e-300)))
        pos = list(ids).index(i)      # position of the fit in the batch (the trace rows)
        row['trajectory_rel_dev'] = [[rel(t[0][pos], g['trace_h'][k][e]), rel(t[1][pos], g['trace_params'][k][e]),
                                      rel(t[2][pos], g['trace_elbo'][k][e])] for e, t in enumerate(trace)]
        row['margins'] = [float(m) for m in cand['margins']]
        rows.append(row)
    out = {
        'fixture': full, 'max_copy_number': cfg['max_copy_number'], 'mstep': mstep, 'wall_s': wall, 'profile': prof,
        'selected_init': {'reference': int(g['best_init_id']), 'b200': None if best is None else int(best),
                          'equal': best is not None and int(best) == int(g['best_init_id'])},
        'failed_inits': {'reference': [int(i) for i in g['failed_ids']], 'b200': sorted(int(i) for i in errors)},
        'summary': summarise(rows),
        'reference_self_sensitivity': reference_self_sensitivity(full),
        'fits': rows,
    }
    sel = [r for r in rows if r['init_id'] == int(g['best_init_id'])]
    out['selected_solution'] = sel[0] if sel else None
    return out


def main():
    args = sys.argv[1:]
    out_path = args[args.index('--out') + 1] if '--out' in args else os.path.join(ROOT, 'profiles', 'r02_fullfit_parity.json')
    mstep = args[args.index('--mstep') + 1] if '--mstep' in args else 'lockstep'
    names = [a for i, a in enumerate(args) if not a.startswith('--') and (i == 0 or args[i - 1] not in ('--out', '--mstep'))]
    names = names or ['small', 'config1']
    report = {'what': __doc__.split('\n\n')[0], 'grids': [run(n, mstep) for n in names]}
    with open(out_path, 'w') as f:
        json.dump(report, f, indent=1)
    for gr in report['grids']:
        print(gr['fixture'], 'wall %.1f s' % gr['wall_s'], json.dumps(gr['selected_init']), json.dumps(gr['failed_inits']))
        print('   b200 vs reference :', json.dumps(gr['summary']))
        if gr['reference_self_sensitivity']:
            print('   reference vs itself:', json.dumps(gr['reference_self_sensitivity']['summary']))


if __name__ == '__main__':
    main()
