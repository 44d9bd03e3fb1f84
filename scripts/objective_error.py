"""How far is the device's M-step objective from the compiled reference's on IDENTICAL inputs?  Both models get the
same posteriors and q-factors (the device's, written into the reference model), then E[LL] on one sample is evaluated
over a sweep of each dispersion parameter.  Prints the size of the difference and of its erratic part (the difference
minus its own smooth trend), which is what can change a Nelder-Mead comparison."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refenv  # noqa: E402
from tests import util  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'config1'
ref = refenv.load_ref_bpmodel()
exp = util.load_experiment(name)
ip = util.init_params(name, 1)
cu = util.make_driver(exp, ip).model
rf = util.make_driver(exp, ip, model_class=ref.RemixtModel).model
cu.variational_update(5)
for a in ('posterior_marginals', 'p_allele_swap', 'p_outlier_total', 'p_outlier_allele'):
    setattr(rf, a, np.ascontiguousarray(getattr(cu, a)))
rng = np.random.RandomState(0)
sample = np.zeros(cu.num_segments, dtype=np.int64)
sample[rng.choice(cu.num_segments, min(200, cu.num_segments // 10), replace=False)] = 1
full = np.ones(cu.num_segments, dtype=np.int64)
for pname, centre in (('negbin_r_0', 700.), ('negbin_r_1', 12.), ('betabin_M_0', 900.), ('betabin_M_1', 15.)):
    for label, s in (('sample', sample), ('all', full)):
        xs = centre * (1. + 1e-4 * np.arange(-20, 21))
        d = []
        for x in xs:
            setattr(cu, pname, float(x))
            setattr(rf, pname, float(x))
            d.append(cu.calculate_expected_log_likelihood(s) - rf.calculate_expected_log_likelihood(s))
        d = np.array(d)
        trend = np.polyval(np.polyfit(np.arange(len(d)), d, 2), np.arange(len(d)))
        f = rf.calculate_expected_log_likelihood(s)
        print('%-12s %-6s E[LL] %.6e  diff mean %.3e  max |diff| %.3e  erratic rms %.3e  (relative %.1e)' % (
            pname, label, f, d.mean(), np.abs(d).max(), np.sqrt(np.mean((d - trend) ** 2)), np.sqrt(np.mean((d - trend) ** 2)) / abs(f)))
    setattr(cu, pname, centre)
    setattr(rf, pname, centre)
