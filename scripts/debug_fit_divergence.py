"""Diagnostic (test infrastructure): run the same seeded full fit on the CUDA model and on the compiled
reference model side by side and print where the two trajectories part."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import util
from oracle import refenv
from remixt_b200 import pipeline, cn_model

name = sys.argv[1] if len(sys.argv) > 1 else 'small'
g = util.golden('fit_%s.npz' % name)
exp = util.load_experiment(name)
ip = util.init_params(name, int(g['init_id']))
cfg = {'max_copy_number': 6, 'num_em_iter': 5, 'num_update_iter': 5}
ref = refenv.load_ref_bpmodel()

log = {}
orig_update_h = cn_model.BreakpointModel.update_h
orig_update_param = cn_model.BreakpointModel.update_param


def traced(tag):
    rec = log.setdefault(tag, [])

    def update_h(self):
        rec.append(('elbo_pre_h', self.model.calculate_elbo()))
        orig_update_h(self)
        rec.append(('h', np.asarray(self.model.h).copy()))

    def update_param(self, nm):
        orig_update_param(self, nm)
        rec.append((nm, getattr(self.model, nm)))
    return update_h, update_param


for tag, cls in (('cuda', None), ('ref', ref.RemixtModel)):
    cn_model.BreakpointModel.update_h, cn_model.BreakpointModel.update_param = traced(tag)
    np.random.seed(1234)
    res = pipeline.fit(exp, ip, cfg, model_class=cls)
    log[tag].append(('final_elbo', res['stats']['elbo']))

for (ka, va), (kb, vb) in zip(log['cuda'], log['ref']):
    va, vb = np.asarray(va, dtype=float), np.asarray(vb, dtype=float)
    rel = np.max(np.abs(va - vb) / np.maximum(np.abs(vb), 1e-300))
    print('%-14s rel=%.3e cuda=%s ref=%s' % (ka, rel, va, vb))
