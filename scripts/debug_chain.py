import sys, os
import numpy as np
sys.path.insert(0, '.')
from tests import util
from remixt_b200 import _lib
exp = util.load_experiment('small')
g = util.golden('steps_small.npz')
ip = util.init_params('small', int(g['init_id']))
cu = util.make_driver(exp, ip).model
cu.update_p_allele_swap()
try:
    cu.update_p_cn()
except Exception as e:
    print('ERR', e)
p = np.asarray(cu.posterior_marginals)
print('N,S', p.shape, 'nan rows', np.where(np.isnan(p).any(axis=1))[0][:20], 'count', int(np.isnan(p).any(axis=1).sum()))
print('is_telomere idx', np.where(np.asarray(cu.is_telomere))[0][:10], 'brk idx', np.where(np.asarray(cu.breakpoint_idx) >= 0)[0][:10])
np.save(sys.argv[1], p)
print('hmm_log_norm_const', cu.hmm_log_norm_const)
