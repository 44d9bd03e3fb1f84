"""One whole chromosome of a fixture as a fixture of its own (tests/golden/exp_<name>chr<c>.npz): the segments of the
chromosome, the adjacencies and breakpoints that lie wholly on it, and the parent's initialisation grid.  Lets the
compiled reference run complete free-running fits on config 2's data (its full genome takes ~50 min per fit):

    python oracle/make_chromosome_fixture.py config2 1
    python oracle/make_golden.py config2chr1 grid --procs 6 --inits 1,9,20,29,40,44
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def main():
    name, chrom = sys.argv[1], sys.argv[2]
    d = np.load(os.path.join(GOLDEN, 'exp_%s.npz' % name), allow_pickle=False)
    on = np.flatnonzero(d['chromosome'].astype(str) == str(chrom))
    assert on.size and np.all(np.diff(on) == 1)
    s0, n = int(on[0]), int(on.size)
    inside = lambda k: s0 <= k < s0 + n
    adj = np.array([(a - s0, b - s0) for a, b in d['adjacencies'] if inside(a) and inside(b)], dtype=np.int64).reshape(-1, 2)
    keep = [i for i, (n1, s1, n2, s2) in enumerate(d['breakpoints'])
            if inside(n1) and inside(n2) and not (n1 == s0 and s1 == 0) and not (n2 == s0 and s2 == 0)]
    brk = d['breakpoints'][keep].copy()
    brk[:, 0] -= s0
    brk[:, 2] -= s0
    out = dict((k, d[k]) for k in d.files if k.startswith('init_') or k == 'true_h')
    out.update(x=d['x'][s0:s0 + n], l=d['l'][s0:s0 + n], adjacencies=adj, breakpoints=brk,
               breakpoint_ids=d['breakpoint_ids'][keep], chromosome=d['chromosome'][s0:s0 + n])
    path = os.path.join(GOLDEN, 'exp_%schr%s.npz' % (name, chrom))
    np.savez_compressed(path, **out)
    print(path, n, 'segments', len(keep), 'breakpoints')


if __name__ == '__main__':
    main()
