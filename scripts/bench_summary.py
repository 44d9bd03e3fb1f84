"""Print the per-kernel lines of a bench.py JSON line (argument: path)."""
import json
import sys

d = json.load(open(sys.argv[1]))
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'])
for k, v in d['kernels'].items():
    print('%-12s %8.3f ms  fp64 %s  hbm %s' % (k, v.get('ms_per_launch'), v.get('fp64_frac'), v.get('hbm_frac')))
