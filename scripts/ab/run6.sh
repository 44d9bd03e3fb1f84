mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py tests/test_mstep_gpu.py tests/test_statespace_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/r02c_bench.json 2>gpurun_out/r02c_bench.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02c_bench.json'))
print('step', d['ms_per_step'], 'E', d['e_step_ms_per_step'], 'M', d['m_step_ms_per_step'], 'fullfit', d['full_fit']['wall_s'], d['full_fit']['mstep_seconds'])
for k in ('lltable','chain','round','qsums'): print(k, d['kernels'][k]['launches'], d['kernels'][k]['ms_per_launch'], d['kernels'][k].get('fp64_frac'))
PY
