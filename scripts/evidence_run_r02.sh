# One evidence run of round 2 on the GPU box (gpurun): everything lands in gpurun_out/r02_*.
set -x
mkdir -p gpurun_out
T=r02
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_gpu_full.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest_gpu_full.log > gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
timeout 600 python bench.py > gpurun_out/${T}_bench_config2.json 2> gpurun_out/${T}_bench_config2.err
timeout 600 python scripts/fullfit_parity.py small config1 small_cn8 small_cn12 config2chr1 --out gpurun_out/${T}_fullfit_parity.json > gpurun_out/${T}_fullfit_parity.log 2>&1
timeout 300 python scripts/fullfit_time.py config2 > gpurun_out/${T}_fullfit_time.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${T}_launches_config2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_lltable2|k_chain_mma2|k_round|k_hist4|k_ell_sample|k_qsums" -c 5 -f -o gpurun_out/${T}_top python bench.py --no-cpu-baseline --fits 8 --steps 1 --warmup 1 > gpurun_out/${T}_ncu_top.log 2>&1
cat gpurun_out/${T}_noalloc_alone.log | tail -3; tail -3 gpurun_out/${T}_pytest_gpu.log; tail -2 gpurun_out/${T}_smoke.log
cut -c1-700 gpurun_out/${T}_bench_config2.json; cut -c1-400 gpurun_out/${T}_bench_reference.json; tail -8 gpurun_out/${T}_fullfit_parity.log; tail -1 gpurun_out/${T}_fullfit_time.log | cut -c1-500
timeout 300 python bench.py --workload config3 --no-cpu-baseline > gpurun_out/${T}_bench_config3.json 2> gpurun_out/${T}_bench_config3.err; cut -c1-300 gpurun_out/${T}_bench_config3.json
