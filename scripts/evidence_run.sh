set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/r01_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r01_smoke.log 2>&1
python bench.py --impl reference > gpurun_out/r01_bench_reference.json 2> gpurun_out/r01_bench_reference.err
python bench.py > gpurun_out/r01_bench_config2.json 2> gpurun_out/r01_bench_config2.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r01_launches_config2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_lltable2|k_chain_mma|k_round|k_hist2" -c 4 -f -o gpurun_out/r01_top4 python bench.py --no-cpu-baseline --fits 8 --steps 1 --warmup 1 > gpurun_out/r01_ncu_top4.log 2>&1
tail -2 gpurun_out/r01_pytest_gpu.log; cat gpurun_out/r01_smoke.log | tail -2; cat gpurun_out/r01_bench_config2.json | cut -c1-600; cat gpurun_out/r01_bench_reference.json | cut -c1-400
