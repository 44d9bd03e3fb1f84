mkdir -p gpurun_out
T=r02
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_gpu_full.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest_gpu_full.log > gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/${T}_bench_config2.json 2> gpurun_out/${T}_bench_config2.err
timeout 300 python scripts/fullfit_time.py config2 > gpurun_out/${T}_fullfit_time.log 2>&1
tail -3 gpurun_out/${T}_noalloc_alone.log; tail -3 gpurun_out/${T}_pytest_gpu.log; tail -2 gpurun_out/${T}_smoke.log; cut -c1-300 gpurun_out/${T}_bench_config2.json
