mkdir -p gpurun_out
timeout 200 python scripts/time_chain.py config2 > gpurun_out/r02c_chain_new.log 2>&1; echo "new rc=$?"; tail -4 gpurun_out/r02c_chain_new.log
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -3
