mkdir -p gpurun_out
REMIXT_B200_LIB=$PWD/scripts/ab/libremixt_b200_prof.so python scripts/ab/prof_chain.py > gpurun_out/r02b_chain_prof.log 2>&1; echo "rc=$?"
grep c2prof gpurun_out/r02b_chain_prof.log | tail -8; tail -1 gpurun_out/r02b_chain_prof.log
python scripts/time_chain.py config2 > gpurun_out/r02b_chain_new.log 2>&1; echo "new rc=$?"; tail -4 gpurun_out/r02b_chain_new.log
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -3
