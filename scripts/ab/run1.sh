mkdir -p gpurun_out
python scripts/time_chain.py config2 > gpurun_out/r02b_chain_new.log 2>&1; echo "new rc=$?"
REMIXT_B200_LIB=$PWD/scripts/ab/libremixt_b200_base.so python scripts/time_chain.py config2 > gpurun_out/r02b_chain_base.log 2>&1; echo "base rc=$?"
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02b_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02b_chain_new.log; tail -4 gpurun_out/r02b_chain_base.log; tail -5 gpurun_out/r02b_pytest.log
