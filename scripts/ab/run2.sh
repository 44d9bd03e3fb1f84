mkdir -p gpurun_out
REMIXT_B200_LIB=$PWD/scripts/ab/libremixt_b200_prof.so python scripts/ab/prof_chain.py > gpurun_out/r02b_chain_prof.log 2>&1; echo "rc=$?"
grep c2prof gpurun_out/r02b_chain_prof.log | tail -8; tail -1 gpurun_out/r02b_chain_prof.log
