mkdir -p gpurun_out
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
timeout 600 python tools/mmq_repro.py q4_0 q4_K > gpurun_out/mmq_repro.log 2>&1; echo "exit $?" >> gpurun_out/mmq_repro.log
cat gpurun_out/mmq_repro.log
