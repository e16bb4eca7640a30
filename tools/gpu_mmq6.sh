export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python tools/mmq_debug.py trace q4_K > gpurun_out/mmq_trace.log 2>&1; echo "exit $?" >> gpurun_out/mmq_trace.log
head -75 gpurun_out/mmq_trace.log
