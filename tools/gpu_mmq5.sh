export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python tools/mmq_debug.py check > gpurun_out/mmq_check.log 2>&1; echo "exit $?" >> gpurun_out/mmq_check.log
grep -v "^ " gpurun_out/mmq_check.log | tail -34
timeout 900 python -m pytest tests/test_gpu_mmq.py -q > gpurun_out/mmq_pytest.log 2>&1; echo "exit $?" >> gpurun_out/mmq_pytest.log
tail -4 gpurun_out/mmq_pytest.log
timeout 300 python tools/mmq_debug.py time q4_K q6_K > gpurun_out/mmq_time.log 2>&1; echo "exit $?" >> gpurun_out/mmq_time.log
cat gpurun_out/mmq_time.log
