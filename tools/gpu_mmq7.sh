export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mmq.py -q -x > gpurun_out/mmq_pytest.log 2>&1; echo "exit $?" >> gpurun_out/mmq_pytest.log
tail -3 gpurun_out/mmq_pytest.log
timeout 300 python tools/mmq_debug.py time q4_K > gpurun_out/mmq_time.log 2>&1; echo "exit $?" >> gpurun_out/mmq_time.log
cat gpurun_out/mmq_time.log
