export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fattn.py -x -q > gpurun_out/fa_pytest.log 2>&1; tail -5 gpurun_out/fa_pytest.log
timeout 600 python tools/fa_micro.py > gpurun_out/fa_micro.log 2>&1; cat gpurun_out/fa_micro.log
