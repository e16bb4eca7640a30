export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_mmvq.py -x -q > gpurun_out/fuse_pytest.log 2>&1; tail -15 gpurun_out/fuse_pytest.log
