export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py -x -q > gpurun_out/fuse_pytest.log 2>&1; tail -3 gpurun_out/fuse_pytest.log
KT_N=12 bash tools/gpu_ktrace.sh
