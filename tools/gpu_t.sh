export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py -x -q -k rope > gpurun_out/t.log 2>&1; tail -25 gpurun_out/t.log
