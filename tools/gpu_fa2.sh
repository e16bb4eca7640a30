export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fattn.py tests/test_gpu_fused.py -x -q > gpurun_out/fa_pytest.log 2>&1; tail -5 gpurun_out/fa_pytest.log
bash tools/gpu_ktrace.sh
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 128 -r 3 2>&1 | grep -E "pp512|tg128"
timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 32 -d 300,4096 -r 2 2>&1 | grep -E "tg32"
