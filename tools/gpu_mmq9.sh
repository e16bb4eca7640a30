# MMQ after the epilogue / single-thread issue rework: parity, per-shape timing, prefill
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mmq.py tests/test_gpu_moe.py tests/test_gpu_fattn.py -q -x 2>&1 | tail -3
timeout 300 python tools/mmq_debug.py time q4_K q6_K 2>&1 | grep -v "M=2048\|M= 128"
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
for i in 1 2; do
  timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 5 2>&1 | grep -E "pp512|pp2048"
done
