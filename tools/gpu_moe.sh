export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe.py -x -q > gpurun_out/moe_pytest.log 2>&1; tail -15 gpurun_out/moe_pytest.log
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 600 $B/test-backend-ops test -b B200_0 -o MUL_MAT_ID > gpurun_out/tbo_mmid.log 2>&1; tail -4 gpurun_out/tbo_mmid.log
python tools/make_gguf.py mixtral-8x7b q4_k_m /tmp/mix.gguf > gpurun_out/genmix.log 2>&1; tail -1 gpurun_out/genmix.log
B200_TIMING=1 timeout 900 $B/llama-bench -m /tmp/mix.gguf -ngl 99 -fa 1 -p 512 -n 128 -r 2 > gpurun_out/lb_mix.log 2>&1; grep -E "pp512|tg128|error|stats" gpurun_out/lb_mix.log
B200_MMQ_ID_MIN_T=100000 timeout 900 $B/llama-bench -m /tmp/mix.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/lb_mix_gemv.log 2>&1; grep -E "pp512|error" gpurun_out/lb_mix_gemv.log
timeout 900 $B/llama-batched-bench -m /tmp/mix.gguf -ngl 99 -fa -npp 128 -ntg 32 -npl 1,2,4,8,16,32,64 -c 16384 > gpurun_out/bb_mix.log 2>&1; grep -E "^\|" gpurun_out/bb_mix.log
