export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fattn.py tests/test_gpu_moe.py -x -q > gpurun_out/fa_moe_pytest.log 2>&1; tail -5 gpurun_out/fa_moe_pytest.log
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 900 $B/test-backend-ops test -b B200_0 -o MUL_MAT > gpurun_out/tbo_mm.log 2>&1; tail -3 gpurun_out/tbo_mm.log
timeout 900 $B/test-backend-ops test -b B200_0 -o FLASH_ATTN_EXT > gpurun_out/tbo_fa.log 2>&1; tail -3 gpurun_out/tbo_fa.log
python tools/make_gguf.py mixtral-8x7b q4_k_m /tmp/mix.gguf > gpurun_out/genmix.log 2>&1
B200_TIMING=1 timeout 900 $B/llama-bench -m /tmp/mix.gguf -ngl 99 -fa 1 -p 512 -n 128 -r 2 > gpurun_out/lb_mix.log 2>&1; grep -E "pp512|tg128|error|stats" gpurun_out/lb_mix.log
timeout 900 $B/llama-batched-bench -m /tmp/mix.gguf -ngl 99 -fa -npp 128 -ntg 32 -npl 1,2,4,8,16,32,64 -c 16384 > gpurun_out/bb_mix.log 2>&1; grep -E "^\|" gpurun_out/bb_mix.log
python tools/make_gguf.py llama3-8b q8_0 /tmp/l8b-q8.gguf > gpurun_out/gen8bq8.log 2>&1
timeout 900 $B/llama-bench -m /tmp/l8b-q8.gguf -ngl 99 -fa 1 -ctk q8_0 -ctv q8_0 -p 512,4096 -n 128 -r 2 > gpurun_out/lb_8b_q8kv.log 2>&1; grep -E "pp|tg|error" gpurun_out/lb_8b_q8kv.log
timeout 900 $B/llama-bench -m /tmp/l8b-q8.gguf -ngl 99 -fa 1 -ctk q8_0 -ctv q8_0 -p 512 -n 32 -d 32768 -r 2 > gpurun_out/lb_8b_q8kv_d32k.log 2>&1; grep -E "pp|tg|error" gpurun_out/lb_8b_q8kv_d32k.log
timeout 900 $B/llama-bench -m /tmp/l8b-q8.gguf -ngl 99 -fa 1 -p 512,4096 -n 128 -r 2 > gpurun_out/lb_8b_q8_f16kv.log 2>&1; grep -E "pp|tg|error" gpurun_out/lb_8b_q8_f16kv.log
