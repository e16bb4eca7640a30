export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python tools/mmq_debug.py check > gpurun_out/mmq_check.log 2>&1; echo "exit $?" >> gpurun_out/mmq_check.log
cat gpurun_out/mmq_check.log | tail -40
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py tiny q4_k_m /tmp/tiny.gguf > /dev/null 2>&1
B200_DEBUG=1 timeout 120 $B/llama-bench -m /tmp/tiny.gguf -ngl 99 -fa 1 -p 32 -n 0 -r 1 2>&1 | grep -v "^/\|^$" | head -20
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_DEBUG=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 1 2>&1 | grep -v "^/\|^$" | head -20
timeout 300 python tools/mmq_debug.py time q4_K q6_K > gpurun_out/mmq_time.log 2>&1; echo "exit $?" >> gpurun_out/mmq_time.log
cat gpurun_out/mmq_time.log
