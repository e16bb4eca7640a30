export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
B=baseline/_ref/bin
( time python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf ) > gpurun_out/gen8b.log 2>&1
python tools/make_gguf.py tiny q4_k_m /tmp/tiny.gguf > /dev/null 2>&1
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 120 $B/llama-bench -m /tmp/tiny.gguf -ngl 99 -fa 1 -p 32 -n 16 -r 1 > gpurun_out/lb_tiny.log 2>&1; echo "exit $?" >> gpurun_out/lb_tiny.log
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 2 > gpurun_out/lb_8b_pp.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp.log
timeout 300 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 0 -p 0 -n 64 -r 1 > gpurun_out/lb_8b_tg_nofa.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg_nofa.log
unset GGML_BACKEND_PATH
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 0 -p 64 -n 16 -r 1 -t $(nproc) > gpurun_out/lb_8b_cpu.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_cpu.log
