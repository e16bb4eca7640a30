export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fused.py -q 2>&1 | tail -5
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
grep -E "tg128|ggml-b200: (stats|cuda)|error|exit" gpurun_out/lb_8b_tg.log
B200_FUSE=0 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128|error"
for f in 1 0; do B200_FUSE=$f timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/fuse$f.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse=$f exit $?"; done
cmp /tmp/fuse1.bin /tmp/fuse0.bin && echo "fused == unfused logits: bit-identical"
unset GGML_BACKEND_PATH
timeout 900 python -m pytest tests/test_gpu_plugin.py -q -k "logits or llama_bench" 2>&1 | grep -E "passed|failed|FAILED|assert " | head
