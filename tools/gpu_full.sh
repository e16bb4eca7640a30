export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 128 -r 3 > gpurun_out/lb_8b.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b.log
grep -E "pp512|tg128|ggml-b200: (stats|cuda)|error|exit" gpurun_out/lb_8b.log
for f in 1 0; do B200_FUSE=$f timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/fuse$f.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse=$f exit $?"; done
cmp /tmp/fuse1.bin /tmp/fuse0.bin && echo "fused == unfused logits: bit-identical"
