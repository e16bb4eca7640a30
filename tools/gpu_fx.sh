export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
timeout 900 python -m pytest tests/test_gpu_fattn.py -x -q > gpurun_out/fa_pytest.log 2>&1; tail -3 gpurun_out/fa_pytest.log
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 128 -r 3 > gpurun_out/lb_8b.log 2>&1
grep -E "pp512|tg128|error|stats|cuda graphs" gpurun_out/lb_8b.log
for f in 1 0; do B200_FUSE=$f timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/fuse$f.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse=$f exit $?"; done
cmp /tmp/fuse1.bin /tmp/fuse0.bin && echo "fused == unfused logits: bit-identical"
timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/long1.bin 99 1 1200 6 > /dev/null 2>&1; echo "long ctx exit $?"
B200_FUSE=0 timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/long0.bin 99 1 1200 6 > /dev/null 2>&1; echo "long ctx unfused exit $?"
cmp /tmp/long1.bin /tmp/long0.bin && echo "long ctx fused == unfused: bit-identical"
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 64 -d 4096,32768 -r 2 > gpurun_out/lb_8b_depth.log 2>&1; grep -E "tg64|error" gpurun_out/lb_8b_depth.log
