export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
for pf in 1 0 1 0; do
B200_L2_PREFETCH=$pf B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_pf$pf.log 2>&1
echo "prefetch=$pf"; grep -E "tg128|error|stats|cuda graphs" gpurun_out/lb_8b_pf$pf.log
done
for pf in 1 0; do B200_L2_PREFETCH=$pf timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/pf$pf.bin 99 1 20 8 > /dev/null 2>&1; echo "logits pf=$pf exit $?"; done
cmp /tmp/pf1.bin /tmp/pf0.bin && echo "prefetch on == off logits: bit-identical"
python tools/make_gguf.py llama3-70b q4_k_m /tmp/l70b.gguf > gpurun_out/gen70b.log 2>&1
for pf in 1 0; do
B200_L2_PREFETCH=$pf timeout 600 $B/llama-bench -m /tmp/l70b.gguf -ngl 99 -fa 1 -p 0 -n 64 -r 2 > gpurun_out/lb_70b_pf$pf.log 2>&1
echo "70b prefetch=$pf"; grep -E "tg64|error" gpurun_out/lb_70b_pf$pf.log
done
