export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3.2-1b q4_k_m /tmp/l1b.gguf > gpurun_out/gen1b.log 2>&1
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
python tools/make_gguf.py llama3-70b q4_k_m /tmp/l70b.gguf > gpurun_out/gen70b.log 2>&1
L=tests/host/build/logits_check
for m in l1b l8b-q4km; do
  timeout 300 $L /tmp/$m.gguf /tmp/$m.none.bin 99 1 40 8 512 none > /dev/null 2>&1
  B200_TIMING=1 timeout 120 $L /tmp/$m.gguf /tmp/$m.row.bin 99 1 40 8 512 row > gpurun_out/split_$m.row.log 2>&1; echo "$m row exit $?"
  grep -E "row-split|cuda graphs|error|failed" gpurun_out/split_$m.row.log | head -5
  python tools/cmp_logits.py /tmp/$m.none.bin /tmp/$m.row.bin
done
B200_CUDA_GRAPHS=0 timeout 120 $L /tmp/l1b.gguf /tmp/l1b.row_ng.bin 99 1 40 8 512 row > gpurun_out/split_l1b.row_ng.log 2>&1; echo "row no graphs exit $?"
python tools/cmp_logits.py /tmp/l1b.none.bin /tmp/l1b.row_ng.bin
for m in l8b-q4km l70b; do
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/$m.gguf -ngl 99 -fa 1 -sm row -p 512 -n 128 -r 3 > gpurun_out/lb_${m}_row_flags.log 2>&1
echo "flags $m"; grep -E "pp512|tg128|error|cuda graphs" gpurun_out/lb_${m}_row_flags.log
done
B200_CUDA_GRAPHS=0 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -sm row -p 0 -n 128 -r 3 > gpurun_out/lb_8b_row_flags_nograph.log 2>&1
grep -E "tg128|error" gpurun_out/lb_8b_row_flags_nograph.log
