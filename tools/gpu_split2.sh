# 2+ GPUs: row split with grouped fork/join and CUDA graphs; 70B
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3.2-1b q4_k_m /tmp/l1b.gguf > gpurun_out/gen1b.log 2>&1
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
L=tests/host/build/logits_check
for m in l1b l8b-q4km; do
  for sm in none row; do
    B200_TIMING=1 timeout 300 $L /tmp/$m.gguf /tmp/$m.$sm.bin 99 1 40 8 512 $sm > gpurun_out/split_$m.$sm.log 2>&1; echo "$m $sm exit $?"
    grep -E "row-split|cuda graphs|error|failed" gpurun_out/split_$m.$sm.log | head -5
  done
  python tools/cmp_logits.py /tmp/$m.none.bin /tmp/$m.row.bin
done
B200_SPLIT_P2P=0 timeout 300 $L /tmp/l1b.gguf /tmp/l1b.row_nop2p.bin 99 1 40 8 512 row > gpurun_out/split_l1b.row_nop2p.log 2>&1; echo "row without p2p exit $?"
python tools/cmp_logits.py /tmp/l1b.none.bin /tmp/l1b.row_nop2p.bin
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -sm row -p 512 -n 128 -r 3 > gpurun_out/lb_8b_sm_row.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_sm_row.log
grep -E "pp512|tg128|error|exit|cuda graphs" gpurun_out/lb_8b_sm_row.log
B200_SPLIT_GRAPHS=0 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -sm row -p 0 -n 128 -r 3 > gpurun_out/lb_8b_sm_row_nograph.log 2>&1
grep -E "tg128|error" gpurun_out/lb_8b_sm_row_nograph.log
B200_SPLIT_XCOPY=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -sm row -p 0 -n 128 -r 3 > gpurun_out/lb_8b_sm_row_xcopy.log 2>&1
grep -E "tg128|error" gpurun_out/lb_8b_sm_row_xcopy.log
( time python tools/make_gguf.py llama3-70b q4_k_m /tmp/l70b.gguf ) > gpurun_out/gen70b.log 2>&1; tail -4 gpurun_out/gen70b.log
for sm in none layer row; do
  timeout 900 $B/llama-bench -m /tmp/l70b.gguf -ngl 99 -fa 1 -sm $sm -p 512 -n 128 -r 2 > gpurun_out/lb_70b_sm_$sm.log 2>&1; echo "exit $?" >> gpurun_out/lb_70b_sm_$sm.log
  grep -E "pp512|tg128|error|exit" gpurun_out/lb_70b_sm_$sm.log
done
