export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3.2-1b q4_k_m /tmp/l1b.gguf > gpurun_out/gen1b.log 2>&1
L=tests/host/build/logits_check
B200_DEBUG=1 timeout 300 $L /tmp/l1b.gguf /tmp/l1b.row.bin 99 1 40 2 512 row > gpurun_out/split_dbg.log 2>&1; echo "dbg exit $?"
grep -E "ggml-b200" gpurun_out/split_dbg.log | head -20


