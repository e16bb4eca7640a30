export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3.2-1b q4_k_m /tmp/l1b.gguf > gpurun_out/gen1b.log 2>&1
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
python tools/make_gguf.py llama3-70b q4_k_m /tmp/l70b.gguf > gpurun_out/gen70b.log 2>&1
L=tests/host/build/logits_check
timeout 300 $L /tmp/l1b.gguf /tmp/l1b.none.bin 99 1 40 8 512 none > /dev/null 2>&1
B200_SPLIT_P2P=0 timeout 300 $L /tmp/l1b.gguf /tmp/l1b.row_nop2p.bin 99 1 40 8 512 row > gpurun_out/split_l1b.row_nop2p.log 2>&1; echo "row without p2p exit $?"
python tools/cmp_logits.py /tmp/l1b.none.bin /tmp/l1b.row_nop2p.bin
B200_SPLIT_XCOPY=1 timeout 300 $L /tmp/l1b.gguf /tmp/l1b.row_xcopy.bin 99 1 40 8 512 row > gpurun_out/split_l1b.row_xcopy.log 2>&1; echo "row xcopy exit $?"
python tools/cmp_logits.py /tmp/l1b.none.bin /tmp/l1b.row_xcopy.bin
for x in 0 1; do for m in l8b-q4km l70b; do
B200_SPLIT_XCOPY=$x timeout 600 $B/llama-bench -m /tmp/$m.gguf -ngl 99 -fa 1 -sm row -p 0 -n 128 -r 2 > gpurun_out/lb_${m}_row_xcopy$x.log 2>&1
echo "xcopy=$x $m"; grep -E "tg128|error" gpurun_out/lb_${m}_row_xcopy$x.log
done; done
