export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_fused.py -x -q ) > gpurun_out/plugin_pytest.log 2>&1; tail -4 gpurun_out/plugin_pytest.log
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
B200_FUSE_ROPE=1 timeout 600 tests/host/build/logits_check $M /tmp/fr1.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse_rope=1 exit $?"
B200_FUSE_ROPE=0 timeout 600 tests/host/build/logits_check $M /tmp/fr0.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse_rope=0 exit $?"
B200_FUSE=0 timeout 600 tests/host/build/logits_check $M /tmp/f0.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse=0 exit $?"
cmp /tmp/fr1.bin /tmp/fr0.bin && echo "rope-in-GEMV == rope kernel logits: bit-identical"
cmp /tmp/fr1.bin /tmp/f0.bin && echo "rope-in-GEMV == unfused logits: bit-identical"
for fr in 1 0; do
echo "== FUSE_ROPE=$fr"; B200_TIMING=1 B200_FUSE_ROPE=$fr timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128|ropes|stats"
done
