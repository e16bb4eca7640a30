export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_plugin.py -x -q ) > gpurun_out/plugin_pytest.log 2>&1; tail -6 gpurun_out/plugin_pytest.log
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
if [ -x tests/host/build/logits_check ]; then
  B200_FUSE_GEMV=1 timeout 600 tests/host/build/logits_check $M /tmp/fg1.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse_gemv=1 exit $?"
  B200_FUSE_GEMV=0 timeout 600 tests/host/build/logits_check $M /tmp/fg0.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse_gemv=0 exit $?"
  B200_FUSE=0 timeout 600 tests/host/build/logits_check $M /tmp/f0.bin 99 1 20 8 > /dev/null 2>&1; echo "logits fuse=0 exit $?"
  cmp /tmp/fg1.bin /tmp/fg0.bin && echo "gemv-fused == glue-fused logits: bit-identical"
  cmp /tmp/fg1.bin /tmp/f0.bin && echo "gemv-fused == unfused logits: bit-identical"
fi
B200_TIMING=1 timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 128 -r 3 2>&1 | grep -E "pp512|tg128|stats|graphs"
KT_N=16 bash tools/gpu_ktrace.sh
