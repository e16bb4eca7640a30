# grouped q/k/v MMQ of a prefill batch: unit test, logits vs reference CPU backend, fused vs unfused logits, prefill speed
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mmq.py -q -x 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -k "logits or tiny" 2>&1 | tail -3
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
L=tests/host/build/logits_check
for cfg in "600 4"; do
  set -- $cfg
  B200_FUSE_MMQ_QKV=0 timeout 600 $L $M /tmp/a.bin 99 1 $1 $2 > /dev/null 2>&1 || echo "run a failed"
  B200_FUSE_MMQ_QKV=1 timeout 600 $L $M /tmp/b.bin 99 1 $1 $2 > /dev/null 2>&1 || echo "run b failed"
  cmp -s /tmp/a.bin /tmp/b.bin && echo "qkv group == separate launches ($cfg)" || { echo "qkv group DIFFERS ($cfg)"; python tools/cmp_bins.py /tmp/a.bin /tmp/b.bin; }
done
B=baseline/_ref/bin
for v in 0 1 0 1; do
  echo "== B200_FUSE_MMQ_QKV=$v"
  B200_FUSE_MMQ_QKV=$v timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 5 2>&1 | grep -E "pp512|pp2048"
done
