# bf16 rows out of the tensor-core attention kernel: unit test, reference op tests, logits identity, prefill speed
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fattn.py -q -x 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -k "logits or tiny" 2>&1 | tail -3
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
L=tests/host/build/logits_check
B200_FUSE_ATTN_XB=0 timeout 600 $L $M /tmp/a.bin 99 1 600 4 > /dev/null 2>&1 || echo "run a failed"
B200_FUSE_ATTN_XB=1 timeout 600 $L $M /tmp/b.bin 99 1 600 4 > /dev/null 2>&1 || echo "run b failed"
B200_FUSE=0 timeout 600 $L $M /tmp/c.bin 99 1 600 4 > /dev/null 2>&1 || echo "run c failed"
cmp -s /tmp/a.bin /tmp/b.bin && echo "attn xb == separate conversion" || { echo "attn xb DIFFERS"; python tools/cmp_bins.py /tmp/a.bin /tmp/b.bin; }
cmp -s /tmp/c.bin /tmp/b.bin && echo "all fusions == unfused" || { echo "fused DIFFERS from unfused"; python tools/cmp_bins.py /tmp/c.bin /tmp/b.bin; }
B=baseline/_ref/bin
for v in 0 1 0 1; do
  echo "== B200_FUSE_ATTN_XB=$v"
  B200_FUSE_ATTN_XB=$v timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 5 2>&1 | grep -E "pp512|pp2048"
done
