# one launch for rope(q) / rope(k) + K store / V store of a prefill batch: unit test, logits identity, speed
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fused.py -q -x 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -k "logits or tiny" 2>&1 | tail -3
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
L=tests/host/build/logits_check
B200_FUSE=0 timeout 600 $L $M /tmp/a.bin 99 1 600 6 > /dev/null 2>&1 || echo "run a failed"
B200_FUSE=1 timeout 600 $L $M /tmp/b.bin 99 1 600 6 > /dev/null 2>&1 || echo "run b failed"
B200_FUSE_ROPE_BATCH=0 timeout 600 $L $M /tmp/c.bin 99 1 600 6 > /dev/null 2>&1 || echo "run c failed"
cmp -s /tmp/a.bin /tmp/b.bin && echo "all fusions == unfused" || { echo "fused DIFFERS from unfused"; python tools/cmp_bins.py /tmp/a.bin /tmp/b.bin; }
cmp -s /tmp/c.bin /tmp/b.bin && echo "rope batch == separate" || echo "rope batch DIFFERS"
B=baseline/_ref/bin
for v in 0 1; do
  echo "== B200_FUSE_ROPE_BATCH=$v"
  B200_FUSE_ROPE_BATCH=$v timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 12 -o json 2>/dev/null | python -c "
import json,sys
for t in json.load(sys.stdin): print(t['n_prompt'], sorted(round(x) for x in t['samples_ts'])[3:-1])"
done
