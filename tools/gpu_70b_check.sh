#!/bin/bash
# gpu_70b_check.sh -- Llama-3-70B Q4_K_M on ONE GPU: fused == unfused logits (20 + 4 tokens) and tg128 / pp512 with the gate/up pair mode on and off
set -u
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:${LD_LIBRARY_PATH:-}
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
M=$(python -c "import bench; print(bench.ensure_model('llama3-70b', 'q4_k_m'))")
L=tests/host/build/logits_check
B200_FUSE=0 timeout 600 $L $M /tmp/base70.bin 99 1 20 4 > /dev/null 2>&1 || echo "base failed"
B200_FUSE=1 timeout 600 $L $M /tmp/fused70.bin 99 1 20 4 > /dev/null 2>&1 || echo "fused failed"
cmp -s /tmp/base70.bin /tmp/fused70.bin && echo "70B: fused == unfused" || { echo "70B: fused DIFFERS from unfused"; python tools/cmp_bins.py /tmp/base70.bin /tmp/fused70.bin; }
for v in 1 0; do
  echo "== B200_FUSE_GEMV_SILU=$v"
  B200_FUSE_GEMV_SILU=$v B200_TIMING=1 timeout 600 baseline/_ref/bin/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 128 -r 2 -t 4 2>&1 | grep -E "pp512|tg128|replay|error|fail" | cut -c1-300
done
