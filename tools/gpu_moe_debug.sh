#!/bin/bash
# gpu_moe_debug.sh -- why the MoE tail fusion does / does not match on the Mixtral model (B200_DEBUG_FUSE lines of a 4-token decode)
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:${LD_LIBRARY_PATH:-}
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
M=$(python -c "import bench; print(bench.ensure_model('mixtral-8x7b', 'q4_k_m'))")
B200_DEBUG_FUSE=1 B200_TIMING=1 timeout 300 baseline/_ref/bin/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 4 -r 1 -t 4 2>&1 | grep -E "moe tail|stats|silu not fused" | sort | uniq -c | sort -rn | head -n 12
