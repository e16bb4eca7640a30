#!/bin/bash
# gpu_ab.sh VAR [llama-bench args] -- A/B of one plugin switch (VAR=1 / VAR=0, alternating, twice) on the bench model's decode:
# llama-bench tok/s and the replayed-graph device time per token (B200_TIMING=1 stats line). Run through gpurun.
# MODEL_FTYPE=q8_0 picks another quantisation, MODEL_PRESET=llama3.2-1b another model (tools/make_gguf.py).
set -u
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:${LD_LIBRARY_PATH:-}
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
var=$1; shift
FT=${MODEL_FTYPE:-q4_k_m}
M=$(python -c "import bench; print(bench.ensure_model('${MODEL_PRESET:-llama3-8b}', '$FT'))")
for v in 1 0 1 0; do
  echo "== $var=$v"
  env $var=$v B200_TIMING=1 timeout 300 baseline/_ref/bin/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 -t 4 "$@" 2>&1 | grep -E "tg128|replay|error|fail" | cut -c1-400
done
