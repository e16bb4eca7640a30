#!/bin/bash
# The reference's OWN CUDA backend (ggml/src/ggml-cuda, unmodified, built by `make -C baseline cuda` for sm_100) on the same box
# and the same GGUFs as our plugin: context numbers, not a target.
export LD_LIBRARY_PATH=baseline/_ref/lib:llamacog_b200
mkdir -p /tmp/b200_models
for m in ${MODELS:-llama3-8b}; do
  f=/tmp/b200_models/$m-${FT:-q4_k_m}.gguf
  [ -f $f ] || python tools/make_gguf.py $m ${FT:-q4_k_m} $f > /dev/null
  echo "=== reference ggml-cuda (sm_100 build): $m ${FT:-q4_k_m}"
  GGML_BACKEND_PATH=$PWD/baseline/_ref/lib/libggml-cuda.so baseline/_ref/bin/llama-bench -m $f -ngl 99 -fa 1 -p 512 -n 128 -r 3 -t 4 ${EXTRA} 2>&1 | grep -v "^$" | tail -n 8
  echo "=== ggml-b200 plugin: $m ${FT:-q4_k_m}"
  GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so baseline/_ref/bin/llama-bench -m $f -ngl 99 -fa 1 -p 512 -n 128 -r 3 -t 4 ${EXTRA} 2>&1 | grep -v "^$" | tail -n 6
done
