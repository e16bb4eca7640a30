#!/bin/bash
# row-split debug: llama-bench -sm row, stderr kept; B200_TIMING=1 prints the graph capture / replay counts
export LD_LIBRARY_PATH=baseline/_ref/lib:llamacog_b200
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
mkdir -p /tmp/b200_models
for m in ${MODELS:-llama3-8b}; do
  [ -f /tmp/b200_models/$m-q4_k_m.gguf ] || python tools/make_gguf.py $m q4_k_m /tmp/b200_models/$m-q4_k_m.gguf > /dev/null
  echo "=== $m row n=5,32 (B200_TIMING=1)"
  B200_TIMING=1 baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm row -ts ${TS:-1/1} -p 0 -n 5,32 -t 4 2>&1 | grep -v "^$" | tail -n 14
  echo "=== $m row n=64"
  baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm row -ts ${TS:-1/1} -p 0 -n 64 -t 4 2>&1 | grep "tg64\|error\|ggml-b200:"
  echo "=== $m row, events instead of flags"
  B200_SPLIT_FLAGS=0 baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm row -ts ${TS:-1/1} -p 0 -n 64 -t 4 2>&1 | grep "tg64\|error\|ggml-b200:"
  echo "=== $m row, unfolded flags"
  B200_SPLIT_FOLD=0 baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm row -ts ${TS:-1/1} -p 0 -n 64 -t 4 2>&1 | grep "tg64\|error\|ggml-b200:"
  echo "=== $m row, shard everything"
  B200_SPLIT_MIN_MB=0 baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm row -ts ${TS:-1/1} -p 0 -n 64 -t 4 2>&1 | grep "tg64\|error\|ggml-b200:"
  echo "=== $m none n=64"
  baseline/_ref/bin/llama-bench -m /tmp/b200_models/$m-q4_k_m.gguf -ngl 99 -fa 1 -r 1 -mg 0 -sm none -p 0 -n 64 -t 4 2>&1 | grep "tg64\|error"
done
