#!/bin/bash
# model load time through the plugin: llama-bench -p 0 -n 1 wall clock, staged pinned upload vs the plain pageable copy
export LD_LIBRARY_PATH=baseline/_ref/lib:llamacog_b200
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
mkdir -p /tmp/b200_models
m=${MODEL:-llama3-8b}
f=/tmp/b200_models/$m-q4_k_m.gguf
[ -f $f ] || python tools/make_gguf.py $m q4_k_m $f > /dev/null
ls -la $f
for rep in 1 2; do
  for mode in 1 0; do
    s=$(date +%s%N)
    B200_STAGED_UPLOAD=$mode baseline/_ref/bin/llama-bench -m $f -ngl 99 -fa 1 -p 0 -n 1 -r 1 -t 4 > /dev/null 2>&1
    e=$(date +%s%N)
    echo "rep $rep B200_STAGED_UPLOAD=$mode : $(( (e - s) / 1000000 )) ms wall (process start, load, 2 tokens)"
  done
done
