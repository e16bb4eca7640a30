#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_fattn.py -m gpu -x -q -k "batched_decode or fattn" 2>&1 | tail -4
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200 GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
M=$(python -c "import bench; print(bench.ensure_model())")
B200_TIMING=1 timeout 300 $B/llama-batched-bench -m $M -ngl 99 -fa -npp 128 -ntg 32 -npl 8,9,12,16,32,64 -c 16384 2>&1 | grep "^|  \|stats"
f=/tmp/b200_models/mixtral-8x7b-q4_k_m.gguf; mkdir -p /tmp/b200_models; python tools/make_gguf.py mixtral-8x7b q4_k_m $f > /dev/null
B200_TIMING=1 timeout 300 $B/llama-batched-bench -m $f -ngl 99 -fa -npp 128 -ntg 32 -npl 1,2,4,8,12,16,32,64 -c 16384 > gpurun_out/r02_batched_bench_mixtral.log 2>&1
grep "^|  \|stats" gpurun_out/r02_batched_bench_mixtral.log
timeout 100 python tools/fa_decode_micro.py q8_0 2>&1 | grep "32/8"
