# host time per decode step inside graph_compute (enqueue) against llama-bench's own per-token time; graph replay checks
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
B200_TIMING=1 timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128|stats|cuda graphs"
timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128"
L=tests/host/build/logits_check
B200_CUDA_GRAPHS=0 timeout 600 $L $M /tmp/a.bin 99 1 300 40 > /dev/null 2>&1 || echo "run a failed"
B200_CUDA_GRAPHS=1 timeout 600 $L $M /tmp/b.bin 99 1 300 40 > /dev/null 2>&1 || echo "run b failed"
cmp -s /tmp/a.bin /tmp/b.bin && echo "graph replay == eager (300 + 40 tokens, crosses a KV padding boundary)" || { echo "graph replay DIFFERS"; python tools/cmp_bins.py /tmp/a.bin /tmp/b.bin; }
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -k "logits or tiny or row_split" 2>&1 | tail -3
