export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
export MICRO_XQ=2 MICRO_SHAPES="4096x4096 q4_K,14336x4096 q4_K,4096x14336 q4_K,4096x14336 q6_K"
( for skb in 28 36 54; do for ns in 8 2; do
  echo "== MMVQR STAGE_KB=$skb NS=$ns"; B200_MMVQR_NS=$ns B200_MMVQR_STAGE_KB=$skb timeout 300 python tools/mmvq_micro.py
done; done ) > gpurun_out/mmvqr_micro3.log 2>&1
cat gpurun_out/mmvqr_micro3.log
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "exit $?" >> gpurun_out/bench.err
tail -3 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
M=$(ls /tmp/*.gguf gpurun_out/*.gguf 2>/dev/null | head -1); echo "model: $M"
