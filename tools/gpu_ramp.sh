export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_mmvq.py -x -q > gpurun_out/fuse_pytest.log 2>&1; tail -3 gpurun_out/fuse_pytest.log
export MICRO_XQ=2 MICRO_SHAPES="4096x4096 q4_K,14336x4096 q4_K,4096x14336 q4_K,4096x14336 q6_K,lm_head"
( for ramp in 0 1; do for skb in 36 54; do
  echo "== MMVQR RAMP=$ramp STAGE_KB=$skb"; B200_MMVQR_RAMP=$ramp B200_MMVQR_STAGE_KB=$skb timeout 300 python tools/mmvq_micro.py
done; done ) > gpurun_out/mmvqr_micro4.log 2>&1
cat gpurun_out/mmvqr_micro4.log
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
for fg in 1 0; do for ramp in 1 0; do
echo "== FUSE_GEMV=$fg RAMP=$ramp"; B200_FUSE_GEMV=$fg B200_MMVQR_RAMP=$ramp timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128"
done; done
KT_N=16 bash tools/gpu_ktrace.sh
