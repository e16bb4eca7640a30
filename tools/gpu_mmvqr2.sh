export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mmvq.py -x -q > gpurun_out/mmvqr_pytest.log 2>&1; tail -5 gpurun_out/mmvqr_pytest.log
export MICRO_XQ=2 MICRO_SHAPES="4096x4096 q4_K,1024x4096 q4_K,14336x4096 q4_K,4096x14336 q4_K,4096x14336 q6_K,lm_head"
( for u in 1 2; do for cps in 1 2; do for skb in 16 36; do
  echo "== MMVQR U=$u CPS=$cps STAGE_KB=$skb"; B200_MMVQR_U=$u B200_MMVQR_CPS=$cps B200_MMVQR_STAGE_KB=$skb timeout 300 python tools/mmvq_micro.py
done; done; done
echo "== MMVQR U=2 CPS=2 SMEM_KB=64 STAGE_KB=16"; B200_MMVQR_U=2 B200_MMVQR_CPS=2 B200_MMVQR_SMEM_KB=64 B200_MMVQR_STAGE_KB=16 timeout 300 python tools/mmvq_micro.py
) > gpurun_out/mmvqr_micro2.log 2>&1
cat gpurun_out/mmvqr_micro2.log
