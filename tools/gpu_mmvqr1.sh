# register-resident GEMV: parity, then micro-benchmark sweep, then the bench line
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mmvq.py -x -q > gpurun_out/mmvqr_pytest.log 2>&1; tail -5 gpurun_out/mmvqr_pytest.log
export MICRO_XQ=2 MICRO_SHAPES="4096x4096 q4_K,1024x4096 q4_K,14336x4096 q4_K,4096x14336 q4_K,4096x14336 q6_K,lm_head"
( echo "== old kernel"; B200_MMVQR=0 timeout 300 python tools/mmvq_micro.py
for cps in 1 2; do for skb in 8 16 24; do
  echo "== MMVQR CPS=$cps STAGE_KB=$skb"; B200_MMVQR_CPS=$cps B200_MMVQR_STAGE_KB=$skb timeout 300 python tools/mmvq_micro.py
done; done
echo "== MMVQR CPS=2 SMEM_KB=56 STAGE_KB=12 (4 CTAs/SM smem-wise)"; B200_MMVQR_CPS=2 B200_MMVQR_SMEM_KB=56 B200_MMVQR_STAGE_KB=12 timeout 300 python tools/mmvq_micro.py
) > gpurun_out/mmvqr_micro.log 2>&1
cat gpurun_out/mmvqr_micro.log
