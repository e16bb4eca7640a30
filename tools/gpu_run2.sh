export MICRO_XQ=1 MICRO_SHAPES="ffn_up/gate,ffn_down 4096x14336 q4_K,attn_q,lm_head"
for kb in 6 8 12 16 24 32; do for ns in 4 6 8; do
echo "== STAGE_KB=$kb NS=$ns"; B200_MMVQ_STAGE_KB=$kb B200_MMVQ_NS=$ns python tools/mmvq_micro.py 1 2>&1 | cut -c1-75
done; done > gpurun_out/sweep.log 2>&1
