export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
for fg in 0 1; do for sg in 0 1; do
echo "== FUSE_GEMV=$fg STAGGER=$sg"; B200_FUSE_GEMV=$fg B200_MMVQR_STAGGER=$sg timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128"
done; done
for ns in 2 3; do for skb in 18 27 36; do
echo "== NS=$ns STAGE_KB=$skb"; B200_MMVQR_NS=$ns B200_MMVQR_STAGE_KB=$skb timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128"
done; done
