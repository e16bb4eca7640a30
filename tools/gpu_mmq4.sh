export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
for preset in tiny tiny-moe; do
  python tools/make_gguf.py $preset q4_k_m /tmp/$preset.gguf > /dev/null 2>&1
  for fa in 0 1; do
    echo "== $preset fa=$fa"
    B200_DEBUG=1 timeout 120 $B/llama-bench -m /tmp/$preset.gguf -ngl 99 -fa $fa -p 48 -n 8 -r 1 2>&1 | grep -E "ggml-b200: op|pp48|tg8|error" | head -5
  done
done
unset GGML_BACKEND_PATH
timeout 900 python -m pytest tests/test_gpu_mmq.py -q > gpurun_out/mmq_pytest.log 2>&1; echo "exit $?" >> gpurun_out/mmq_pytest.log
tail -4 gpurun_out/mmq_pytest.log
python tools/prof_mmq.py up q > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mmq_kernel -s 2 -c 1 -o gpurun_out/prof_mmq_up python tools/prof_mmq.py up > gpurun_out/ncu_mmq.log 2>&1
tail -3 gpurun_out/ncu_mmq.log
