# prefill glue A/B: rope heads kernel and vector cpy on / off; new unit tests; reference ROPE / CPY op tests
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fused.py -x -q -k "rope_prefill or cpy_vector" 2>&1 | tail -5
timeout 600 python -m pytest tests/test_gpu_fattn.py -x -q 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -k "ROPE or CPY or CONT or logits" 2>&1 | tail -3
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
for cfg in "B200_ROPE_HEADS=0 B200_CPY_VEC=1" "B200_ROPE_HEADS=1 B200_CPY_VEC=1" "B200_ROPE_HEADS=0 B200_CPY_VEC=1" "B200_ROPE_HEADS=1 B200_CPY_VEC=1"; do
  echo "== $cfg"
  env $cfg timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 5 2>&1 | grep -E "pp512|pp2048"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_pp.csv $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/ncu_pp.log 2>&1
wc -l gpurun_out/launches_pp.csv
