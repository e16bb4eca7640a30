# run-to-run spread of pp512 inside one llama-bench process (chronological samples)
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 12 -o json 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
print('pp512 chronological:', [round(x) for x in d[-1]['samples_ts']])"
B200_TIMING=1 timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 12 2>&1 | grep -E "pp512|stats"
