export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
B200_TIMING=1 timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 512,2048 -n 0 -r 3 2>&1 | grep -E "pp512|pp2048|stats"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_pp.csv $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/ncu_pp.log 2>&1
wc -l gpurun_out/launches_pp.csv
