export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 32 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
CMD="baseline/_ref/bin/llama-bench -m /tmp/b200_models/llama3-8b-q4_k_m.gguf -ngl 99 -fa 1 -p 0 -n 3 -r 1"
$CMD > gpurun_out/plain_lb.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_lb.csv $CMD > gpurun_out/ncu_lb.log 2>&1
