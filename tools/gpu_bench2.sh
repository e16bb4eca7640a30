export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
( time python bench.py ) > gpurun_out/bench.json 2> gpurun_out/bench.err
tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json | head -c 3000; echo
python bench.py --impl reference --steps 32 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
cat gpurun_out/bench_ref.json | head -c 800; echo
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
# launch lists: decode (graphs off so that every kernel is listed as launched) and prefill
CMD="baseline/_ref/bin/llama-bench -m /tmp/b200_models/llama3-8b-q4_k_m.gguf -ngl 99 -fa 1 -p 0 -n 3 -r 1"
B200_CUDA_GRAPHS=0 $CMD > gpurun_out/plain_lb.log 2>&1 &&
B200_CUDA_GRAPHS=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_tg.csv $CMD > gpurun_out/ncu_lb.log 2>&1
CMD2="baseline/_ref/bin/llama-bench -m /tmp/b200_models/llama3-8b-q4_k_m.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 1"
$CMD2 > gpurun_out/plain_pp.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_pp512.csv $CMD2 > gpurun_out/ncu_pp.log 2>&1
# one full capture each of the two tensor-core kernels
unset GGML_BACKEND_PATH
python tools/prof_mmq.py up > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mmq_kernel -s 2 -c 1 -o gpurun_out/prof_mmq_v3 python tools/prof_mmq.py up > gpurun_out/ncu_mmq.log 2>&1
tail -2 gpurun_out/ncu_mmq.log
