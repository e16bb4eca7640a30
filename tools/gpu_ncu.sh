# ncu evidence for profiles/: (1) full capture of the batch-1 GEMV on the four Q4_K shapes of a layer, (2) launch list of a decode step
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
MICRO_XQ=2 timeout 300 python tools/prof_mmvq.py q4k kv4k up4k down4k > gpurun_out/prof_plain.log 2>&1; tail -1 gpurun_out/prof_plain.log
MICRO_XQ=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mmvqr -c 12 -f -o gpurun_out/prof_mmvqr python tools/prof_mmvq.py q4k kv4k up4k down4k > gpurun_out/prof_ncu.log 2>&1; tail -2 gpurun_out/prof_ncu.log
ncu -i gpurun_out/prof_mmvqr.ncu-rep --page raw --csv > gpurun_out/prof_mmvqr_raw.csv 2>/dev/null
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
B200_CUDA_GRAPHS=0 timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 3 -r 1 > gpurun_out/lb_nograph.log 2>&1; tail -3 gpurun_out/lb_nograph.log | head -1
B200_CUDA_GRAPHS=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_tg.csv $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 3 -r 1 > gpurun_out/ncu_lb.log 2>&1
wc -l gpurun_out/launches_tg.csv
ls -la gpurun_out/prof_mmvqr.ncu-rep
