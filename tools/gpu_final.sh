# end-of-session evidence: full GPU validation, then ncu of the prefill (launch list of pp512, full capture of the MMQ kernel)
bash tools/gpu_validate.sh
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_pp.csv $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/ncu_pp.log 2>&1
wc -l gpurun_out/launches_pp.csv
timeout 300 python tools/prof_mmq.py up q down > gpurun_out/prof_mmq_plain.log 2>&1; tail -1 gpurun_out/prof_mmq_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mmq_kernel -c 9 -f -o gpurun_out/prof_mmq_v4 python tools/prof_mmq.py up q down > gpurun_out/prof_mmq_ncu.log 2>&1; tail -2 gpurun_out/prof_mmq_ncu.log
ncu -i gpurun_out/prof_mmq_v4.ncu-rep --page raw --csv > gpurun_out/prof_mmq_v4_raw.csv 2>/dev/null
ls -la gpurun_out/prof_mmq_v4.ncu-rep
