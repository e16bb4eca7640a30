export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
B200_KTRACE=/tmp/ktrace.txt timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 64 -r 1 > gpurun_out/kt_lb.log 2>&1; tail -3 gpurun_out/kt_lb.log | head -1
python tools/ktrace_report.py /tmp/ktrace.txt 0 ${KT_N:-20} > gpurun_out/ktrace_report.txt 2>&1; cat gpurun_out/ktrace_report.txt
