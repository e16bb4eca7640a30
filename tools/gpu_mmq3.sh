export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python tools/mmq_debug.py check > gpurun_out/mmq_check.log 2>&1; echo "exit $?" >> gpurun_out/mmq_check.log
grep -v "^ " gpurun_out/mmq_check.log | tail -40
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 600 $B/test-backend-ops test -b B200_0 -o MUL_MAT > gpurun_out/tbo_mul_mat.log 2>&1; echo "exit $?" >> gpurun_out/tbo_mul_mat.log
tail -4 gpurun_out/tbo_mul_mat.log
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 3 > gpurun_out/lb_8b_pp.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp.log
tail -5 gpurun_out/lb_8b_pp.log
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 2048 -n 0 -ub 2048 -b 2048 -r 2 > gpurun_out/lb_8b_pp2048.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp2048.log
tail -5 gpurun_out/lb_8b_pp2048.log
timeout 300 python tools/mmq_debug.py time q6_K q8_0 > gpurun_out/mmq_time2.log 2>&1; echo "exit $?" >> gpurun_out/mmq_time2.log
cat gpurun_out/mmq_time2.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_pp512.csv $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/ncu_pp512.log 2>&1; echo "exit $?" >> gpurun_out/ncu_pp512.log
tail -3 gpurun_out/ncu_pp512.log
