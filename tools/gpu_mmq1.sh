# first GPU contact of the MMQ kernel: diagnostics first (bounded by timeout), then parity tests, timing, llama-bench pp512
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python tools/mmq_debug.py check > gpurun_out/mmq_check.log 2>&1; echo "exit $?" >> gpurun_out/mmq_check.log
tail -5 gpurun_out/mmq_check.log
timeout 600 python -m pytest tests/test_gpu_mmq.py -q -x > gpurun_out/mmq_pytest.log 2>&1; echo "exit $?" >> gpurun_out/mmq_pytest.log
tail -5 gpurun_out/mmq_pytest.log
timeout 300 python tools/mmq_debug.py time q4_K q6_K > gpurun_out/mmq_time.log 2>&1; echo "exit $?" >> gpurun_out/mmq_time.log
cat gpurun_out/mmq_time.log
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_mmq.py > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 600 $B/test-backend-ops test -b B200_0 -o MUL_MAT > gpurun_out/tbo_mul_mat.log 2>&1; echo "exit $?" >> gpurun_out/tbo_mul_mat.log
tail -4 gpurun_out/tbo_mul_mat.log
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 3 > gpurun_out/lb_8b_pp.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp.log
cat gpurun_out/lb_8b_pp.log | tail -6
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
cat gpurun_out/lb_8b_tg.log | tail -4
