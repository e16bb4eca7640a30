export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fattn.py -q > gpurun_out/fa_pytest.log 2>&1; echo "exit $?" >> gpurun_out/fa_pytest.log
grep -E "passed|failed|FAILED|Error|assert " gpurun_out/fa_pytest.log | head -30
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
timeout 900 $B/test-backend-ops test -b B200_0 -o FLASH_ATTN_EXT > gpurun_out/tbo_fa.log 2>&1; echo "exit $?" >> gpurun_out/tbo_fa.log
grep -c "OK" gpurun_out/tbo_fa.log; grep "FAIL" gpurun_out/tbo_fa.log | head -5; tail -3 gpurun_out/tbo_fa.log
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 512 -n 0 -r 3 > gpurun_out/lb_8b_pp.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp.log
tail -4 gpurun_out/lb_8b_pp.log
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 2048 -n 0 -ub 2048 -b 2048 -r 2 > gpurun_out/lb_8b_pp2048.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_pp2048.log
tail -4 gpurun_out/lb_8b_pp2048.log
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
tail -4 gpurun_out/lb_8b_tg.log
unset GGML_BACKEND_PATH
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
