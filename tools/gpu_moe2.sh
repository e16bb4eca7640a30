export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe.py -x -q > gpurun_out/moe_pytest.log 2>&1; tail -3 gpurun_out/moe_pytest.log
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py mixtral-8x7b q4_k_m /tmp/mix.gguf > gpurun_out/genmix.log 2>&1
CMD="baseline/_ref/bin/llama-bench -m /tmp/mix.gguf -ngl 99 -fa 1 -p 0 -n 3 -r 1"
B200_CUDA_GRAPHS=0 $CMD > gpurun_out/plain_mix.log 2>&1 &&
B200_CUDA_GRAPHS=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_mix_tg.csv $CMD > gpurun_out/ncu_mix.log 2>&1
tail -2 gpurun_out/ncu_mix.log
