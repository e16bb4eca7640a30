export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py -q -k "logits or llama_bench" > gpurun_out/plugin_pytest.log 2>&1; echo "exit $?" >> gpurun_out/plugin_pytest.log
grep -E "passed|failed|FAILED|Error|assert|nmse" gpurun_out/plugin_pytest.log | head -30
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_TIMING=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
grep -E "tg128|ggml-b200: (stats|cuda)|error|exit" gpurun_out/lb_8b_tg.log
B200_CUDA_GRAPHS=0 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128|error"
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 2 2>&1 | grep -E "tg128|error"
timeout 600 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/gpu8b.bin 99 1 100 8 > /dev/null 2>&1; echo "gpu logits exit $?"
GGML_BACKEND_PATH= timeout 900 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/cpu8b.bin 0 1 100 8 > /dev/null 2>&1; echo "cpu logits exit $?"
python - <<'PY'
import numpy as np
a=np.fromfile('/tmp/gpu8b.bin',dtype=np.uint8); b=np.fromfile('/tmp/cpu8b.bin',dtype=np.uint8)
nv=int(a[:4].view(np.int32)[0]); A=a[4:].view(np.float32).reshape(-1,nv); B=b[4:].view(np.float32).reshape(-1,nv)
for i in range(A.shape[0]):
    print("8B logits step %d nmse %.3g argmax gpu %d cpu %d" % (i, ((A[i].astype(np.float64)-B[i])**2).sum()/(B[i].astype(np.float64)**2).sum(), A[i].argmax(), B[i].argmax()))
PY
