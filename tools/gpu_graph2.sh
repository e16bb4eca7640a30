export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
timeout 600 python tools/mmq_repro.py q4_K q5_K 2>&1 | grep -E "2048 K= 2048 M= 200|8192 M=" 
timeout 600 python -m pytest tests/test_gpu_mmq.py -q -x 2>&1 | tail -2
B=baseline/_ref/bin
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > gpurun_out/gen8b.log 2>&1
B200_TIMING=1 B200_DEBUG_GRAPH=1 timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 > gpurun_out/lb_8b_tg.log 2>&1; echo "exit $?" >> gpurun_out/lb_8b_tg.log
grep -E "tg128|ggml-b200: (stats|cuda)|error|exit" gpurun_out/lb_8b_tg.log; grep "signature differs" gpurun_out/lb_8b_tg.log | sort | uniq -c | head
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 1 -p 0 -n 128 -r 3 2>&1 | grep -E "tg128|error"
timeout 600 $B/llama-bench -m /tmp/l8b-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 2 2>&1 | grep -E "tg128|error"
unset GGML_BACKEND_PATH
timeout 900 python -m pytest tests/test_gpu_plugin.py -q -k "logits" 2>&1 | grep -E "passed|failed|FAILED|assert " | head
python - <<'PY'
import sys, subprocess
sys.path.insert(0, 'tools')
import logits_diag as L
model = "/tmp/diag-1b-q4km.gguf"
subprocess.check_call([sys.executable, "tools/make_gguf.py", "llama3.2-1b", "q4_k_m", model], stdout=subprocess.DEVNULL)
ref = L.run(model, "/tmp/c.bin", 0, 1, 64, 2)
a = L.run(model, "/tmp/g1.bin", 99, 1, 64, 2)
b = L.run(model, "/tmp/g2.bin", 99, 1, 64, 2, {"B200_NO_MMQ": "1"})
for i in range(3):
    print("1b q4_k_m step %d: mmq vs cpu %.3g | gemv-passes vs cpu %.3g | mmq vs gemv-passes %.3g" % (i, L.nmse(a[i], ref[i]), L.nmse(b[i], ref[i]), L.nmse(a[i], b[i])))
PY
