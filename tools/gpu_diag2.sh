mkdir -p gpurun_out
python - <<'PY' > gpurun_out/diag2.log 2>&1
import os, sys, subprocess
sys.path.insert(0, 'tools')
import logits_diag as L
for preset in ("tiny", "llama3.2-1b"):
    model = "/tmp/diag-%s-q4_0.gguf" % preset
    subprocess.check_call([sys.executable, "tools/make_gguf.py", preset, "q4_0", model], stdout=subprocess.DEVNULL)
    for n_prompt in (1, 8, 9, 64):
        ref = L.run(model, "/tmp/c.bin", 0, 1, n_prompt, 2)
        for name, env in [("default", {}), ("debug-sync", {"B200_DEBUG": "1"}), ("no-pdl", {"B200_PDL": "0", "B200_CUDA_GRAPHS": "0"})]:
            got = L.run(model, "/tmp/g.bin", 99, 1, n_prompt, 2, env)
            print(preset, "n_prompt", n_prompt, name, " ".join("%.3g" % L.nmse(got[i], ref[i]) for i in range(got.shape[0])), flush=True)
PY
cat gpurun_out/diag2.log
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so timeout 300 baseline/_ref/bin/test-backend-ops test -b B200_0 -o MUL_MAT -p "type_a=q4_0" 2>&1 | grep -v "not supported" | tail -30
