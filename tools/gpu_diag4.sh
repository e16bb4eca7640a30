mkdir -p gpurun_out
python - <<'PY' > gpurun_out/diag4.log 2>&1
import os, sys, subprocess
sys.path.insert(0, 'tools')
import logits_diag as L
model = "/tmp/diag-tiny-q4_0.gguf"
subprocess.check_call([sys.executable, "tools/make_gguf.py", "tiny", "q4_0", model], stdout=subprocess.DEVNULL)
for n_prompt in (16, 31, 32, 33, 48, 64, 65, 128, 129):
    ref = L.run(model, "/tmp/c.bin", 0, 1, n_prompt, 1)
    for name, env in [("default", {}), ("no-mmq", {"B200_NO_MMQ": "1"}), ("no-xb-cache", {"B200_NO_XB_CACHE": "1"}), ("no-fattn-tc", {"B200_FATTN_TC_MIN_ROWS": "100000"})]:
        got = L.run(model, "/tmp/g.bin", 99, 1, n_prompt, 1, env)
        print("tiny q4_0 n_prompt", n_prompt, name, " ".join("%.3g" % L.nmse(got[i], ref[i]) for i in range(got.shape[0])), flush=True)
    ref0 = L.run(model, "/tmp/c0.bin", 0, 0, n_prompt, 1)
    print("tiny q4_0 n_prompt", n_prompt, "cpu fa0 vs fa1", " ".join("%.3g" % L.nmse(ref0[i], ref[i]) for i in range(ref.shape[0])), flush=True)
PY
cat gpurun_out/diag4.log
