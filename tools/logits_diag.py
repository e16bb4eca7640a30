#!/usr/bin/env python
"""End-to-end logits parity diagnostics: reference CPU backend vs the B200 plugin over presets / file types /
kernel-selection switches (test infrastructure; uses tests/host/build/logits_check)."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "host", "build", "logits_check")


def run(model, out, ngl, fa, n_prompt, n_decode, extra_env=None):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.join(ROOT, "baseline", "_ref", "lib") + ":" + os.path.join(ROOT, "llamacog_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
    if ngl > 0:
        env["GGML_BACKEND_PATH"] = os.path.join(ROOT, "llamacog_b200", "libggml-b200.so")
    else:
        env.pop("GGML_BACKEND_PATH", None)
    env.update(extra_env or {})
    r = subprocess.run([EXE, model, out, str(ngl), str(fa), str(n_prompt), str(n_decode)], capture_output=True, text=True, env=env)
    if r.returncode != 0:
        print(r.stderr[-1500:]); raise SystemExit(1)
    raw = np.fromfile(out, dtype=np.uint8)
    nv = int(raw[:4].view(np.int32)[0])
    return raw[4:].view(np.float32).reshape(-1, nv)


def nmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / (b.astype(np.float64) ** 2).sum())


def main():
    preset = sys.argv[1] if len(sys.argv) > 1 else "llama3.2-1b"
    ftypes = sys.argv[2].split(",") if len(sys.argv) > 2 else ["q8_0", "q4_0", "q4_k_m", "f16"]
    n_prompt, n_decode = 64, 3
    for ft in ftypes:
        model = "/tmp/diag-%s-%s.gguf" % (preset, ft)
        if not os.path.exists(model):
            subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), preset, ft, model], stdout=subprocess.DEVNULL)
        ref = run(model, "/tmp/diag_cpu.bin", 0, 1, n_prompt, n_decode)
        for name, env in [("default", {}), ("no-mmq", {"B200_MMQ_MIN_M": "100000"}), ("no-fattn-tc", {"B200_FATTN_TC_MIN_ROWS": "100000"}),
                          ("no-mmq,no-fattn-tc", {"B200_MMQ_MIN_M": "100000", "B200_FATTN_TC_MIN_ROWS": "100000"}), ("no-graphs", {"B200_CUDA_GRAPHS": "0"})]:
            got = run(model, "/tmp/diag_gpu.bin", 99, 1, n_prompt, n_decode, env)
            print("%-12s %-8s %-20s prefill nmse %.3g   decode nmse %s" % (preset, ft, name, nmse(got[0], ref[0]),
                  " ".join("%.3g" % nmse(got[i], ref[i]) for i in range(1, got.shape[0]))), flush=True)
        cpu2 = run(model, "/tmp/diag_cpu2.bin", 0, 0, n_prompt, n_decode)
        print("%-12s %-8s %-20s prefill nmse %.3g   decode nmse %s" % (preset, ft, "cpu fa0 vs cpu fa1", nmse(cpu2[0], ref[0]),
              " ".join("%.3g" % nmse(cpu2[i], ref[i]) for i in range(1, cpu2.shape[0]))), flush=True)


if __name__ == "__main__":
    main()
