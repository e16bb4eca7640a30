"""smoke() hook: a 2-layer random Llama GGUF through unmodified llama-bench over the plugin on cuda:0
(decode + prefill, flash attention on).  Fails loudly if the plugin or the host tools are missing."""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
bin_dir = os.path.join(ROOT, "baseline", "_ref", "bin")
plugin = os.path.join(ROOT, "llamacog_b200", "libggml-b200.so")
if not os.path.exists(plugin) or not os.path.exists(os.path.join(bin_dir, "llama-bench")):
    raise RuntimeError("libggml-b200.so or baseline/_ref missing: run __graft_entry__.build() in the build container")
env = dict(os.environ)
env["LD_LIBRARY_PATH"] = os.path.join(ROOT, "baseline", "_ref", "lib") + ":" + os.path.join(ROOT, "llamacog_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
env["GGML_BACKEND_PATH"] = plugin
with tempfile.TemporaryDirectory() as d:
    model = os.path.join(d, "tiny.gguf")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), "tiny", "q4_k_m", model], stdout=subprocess.DEVNULL)
    r = subprocess.run([os.path.join(bin_dir, "llama-bench"), "-m", model, "-ngl", "99", "-fa", "1", "-p", "32", "-n", "8", "-r", "1", "-o", "json"],
                       capture_output=True, text=True, env=env, timeout=600)
    if r.returncode != 0:
        raise RuntimeError("llama-bench over the plugin failed: " + r.stderr[-2000:])
    res = json.loads(r.stdout[r.stdout.index("["):])
    assert all("B200" in x["backends"] and x["avg_ts"] > 0 for x in res), res
    print("plugin smoke ok:", [(x["n_prompt"], x["n_gen"], round(x["avg_ts"], 1)) for x in res])
