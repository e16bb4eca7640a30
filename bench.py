#!/usr/bin/env python
"""bench.py -- headline measurement of the hot path.

Workload (BASELINE.json configs[1]): Llama-3-8B Q4_K_M, batch-1 decode (tg), flash attention on, f16 KV cache.
One STEP = one decoded token = one pass of the whole quantised forward graph (32 layers + lm_head, 4.616 GB of
Q4_K/Q6_K weights streamed from HBM, ~420 kernels) through the reference's own front end: the UNMODIFIED
llama-bench of the reference (baseline/_ref/bin) loads the plugin libggml-b200.so through GGML_BACKEND_PATH
(ggml/src/ggml-backend-reg.cpp:586-590) and drives libllama -> ggml_backend_sched -> our graph_compute.
The model is a random-init GGUF of that architecture written by tools/make_gguf.py (reference quantiser).

  e2e     llama-bench's own tokens/s: every step copies token id / positions / mask from host memory, runs the
          graph on the GPU and reads the 128256 logits back to the host (the call a user makes)
  value   tokens / device time of the graph_compute calls of the same run, taken with CUDA events on the backend's
          stream (B200_TIMING=1 makes the plugin accumulate them); inputs are resident in HBM at that point
  roofline  the dominant kernel, mmvqr_kernel<Q4_K> (mmvq_r.cuh): algorithmic bytes (N*K*144/256 per matrix) over its mean launch
          duration, timed live in this process with CUDA events on a CUDA graph holding exactly the 192 Q4_K
          GEMVs of one token (distinct weight matrices, 3.36 GB > L2), through the C ABI of include/b200_ops.h
  cpu_baseline  the reference CPU backend on the same GGUF (llama-bench -ngl 0, all host threads), bounded to tg16

  value   tokens / device time of the REPLAYED graph_compute calls of the same workload (CUDA events on the backend's
          stream; the plugin keeps the first eager run and the capture + instantiate call of a graph out of this sum)

`--impl reference` times the reference's own CPU implementation (same llama-bench, no plugin) on the same model.

N > 1 (torchrun, one rank per GPU): the path shards by WEIGHT ROWS (--split-mode row): rank 0, with all N GPUs
visible, runs the north-star multi-GPU workload -- Llama-3-70B Q4_K_M through unmodified llama-bench with -sm row -ts 1/../1
(and -sm layer, and -sm none on one GPU as the reference point) -- after checking IN THE SAME RUN that split and single-GPU
logits of the 1B model are bit-identical (tests/host/build/logits_check); the other ranks take part in the replica side
measurement and then wait at the barrier (libllama drives all GPUs of a split model from one process: the ggml backend ABI has
one host thread per model, not one per GPU).  value / e2e = 70B row-split tg tokens/s, scaling "strong".
"""
import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

REF_BIN = os.path.join(ROOT, "baseline", "_ref", "bin")
REF_LIB = os.path.join(ROOT, "baseline", "_ref", "lib")
PLUGIN = os.path.join(ROOT, "llamacog_b200", "libggml-b200.so")
MODEL_DIR = os.environ.get("B200_MODEL_DIR", "/tmp/b200_models")
METRIC = "llama-bench tg tokens/s, Llama-3-8B Q4_K_M (batch-1 decode)"

N_EMBD, N_FF, N_LAYER, N_KV, N_VOCAB = 4096, 14336, 32, 1024, 128256
Q4_K, Q6_K = 12, 14
BPW = {Q4_K: 144 / 256, Q6_K: 210 / 256}


def use_more_bits(i, n):   # reference src/llama-quant.cpp:185-187
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def step_plan():
    """[(name, type, N, K, input, output)] of the quantised MUL_MATs of one decode token, Q4_K_M mix."""
    plan = []
    for il in range(N_LAYER):
        more = use_more_bits(il, N_LAYER)
        plan += [("wq", Q4_K, N_EMBD, N_EMBD, "h", "q"), ("wk", Q4_K, N_KV, N_EMBD, "h", "k"),
                 ("wv", Q6_K if more else Q4_K, N_KV, N_EMBD, "h", "v"), ("wo", Q4_K, N_EMBD, N_EMBD, "q", "h2"),
                 ("gate", Q4_K, N_FF, N_EMBD, "h2", "g"), ("up", Q4_K, N_FF, N_EMBD, "h2", "u"),
                 ("down", Q6_K if more else Q4_K, N_EMBD, N_FF, "g", "h")]
    plan.append(("output", Q6_K, N_VOCAB, N_EMBD, "h", "logits"))
    return plan


def plan_bytes(plan, only=None):
    return sum(int(N * K * BPW[t]) for _, t, N, K, _, _ in plan if only is None or t == only)


class ClockSampler(threading.Thread):
    """Polls SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.stop_flag, self.ok = [], False, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.max_mhz = None

    def run(self):
        while self.ok and not self.stop_flag:
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                try:
                    rs = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                util = self.nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                self.samples.append((time.time(), mhz, rs, util))
            except Exception:
                pass
            time.sleep(0.01)

    def summary(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        loaded = [s for s in inside if s[3] > 0] or inside or self.samples[-3:]
        mhz = sorted(s[1] for s in loaded)
        bits = 0
        for s in loaded:
            bits |= s[2]
        names = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
                 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
                 0x100: "display_clock_setting"}
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz,
                "reasons": [n for b, n in names.items() if bits & b], "samples": len(loaded)}


def measured_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def tc_issue_peaks():
    """tcgen05.mma issue-rate peaks (bf16 / int8 / fp8): measured live with tools/tc_peak.bin when it is there (0.1 s), else the
    committed measurement of the same program on this pool's B200 (profiles/r02_tc_peak.log)."""
    def parse(out):
        return json.loads([l for l in out.split("\n") if l.startswith('{"tc_peak"')][-1])["tc_peak"]
    try:
        d = parse(subprocess.run([os.path.join(ROOT, "tools", "tc_peak.bin")], capture_output=True, text=True, timeout=120).stdout)
        d["source"] = "measured in this run (tools/tc_peak.bin)"
        return d
    except Exception:  # noqa: BLE001
        pass
    try:
        d = parse(open(os.path.join(ROOT, "profiles", "r02_tc_peak.log")).read())
        d["source"] = "profiles/r02_tc_peak.log"
        return d
    except Exception:  # noqa: BLE001
        return None


def ensure_model(preset="llama3-8b", ftype="q4_k_m"):
    os.makedirs(MODEL_DIR, exist_ok=True)
    path = os.path.join(MODEL_DIR, "%s-%s.gguf" % (preset, ftype))
    if not os.path.exists(path):
        tmp = path + ".tmp.%d" % os.getpid()
        subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), preset, ftype, tmp],
                              stdout=subprocess.DEVNULL)
        os.replace(tmp, path)
    return path


GPU_ARM_THREADS = 4      # host threads of a GPU-arm llama-bench (-t): libllama's CPU work per token is tiny; the default (all cores) makes replicas fight


def llama_bench(model, extra, plugin, timing=False, device=None, timeout=3600):
    """Run the reference's llama-bench; returns (list of json results, stats dict from the plugin)."""
    exe = os.path.join(REF_BIN, "llama-bench")
    if not os.path.exists(exe):
        raise RuntimeError("baseline/_ref/bin/llama-bench is missing: run __graft_entry__.build() where /root/reference exists")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = REF_LIB + ":" + os.path.join(ROOT, "llamacog_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
    env.pop("GGML_BACKEND_PATH", None)
    if plugin:
        if not os.path.exists(PLUGIN):
            raise RuntimeError("llamacog_b200/libggml-b200.so is not built; there is no CPU fallback")
        env["GGML_BACKEND_PATH"] = PLUGIN
        if timing:
            env["B200_TIMING"] = "1"
        if device is not None:
            env["CUDA_VISIBLE_DEVICES"] = str(device)
    if plugin and "-t" not in extra:
        extra = extra + ["-t", str(GPU_ARM_THREADS)]
    r = subprocess.run([exe, "-m", model, "-o", "json"] + extra, capture_output=True, text=True, env=env, timeout=timeout)
    logdir = os.environ.get("B200_BENCH_LOGDIR")            # keep the raw llama-bench output (profiles/ evidence)
    if logdir and int(os.environ.get("RANK", "0")) == 0:
        os.makedirs(logdir, exist_ok=True)
        with open(os.path.join(logdir, "llama_bench_calls.log"), "a") as f:
            f.write("$ %s%s\n%s\n--- stderr (tail) ---\n%s\n\n" % ("GGML_BACKEND_PATH=libggml-b200.so " if plugin else "", " ".join([os.path.basename(exe), "-m", os.path.basename(model), "-o", "json"] + extra),
                                                                 r.stdout[-6000:], r.stderr[-1500:]))
    if r.returncode != 0:
        raise RuntimeError("llama-bench failed (%d): %s" % (r.returncode, r.stderr[-2000:]))
    res = json.loads(r.stdout[r.stdout.index("["):])
    stats = {}
    m = re.search(r"ggml-b200: stats graphs=(\d+) device_ms=([0-9.]+) kernel_launches=(\d+) host_enqueue_ms=([0-9.]+) replay_graphs=(\d+) replay_ms=([0-9.]+)", r.stderr)
    if m:
        stats = {"graphs": int(m.group(1)), "device_ms": float(m.group(2)), "launches": int(m.group(3)), "host_ms": float(m.group(4)),
                 "replay_graphs": int(m.group(5)), "replay_ms": float(m.group(6))}
    stats["stderr_tail"] = r.stderr[-1500:]
    if plugin and "B200" not in r.stdout + r.stderr:
        raise RuntimeError("llama-bench did not load the B200 backend: " + r.stderr[-1000:])
    return res, stats


def issue_groups(plan):
    """The Q4_K mat-vecs of a token as the plugin issues them (ggml-b200-compute.inc: try_attn_frontend / try_group_gemv): consecutive
    matrices of one type over the same activations share a launch (q/k/v -- v only when it is Q4_K too --, gate/up); wo and down
    are launches of their own."""
    groups = []
    for e in plan:
        if groups and groups[-1][-1][4] == e[4] and groups[-1][-1][3] == e[3] and len(groups[-1]) < 3 and groups[-1][-1][0] != "wo":
            groups[-1].append(e)
        else:
            groups.append([e])
    return groups


def kernel_roofline(steps, warmup, dev_index, grouped=True):
    """Dominant kernel timed live: the Q4_K GEMVs of one token from a CUDA graph, CUDA events on its stream.  grouped: the launches
    the plugin issues in the timed region (128: q/k/v and gate/up share a launch); else every matrix as a launch of its own (192,
    the round-1 form of this figure)."""
    import ctypes as C
    import torch
    from llamacog_b200 import ops
    from tools.mmvq_micro import rand_weights
    torch.cuda.set_device(dev_index)
    dev = torch.device("cuda", dev_index)
    L = ops.lib()
    plan = [p for p in step_plan() if p[1] == Q4_K]
    weights = [rand_weights(t, N * ops.row_bytes(t, K), dev) for _, t, N, K, _, _ in plan]
    X = {4096: torch.rand(1, 4096, device=dev) * 2 - 1, 14336: torch.rand(1, 14336, device=dev) * 2 - 1}
    xq = {K: ops.quantize_acts(x) for K, x in X.items()}
    launches = []
    i = 0
    for g in (issue_groups(plan) if grouped else [[e] for e in plan]):
        ws = weights[i:i + len(g)]
        i += len(g)
        extra = None
        if grouped and len(g) == 2 and g[0][0] == "gate":      # as issued: pair mode, SILU(gate) * up and its q8 activations in the same launch
            N = g[0][2]
            extra = (torch.zeros(1, N, device=dev), torch.zeros(1, N, device=dev), torch.zeros(N // 256 * 400, dtype=torch.uint8, device=dev))
        launches.append((g, ws, [torch.zeros(1, e[2], device=dev) for e in g], extra))

    def run():
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for g, ws, outs, extra in launches:
            t, K, ng = g[0][1], g[0][3], len(g)
            a = ops.MmvqArgs(type=t, W=ws[0].data_ptr(), nb1=ops.row_bytes(t, K), X=X[K].data_ptr(), xs=K, dst=outs[0].data_ptr(),
                             ds=g[0][2], N=g[0][2], K=K, M=1, ny=1, nz=1, pdl=-1, dbg_sumi=None, Xq=xq[K].data_ptr())
            if extra is not None:
                a.sm_op, a.sm_act, a.sm_out, a.sm_xq = 10, extra[0].data_ptr(), extra[1].data_ptr(), extra[2].data_ptr()
            if ng == 1:
                rc = L.b200_mul_mat_vec_q_ex(C.byref(a), st)
            else:
                Wp = (C.c_void_p * ng)(*[w.data_ptr() for w in ws]); Dp = (C.c_void_p * ng)(*[o.data_ptr() for o in outs])
                Np = (C.c_int * ng)(*[e[2] for e in g]); dsp = (C.c_int64 * ng)(*[e[2] for e in g])
                rc = L.b200_mul_mat_vec_q_group(C.byref(a), C.c_int(ng), Wp, Dp, Np, dsp, st)
            if rc:
                raise RuntimeError("b200_mul_mat_vec_q rc=%d" % rc)

    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            run()
        for _ in range(max(3, warmup)):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(steps):
            g.replay()
        e1.record(st)
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    nbytes = plan_bytes(plan)
    return {"launches": len(launches), "us_per_launch": ms * 1e3 / len(launches), "bytes": nbytes, "gbps": nbytes / (ms * 1e-3) / 1e9}


def mmq_roofline(steps, warmup, dev_index):
    """Prefill MMQ (tcgen05) on the seven Q4_K_M projection shapes of one layer at 512 tokens, through the C ABI: TFLOP/s
    (2*N*K*M per mat-mul, activation conversion included) against the measured dense bf16 peak."""
    import torch
    from llamacog_b200 import ops
    from tools.mmvq_micro import rand_weights
    dev = torch.device("cuda", dev_index)
    M = 512
    shapes = [(Q4_K, N_EMBD, N_EMBD), (Q4_K, N_KV, N_EMBD), (Q6_K, N_KV, N_EMBD), (Q4_K, N_EMBD, N_EMBD),
              (Q4_K, N_FF, N_EMBD), (Q4_K, N_FF, N_EMBD), (Q6_K, N_EMBD, N_FF)]
    with torch.cuda.device(dev):
        Ws = [rand_weights(t, N * ops.row_bytes(t, K) + 256, dev) for t, N, K in shapes]
        Xs = {K: torch.randn(M, K, device=dev) for K in (N_EMBD, N_FF)}

        def run():
            for (t, N, K), W in zip(shapes, Ws):
                ops.mul_mat_q(t, W, Xs[K], N, K)
        for _ in range(max(3, warmup)):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            run()
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    flops = sum(2.0 * N * K * M for _, N, K in shapes)
    return {"tflops": flops / (ms * 1e-3) / 1e12, "ms_per_layer": ms, "flops_per_layer": flops}


def cpu_reference_run(model, n_tokens, threads):
    res, _ = llama_bench(model, ["-ngl", "0", "-p", "0", "-n", str(n_tokens), "-r", "1", "-t", str(threads)], plugin=False)
    return float(res[-1]["avg_ts"])


WORKLOAD_1 = "llama3-8b-q4_k_m tg, batch 1"          # config.workload: the same string in both arms
WORKLOAD_N = "llama3-70b-q4_k_m tg, batch 1"


def reference_arm(args):
    """The reference's own CPU implementation of the path on the host cores (unmodified llama-bench, no plugin)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    multi = args.gpus > 1
    model = ensure_model("llama3-70b") if multi else ensure_model()
    threads = os.cpu_count() or 1
    n = max(4, min(args.steps, 8 if multi else 64))          # bounded sample: 70B on the host cores runs at ~2 tokens/s
    v = cpu_reference_run(model, n, threads)
    line = {"impl": "reference", "metric": METRIC_N if multi else METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "strong" if multi else "weak", "vs_baseline": None,
            "dtype": "q4_K/q5_K/q6_K weights x q8_K activations (int8 dot, f32 accumulate)", "data": "synthetic",
            "config": {"workload": WORKLOAD_N if multi else WORKLOAD_1,
                       "harness": "reference ggml CPU backend via llama-bench -ngl 0 (no GPU is used by this arm)"},
            "cpu_baseline": {"value": v, "unit": "tokens/s", "cores": threads, "kind": "reference",
                             "sample": "llama-bench -ngl 0 -p 0 -n %d -r 1 -t %d (after its 1-token warm-up)" % (n, threads)},
            "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


METRIC_N = "llama-bench tg tokens/s, Llama-3-70B Q4_K_M --split-mode row over N GPUs (batch-1 decode)"


def device_step_ms(stats, n_tok):
    """ms of device time per decoded token from the plugin's B200_TIMING line: replayed CUDA-graph calls only (the first eager
    sighting of a graph and its capture + instantiate call carry host milliseconds and are left out)."""
    if not stats or stats.get("graphs", 0) <= 0:
        raise RuntimeError("the plugin reported no graph_compute calls: the CUDA path did not run")
    gpt = stats["graphs"] / n_tok                      # graph_compute calls per token (one per backend that owns layers)
    if stats.get("replay_graphs", 0) > 0:
        return stats["replay_ms"] / stats["replay_graphs"] * gpt, gpt, "replayed graphs only"
    return stats["device_ms"] / stats["graphs"] * gpt, gpt, "all calls (no graph was replayed)"


def split_parity(n_gpus):
    """--split-mode row / layer against the single-GPU run of the same model (Llama-3.2-1B Q4_K_M, prefill 40 + 8 decode steps):
    a weight row's dot product does not depend on the device that holds it, so the logits must be bit-identical."""
    import numpy as np
    exe = os.path.join(ROOT, "tests", "host", "build", "logits_check")
    if not os.path.exists(exe):
        return "not run (tests/host/build/logits_check missing)"
    model = ensure_model("llama3.2-1b")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = REF_LIB + ":" + os.path.join(ROOT, "llamacog_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
    env["GGML_BACKEND_PATH"] = PLUGIN
    env["CUDA_VISIBLE_DEVICES"] = visible_devices(n_gpus)
    outs = {}
    # row-all: every matrix sharded (B200_SPLIT_MIN_MB=0); row: the default placement (small matrices stay whole on the main device)
    for name, mode, extra in (("none", "none", {}), ("row", "row", {}), ("row-all", "row", {"B200_SPLIT_MIN_MB": "0"}), ("layer", "layer", {})):
        out = os.path.join(MODEL_DIR, "parity_%s_%d.bin" % (name, os.getpid()))
        r = subprocess.run([exe, model, out, "99", "1", "40", "8", "512", mode], capture_output=True, text=True, env=dict(env, **extra), timeout=900)
        if r.returncode != 0:
            return "FAILED to run (%s): %s" % (name, r.stderr[-300:])
        outs[name] = np.fromfile(out, dtype=np.uint8)
        os.remove(out)
    bad = [m for m in ("row", "row-all", "layer") if not np.array_equal(outs[m], outs["none"])]
    return "bit-identical" if not bad else "MISMATCH: " + ",".join(bad)


def visible_devices(n):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    ids = vis.split(",") if vis else [str(i) for i in range(64)]
    return ",".join(ids[:n])


def split_arm(args, world, sampler):
    """Rank 0, all `world` GPUs visible: Llama-3-70B Q4_K_M over unmodified llama-bench, -sm none (one GPU) / row / layer."""
    model = ensure_model("llama3-70b")
    dev = visible_devices(world)
    ts = "/".join(["1"] * world)
    common = ["-ngl", "99", "-fa", "1", "-r", "1", "-mg", "0"]
    out = {"model": "llama3-70b-q4_k_m (random init, tools/make_gguf.py)", "tensor_split": ts, "logs": {}}
    t0 = time.time()
    # e2e: tg K (after W untimed tokens) and pp512, the three placements in one llama-bench invocation each
    def run(sm, extra, timing=False):
        return llama_bench(model, common + ["-sm", sm] + (["-ts", ts] if sm != "none" else []) + extra, plugin=True, timing=timing, device=dev)
    res = {}
    for sm in ("none", "row", "layer"):
        r, st = run(sm, ["-p", "0", "-n", "%d,%d" % (args.warmup, args.steps)])
        rp, _ = run(sm, ["-p", "512", "-n", "0", "-r", "2"])
        res[sm] = {"tg_tokens_per_s": float(r[-1]["avg_ts"]), "pp512_tokens_per_s": float(rp[-1]["avg_ts"])}
    t1 = time.time()
    # device time of the row-split decode step (B200_TIMING adds an event synchronisation per graph: its own run)
    _, st = run("row", ["-p", "0", "-n", "%d,%d" % (args.warmup, args.steps)], timing=True)
    n_tok = args.warmup + args.steps + 2
    ms_step, gpt, how = device_step_ms(st, n_tok)
    out.update({"one_gpu": res["none"], "row": res["row"], "layer": res["layer"],
                "row_tg_speedup_vs_one_gpu": res["row"]["tg_tokens_per_s"] / res["none"]["tg_tokens_per_s"],
                "row_device_ms_per_token": ms_step, "device_time": how, "launches_per_token": st["launches"] / n_tok})
    return out, ms_step, (t0, t1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-split", action="store_true", help="N > 1: replicas only (skip the 70B row-split arm)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return reference_arm(args)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import datetime
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        # rank 0 spends minutes in the split arm while the others wait at a barrier: no 10-minute watchdog
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(hours=3))
        # The long wait (rank 0 alone drives all GPUs through the split arm) must be a HOST-side wait: a rank parked in an NCCL
        # barrier keeps a spinning kernel resident on its GPU, the split model's kernels on that GPU are then time-sliced against it
        # and every device-to-device hand-over takes milliseconds (seen: the 30 s spin budget of a hand-over ran out).
        cpu_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=3))
    if rank == 0:
        model = ensure_model()
    if world > 1:
        dist.barrier()
    model = ensure_model()

    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    device = vis.split(",")[local] if vis else str(local)
    lb = ["-ngl", "99", "-fa", "1", "-p", "0", "-n", "%d,%d" % (args.warmup, args.steps), "-r", "1"]

    sampler = ClockSampler(local)
    sampler.start()
    if world > 1:
        dist.barrier()
    t0 = time.time()
    res, _ = llama_bench(model, lb, plugin=True, device=device)                       # e2e: host buffers in and out
    t1 = time.time()
    clocks = sampler.summary(t0, t1)
    res_t, stats = llama_bench(model, lb, plugin=True, timing=True, device=device)    # device time of graph_compute
    e2e_rank = float(res[-1]["avg_ts"])
    n_tok = args.warmup + args.steps + 2            # llama-bench decodes one warm-up token in front of each test
    ms_step, graphs_per_tok, dev_how = device_step_ms(stats, n_tok)
    launches_per_step = stats["launches"] / n_tok

    if world > 1:
        import torch
        tt = torch.tensor([ms_step, 1e3 / e2e_rank], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_step, e2e_ms = float(tt[0]), float(tt[1])
    else:
        e2e_ms = 1e3 / e2e_rank
    value = world * 1e3 / ms_step
    e2e_val = world * 1e3 / e2e_ms

    if rank == 0:
        n_kv = 256
        line = {"metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "q4_K/q6_K weights x q8_1 activations (int8 dp4a, f32 accumulate), f16 KV", "data": "synthetic",
                "config": {"workload": WORKLOAD_1,
                           "harness": "tg%d: unmodified reference llama-bench -ngl 99 -fa 1 -p 0 -t %d over the libggml-b200.so plugin; random-init GGUF "
                                      "(tools/make_gguf.py, reference quantiser, no_vocab)" % (args.steps, GPU_ARM_THREADS),
                           "l2": "inputs larger than L2 (4.6 GB of weights streamed per step)",
                           "parallelism": "replicas x%d" % world, "graph_compute_calls_per_token": graphs_per_tok,
                           "device_time": dev_how},
                "clocks": clocks,
                "e2e": {"value": e2e_val, "unit": "tokens/s", "h2d_bytes_per_step": 12 + n_kv * 64 * 4, "d2h_bytes_per_step": N_VOCAB * 4},
                "gpu_launches": int(round(launches_per_step * args.steps))}
        if world > 1 and not args.no_split:
            # the north-star multi-GPU workload replaces the replica figures as the headline; they stay as a side key
            replicas = {"metric": METRIC, "value": value, "e2e": e2e_val, "ms_per_step": ms_step, "parallelism": "replicas x%d" % world}
            try:
                parity = split_parity(world)
                sp, ms_row, (s0, s1) = split_arm(args, world, sampler)
                line.update({"metric": METRIC_N, "value": 1e3 / ms_row, "ms_per_step": ms_row, "scaling": "strong",
                             "dtype": "q4_K/q5_K/q6_K weights x q8_1 activations (int8 dp4a, f32 accumulate), f16 KV",
                             "clocks": sampler.summary(s0, s1),
                             "e2e": {"value": sp["row"]["tg_tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 12 + n_kv * 64 * 4,
                                     "d2h_bytes_per_step": N_VOCAB * 4},
                             "gpu_launches": int(round(sp["launches_per_token"] * args.steps)),
                             "split": sp, "split_parity": parity, "replicas": replicas})
                line["config"] = {"workload": WORKLOAD_N,
                                  "harness": "tg%d: unmodified reference llama-bench -ngl 99 -fa 1 -p 0 -sm row -ts %s -t %d over the libggml-b200.so plugin, "
                                             "driven from rank 0 with %d GPUs visible; random-init GGUF" % (args.steps, sp["tensor_split"], GPU_ARM_THREADS, world),
                                  "l2": "inputs larger than L2 (42 GB of weights streamed per step)",
                                  "parallelism": "row split x%d" % world, "device_time": sp["device_time"]}
            except Exception as e:  # noqa: BLE001
                line["split"] = {"error": str(e)[-600:]}
                line["replicas"] = replicas
        sampler.stop_flag = True
        if world > 1:
            dist.barrier(group=cpu_group)
        if not args.no_roofline:
            try:
                k = kernel_roofline(max(20, min(args.steps, 200)), args.warmup, local)
                peak, how = measured_peak()
                traffic = None      # ncu --set full capture of the four shapes, averaged with their launch counts per token
                tpath = os.path.join(ROOT, "profiles", "r01_mmvqr_q4k_traffic.json")
                if os.path.exists(tpath):
                    traffic = json.load(open(tpath)).get("dram_bytes_per_launch_token_mix")
                k1 = kernel_roofline(max(20, min(args.steps, 200)), args.warmup, local, grouped=False)
                pair = None         # the gate / up pair launch itself under ncu --set full (tools/prof_mmvq.py pair4k, round 2)
                ppath = os.path.join(ROOT, "profiles", "r02_mmvqr_pair_traffic.json")
                if os.path.exists(ppath):
                    pl = json.load(open(ppath)).get("launches", [])
                    if pl:
                        pair = {"dram_bytes_per_launch": int(sorted(x["dram_read_bytes"] + x["dram_write_bytes"] for x in pl)[len(pl) // 2]),
                                "algorithmic_bytes": 2 * N_FF * N_EMBD * 144 // 256, "source": "profiles/r02_mmvqr_pair_traffic.json"}
                line["roofline"] = {"bound": "hbm", "kernel": "mmvqr_kernel<Q4_K> as the plugin issues it in the timed region (%d launches per token: q/k/v group, wo, gate/up pair launch with SILU*up, down)" % k["launches"],
                                    "achieved": k["gbps"], "peak": peak, "unit": "GB/s", "frac": k["gbps"] / peak,
                                    "traffic": int(traffic * k1["launches"] / k["launches"]) if traffic else None,
                                    "peak_source": how, "launches": k["launches"], "us_per_launch": k["us_per_launch"],
                                    "algorithmic_bytes_per_launch": k["bytes"] / k["launches"],
                                    "traffic_note": "ncu --set full of the four Q4_K shapes as single-matrix launches (profiles/r01_mmvqr_q4k_traffic.json: DRAM bytes = 1.002 x algorithmic, weights are read once however the matrices are grouped), re-expressed per issued launch",
                                    "traffic_gate_up_pair_launch": pair,
                                    "single_launches": {"note": "the same matrices, every one a launch of its own (the round-1 form of this figure; the step never issues wk / wv alone)",
                                                        "launches": k1["launches"], "us_per_launch": k1["us_per_launch"], "achieved": k1["gbps"], "frac": k1["gbps"] / peak,
                                                        "algorithmic_bytes_per_launch": k1["bytes"] / k1["launches"]},
                                    "whole_step_GBps": plan_bytes(step_plan()) / (ms_step * 1e-3) / 1e9 if world == 1 else None}
            except Exception as e:  # noqa: BLE001
                line["roofline"] = {"error": str(e)[:200]}
        if not args.no_roofline and world == 1:
            # the prefill side of the headline metric: pp512 through the same unmodified llama-bench, and the MMQ tensor-pipe fraction
            try:
                resp, _ = llama_bench(model, ["-ngl", "99", "-fa", "1", "-p", "512", "-n", "0", "-r", "9"], plugin=True, device=device)
                samples = sorted(float(x) for x in resp[-1].get("samples_ts", [])) or [float(resp[-1]["avg_ts"])]
                pre = {"pp512_tokens_per_s": samples[len(samples) // 2],      # median of 9 passes (12 ms each): the first passes of a process run 6 % below the rest, a fresh box occasionally stalls one
                       "pp512_samples": samples,
                       "workload": "llama3-8b-q4_k_m pp512, llama-bench -ngl 99 -fa 1 (MMQ tcgen05 + tensor-core flash attention)"}
                m = mmq_roofline(20, args.warmup, local)
                peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
                tpeak = float(peaks.get("bf16_tflops", 1590.0))
                pre["mmq_roofline"] = {"bound": "tensor", "kernel": "mmq_kernel (7 projections of one layer, M=512)", "achieved": m["tflops"],
                                       "peak": tpeak, "unit": "TFLOP/s", "frac": m["tflops"] / tpeak,
                                       "peak_source": "measured bf16 burst (MEASURED_PEAKS.json)" if peaks else "fallback 1.59 PFLOP/s",
                                       "ms_per_layer": m["ms_per_layer"]}
                tc = tc_issue_peaks()      # tcgen05 issue-rate ceilings of this box (tools/tc_peak.cu): what an int8 / fp8 operand route could reach at most
                if tc:
                    pre["mmq_roofline"].update({"frac_int8_peak": m["tflops"] / tc["int8_tops"], "frac_fp8_peak": m["tflops"] / tc["fp8_tflops"],
                                                "frac_bf16_issue_peak": m["tflops"] / tc["bf16_tflops"], "tc_issue_peaks": tc})
                tp = os.path.join(ROOT, "profiles", "r01_mmq_traffic.json")      # ncu --set full of the three large shapes: DRAM bytes per launch
                if os.path.exists(tp):
                    pre["mmq_roofline"]["traffic"] = {x["shape"]: x["dram_read_bytes"] + x["dram_write_bytes"] for x in json.load(open(tp))["launches"]}
                line["prefill"] = pre
            except Exception as e:  # noqa: BLE001
                line["prefill"] = {"error": str(e)[:200]}
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            try:
                v = cpu_reference_run(model, 16, threads)
                line["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": threads, "kind": "reference",
                                        "sample": "reference llama-bench -ngl 0 -p 0 -n 16 -r 1 -t %d on the same GGUF" % threads}
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": str(e)[:200]}
        print(json.dumps(line), flush=True)
    elif world > 1:
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)            # parked on the host until rank 0 is through with the split arm
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
