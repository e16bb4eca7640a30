#!/usr/bin/env python
"""bench.py -- headline measurement of the hot path.

Workload (BASELINE.json configs[1]): Llama-3-8B Q4_K_M, batch-1 decode.  One STEP = one token = every
quantised MUL_MAT of the model (32 layers x {wq,wk,wv,wo,gate,up,down} + lm_head, 225 GEMVs, 4.616 GB
of weights in the Q4_K_M tensor-type mix of reference src/llama-quant.cpp:185-224,291-353), run through
the C ABI of include/b200_ops.h from one CUDA graph, the way the backend's graph_compute replays a
decode step.  Weights are synthetic random-init blocks (random payload bits, sane fp16 scales), every
matrix a distinct allocation, so a step streams 4.6 GB from HBM (inputs >> 126 MB L2; no L2 flush
needed).

  value   tokens/s, everything resident in HBM, CUDA events on the launching stream, max over ranks
  e2e     tokens/s through the same C-ABI calls with HOST buffers: each step copies the activation
          vector from pinned host memory, replays the step and reads the logits back
  roofline  the dominant kernel (mmvq_kernel<Q4_K>): algorithmic bytes (N*K*144/256 per launch) over its
          mean launch duration, timed live with CUDA events on a graph holding only those launches
  cpu_baseline  the oracle port (single thread, scalar C) on a bounded sample of the same GEMVs

`--impl reference` times the reference's own CPU code for the path on the host cores: the unmodified
ggml CPU backend from baseline/_ref (test-backend-ops perf -b CPU, all threads) when it is present,
else the oracle port.

N > 1 (torchrun): the path shards by weight rows with no data-path collective at this level (each
rank holds its own replica of the step: weak scaling, value = sum of ranks' tokens / max time).
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

N_EMBD, N_FF, N_LAYER, N_KV, N_VOCAB = 4096, 14336, 32, 1024, 128256
Q4_K, Q6_K = 12, 14
BPW = {Q4_K: 144 / 256, Q6_K: 210 / 256}


def use_more_bits(i, n):   # reference src/llama-quant.cpp:185-187
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def step_plan():
    """[(name, type, N, K, input, output)] of one decode token, Q4_K_M mix."""
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
                self.samples.append((time.time(), mhz, rs))
            except Exception:
                pass
            time.sleep(0.004)

    def summary(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        inside = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples[-3:]
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
                 0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz,
                "reasons": [n for b, n in names.items() if bits & b and n != "gpu_idle"], "samples": len(inside)}


def measured_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_port_baseline(budget_s=12.0):
    """Oracle port, 1 thread, on a bounded sample: rows of the ffn_down Q4_K and Q6_K GEMVs."""
    import numpy as np
    import oracle as O
    rng = np.random.default_rng(0)
    K = N_EMBD
    done_bytes, t_used, rows = 0, 0.0, 256
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    for t in (O.Q4_K, O.Q6_K):
        W = O.random_quant_bytes(t, rows, K, rng)
        O.mul_mat_vec_q(t, W[:8], x, 8, K, 1)
        reps, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < budget_s / 2:
            O.mul_mat_vec_q(t, W, x, rows, K, 1)
            reps += 1
        t_used += time.perf_counter() - t0
        done_bytes += reps * rows * O.row_bytes(t, K)
    tok_s = (done_bytes / t_used) / plan_bytes(step_plan())
    return {"value": tok_s, "unit": "tokens/s", "cores": 1, "kind": "port",
            "sample": "%d-row slices of 4096-wide Q4_K and Q6_K GEMVs (fused q8_1 quantise + int dot), %.0f s, "
                      "scaled by bytes to the 4.616 GB token" % (rows, t_used)}


def reference_arm(args):
    """The reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    plan = step_plan()
    tbo = os.path.join(ROOT, "baseline", "_ref", "bin", "test-backend-ops")
    cores = os.cpu_count() or 1
    out = None
    if os.path.exists(tbo):
        # unmodified reference CPU backend: MUL_MAT m=4096,k=14336,n=1 perf cases of its own test tool
        # (tests/test-backend-ops.cpp:4622-4628), all host threads
        try:
            env = dict(os.environ)
            env["LD_LIBRARY_PATH"] = os.path.join(ROOT, "baseline", "_ref", "lib") + ":" + env.get("LD_LIBRARY_PATH", "")
            r = subprocess.run([tbo, "perf", "-b", "CPU", "-o", "MUL_MAT", "-p", "type_a=(q4_K|q6_K),type_b=f32,m=4096,n=1,k=14336"], capture_output=True, text=True, timeout=900, env=env)
            us = {}
            for line in r.stdout.splitlines():
                m = re.search(r"MUL_MAT\(type_a=(q4_K|q6_K),type_b=f32,m=4096,n=1,k=14336.*?([0-9.]+) us/run", line)
                if m:
                    us[m.group(1)] = float(m.group(2))
            if "q4_K" in us and "q6_K" in us:
                b4, b6 = 4096 * 14336 * BPW[Q4_K], 4096 * 14336 * BPW[Q6_K]
                t_tok = plan_bytes(plan, Q4_K) / b4 * us["q4_K"] * 1e-6 + plan_bytes(plan, Q6_K) / b6 * us["q6_K"] * 1e-6
                out = {"value": 1.0 / t_tok, "kind": "reference", "cores": cores,
                       "sample": "reference test-backend-ops perf -b CPU -o MUL_MAT, m=4096 k=14336 n=1: q4_K %.0f us, q6_K %.0f us "
                                 "per GEMV, scaled by bytes to the 4.616 GB token" % (us["q4_K"], us["q6_K"])}
        except Exception as e:  # noqa: BLE001
            sys.stderr.write("reference tool failed (%s); falling back to the oracle port\n" % e)
    if out is None:
        b = cpu_port_baseline(20.0)
        out = {"value": b["value"], "kind": "port", "cores": 1, "sample": b["sample"]}
    line = {"impl": "reference", "metric": "llama3-8b-q4_k_m tg (batch-1 decode) tokens/s", "value": out["value"],
            "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 / out["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "q4_K/q6_K x q8 (int8 dot, f32 accumulate)", "data": "synthetic",
            "config": {"workload": "llama3-8b-q4_k_m-decode-b1: 225 quantised MUL_MAT GEMVs per token"},
            "cpu_baseline": {"value": out["value"], "unit": "tokens/s", "cores": out["cores"], "kind": out["kind"], "sample": out["sample"]},
            "e2e": {"value": out["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from llamacog_b200 import ops
    from tools.mmvq_micro import rand_weights

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ops.lib()

    plan = step_plan()
    torch.manual_seed(1234 + rank)
    weights = [rand_weights(t, N * ops.row_bytes(t, K), dev) for _, t, N, K, _, _ in plan]
    bufs = {"h": torch.rand(1, N_EMBD, device=dev) * 2 - 1}
    for _, _, N, _, _, o in plan:
        if o not in bufs:
            bufs[o] = torch.zeros(1, N, device=dev)
    host_x = torch.rand(1, N_EMBD).pin_memory()
    host_logits = torch.empty(1, N_VOCAB).pin_memory()
    x_in = torch.empty(1, N_EMBD, device=dev)

    import ctypes as C

    # where the q8_1 quantisation of the activations runs: "none" = fused prologue of every GEMV;
    # "down" = one quantise kernel in front of ffn_down (K=14336), "all" = one per distinct activation
    # vector (q/k/v share one, gate/up share one), the GEMVs then copy the quantised vector
    xq_mode = os.environ.get("B200_BENCH_XQ", "all")
    xq_bufs = {}

    def launch(t, W, X, dst, N, K, name="", src=""):
        xq = None
        if xq_mode == "all" or (xq_mode == "down" and K > 8192):
            fresh = name in ("wq", "wo", "gate", "down", "output")     # first consumer of a new activation vector
            if fresh or src not in xq_bufs:
                if src not in xq_bufs:
                    xq_bufs[src] = torch.empty(((K + 255) // 256 * 400,), dtype=torch.uint8, device=dev)
                rc = ops.lib().b200_quantize_acts(C.c_void_p(X.data_ptr()), C.c_int64(K), C.c_void_p(xq_bufs[src].data_ptr()),
                                                  C.c_int(K), C.c_int(1), C.c_void_p(torch.cuda.current_stream().cuda_stream))
                if rc:
                    raise RuntimeError("b200_quantize_acts rc=%d" % rc)
            xq = xq_bufs[src].data_ptr()
        a = ops.MmvqArgs(type=t, W=W.data_ptr(), nb1=ops.row_bytes(t, K), X=X.data_ptr(), xs=K, dst=dst.data_ptr(), ds=N,
                         N=N, K=K, M=1, ny=1, nz=1, pdl=-1, dbg_sumi=None, Xq=xq)
        rc = ops.lib().b200_mul_mat_vec_q_ex(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            raise RuntimeError("b200_mul_mat_vec_q_ex rc=%d" % rc)

    def run_step(only=None, first_input=None):
        for (name, t, N, K, i, o), W in zip(plan, weights):
            if only is not None and t != only:
                continue
            X = first_input if (first_input is not None and name == "wq" and W is weights[0]) else bufs[i]
            launch(t, W, X, bufs[o], N, K, name, i)

    st = torch.cuda.Stream()
    graphs = {}
    with torch.cuda.stream(st):
        for _ in range(2):
            run_step()
        torch.cuda.synchronize()
        for key, only, first in (("all", None, None), ("q4k", Q4_K, None), ("e2e", None, x_in)):
            g = torch.cuda.CUDAGraph()
            n0 = ops.kernel_launch_count()
            with torch.cuda.graph(g, stream=st):
                run_step(only, first)
            graphs[key] = (g, ops.kernel_launch_count() - n0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        with torch.cuda.stream(st):
            for _ in range(warmup):
                fn()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.time()
            e0.record(st)
            for _ in range(steps):
                fn()
            e1.record(st)
            barrier()
            t1 = time.time()
        ms = e0.elapsed_time(e1)
        if world > 1:
            tt = torch.tensor([ms], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms, t0, t1

    sampler = ClockSampler(local)
    sampler.start()
    g_all, launches_per_step = graphs["all"]
    ms, t0, t1 = timed(g_all.replay, args.steps, args.warmup)
    clocks = sampler.summary(t0, t1)
    ms_step = ms / args.steps
    value = world * 1e3 / ms_step

    # dominant kernel, timed live: a graph with only the Q4_K launches of the step
    g_q4, n_q4 = graphs["q4k"]
    ms_q4, _, _ = timed(g_q4.replay, args.steps, args.warmup)
    q4_bytes = plan_bytes(plan, Q4_K)
    ach = q4_bytes / (ms_q4 / args.steps * 1e-3) / 1e9
    peak, how = measured_peak()
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r01_mmvq_q4k_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None

    # end to end: host buffers in, logits out, every step
    g_e2e, _ = graphs["e2e"]

    def e2e_step():
        x_in.copy_(host_x, non_blocking=True)
        g_e2e.replay()
        host_logits.copy_(bufs["logits"], non_blocking=True)
        st.synchronize()

    ms_e2e, _, _ = timed(e2e_step, args.steps, args.warmup)
    sampler.stop_flag = True
    e2e_val = world * 1e3 / (ms_e2e / args.steps)

    if rank == 0:
        line = {"metric": "llama3-8b-q4_k_m tg (batch-1 decode) tokens/s", "value": value, "unit": "tokens/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "q4_K/q6_K x q8 (int8 dot, f32 accumulate)", "data": "synthetic",
                "config": {"workload": "llama3-8b-q4_k_m-decode-b1: 225 quantised MUL_MAT GEMVs per token (4.616 GB weights, "
                                       "Q4_K_M type mix), fused q8_1 prologue; non-matmul glue ops not yet in the timed step",
                           "l2": "inputs larger than L2 (4.6 GB streamed per step)", "activation_quantise": xq_mode, "parallelism": "replicas x%d" % world},
                "clocks": clocks,
                "e2e": {"value": e2e_val, "unit": "tokens/s", "h2d_bytes_per_step": N_EMBD * 4, "d2h_bytes_per_step": N_VOCAB * 4},
                "gpu_launches": launches_per_step * args.steps,
                "roofline": {"bound": "hbm", "kernel": "mmvq_kernel<Q4_K>", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": traffic, "peak_source": how,
                             "launches": n_q4, "algorithmic_bytes_per_step": q4_bytes,
                             "whole_step_GBps": plan_bytes(plan) / (ms_step * 1e-3) / 1e9}}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_port_baseline()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
