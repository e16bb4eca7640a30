#!/bin/bash
# gpu_run.sh <task> [args] -- the recurring GPU-box procedures of this repo in one place (run through gpurun; outputs under gpurun_out/).
#   tests [pytest args]     the -m gpu suite
#   bench [bench.py args]   bench.py (N = 1)
#   benchn <N> [args]       bench.py under torchrun on N GPUs (70B row split arm)
#   launches [tg|pp]        ncu launch list of a decode step / a pp512 pass (graphs off: every kernel is a launch) -> gpurun_out/launches_*.csv
#   ncu <regex> <cmd...>    one `ncu --set full` capture of the kernels matching <regex> while <cmd> runs -> gpurun_out/ncu_<regex>.ncu-rep
#   tbo                     the reference's test-backend-ops against the plugin, all ops -> gpurun_out/tbo_all.log
#   bitident                logits with every fusion on / off (must be identical)
#   ktrace                  kernel timeline of a decode step (needs a B200_KTRACE=1 build)
#   split [models]          llama-bench -sm row variants on the visible GPUs (tools/gpu_split.sh)
#   refcuda [models]        the reference's own ggml-cuda (sm_100 build) next to the plugin (tools/gpu_ref_cuda.sh)
#   loadtime                model load wall time, staged vs plain upload (tools/gpu_load_time.sh)
#   config3                 Llama-3-8B Q8_0 with f16 and q8_0 KV cache: pp512 / tg128 at depth 0 and at 28672
#   cpu                     the reference CPU backend on the box's host cores: 1B Q4_0 pp512 / tg128 (config 1), 8B Q4_K_M pp64 / tg16
#   moe                     Mixtral-8x7B Q4_K_M pp512 / tg128 and llama-batched-bench
#   ab <VAR> [args]         A/B of one plugin switch on tg128 (VAR=1 / VAR=0, twice): tok/s and replayed-graph device time (tools/gpu_ab.sh)
#   moedebug                why the MoE fusions match or not on the Mixtral model (B200_DEBUG_FUSE lines of a 4-token decode)
#   check70b                Llama-3-70B on one GPU: fused == unfused logits, tg128 / pp512 with the gate/up pair mode on and off
#   final                   the round's evidence in one box session (tools/gpu_final.sh)
# MODEL_PRESET / MODEL_FTYPE select the model of bitident / launches / ab (tools/make_gguf.py presets and ftypes).
set -u
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:${LD_LIBRARY_PATH:-}
PLUGIN=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
mkdir -p gpurun_out
task=${1:-tests}; shift || true
model() { python -c "import bench; print(bench.ensure_model('${MODEL_PRESET:-llama3-8b}', '${MODEL_FTYPE:-q4_k_m}'))"; }      # MODEL_FTYPE=q8_0 / MODEL_PRESET=mixtral-8x7b ...: another quantisation / model
case $task in
  tests)   python -m pytest tests -m gpu -x -q "$@" ;;
  bench)   python bench.py "$@" ;;
  benchn)  n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $n "$@" ;;
  launches)
    M=$(model); export GGML_BACKEND_PATH=$PLUGIN
    if [ "${1:-tg}" = pp ]; then A="-p 512 -n 0"; else A="-p 0 -n 3"; fi
    B200_CUDA_GRAPHS=0 timeout 300 $B/llama-bench -m $M -ngl 99 -fa 1 $A -r 1 -t 4 > gpurun_out/launches_${1:-tg}_plain.log 2>&1 || exit 1
    B200_CUDA_GRAPHS=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_${1:-tg}.csv \
        $B/llama-bench -m $M -ngl 99 -fa 1 $A -r 1 -t 4 > gpurun_out/launches_${1:-tg}_ncu.log 2>&1
    python tools/launch_agg.py gpurun_out/launches_${1:-tg}.csv | head -n 30 ;;
  ncu)     re=$1; shift; "$@" > gpurun_out/ncu_plain.log 2>&1 || exit 1
           timeout 900 ncu --set full --clock-control none --import-source on -k regex:$re -c ${NCU_COUNT:-4} -f -o gpurun_out/ncu_$re "$@" > gpurun_out/ncu_$re.log 2>&1; tail -n 3 gpurun_out/ncu_$re.log ;;
  tbo)     export GGML_BACKEND_PATH=$PLUGIN; timeout 900 $B/test-backend-ops test -b B200_0 > gpurun_out/tbo_all.log 2>&1; echo "exit $?" >> gpurun_out/tbo_all.log
           sed 's/\x1b\[[0-9;]*m//g' gpurun_out/tbo_all.log | grep -E "FAIL|not supported|exit|backends passed|tests passed" | awk '{print $1, $NF}' | sort | uniq -c | sort -rn | head -n 20 ;;
  bitident)
    M=$(model); export GGML_BACKEND_PATH=$PLUGIN; L=tests/host/build/logits_check
    run() { name=$1; shift; env "$@" timeout 600 $L $M /tmp/$name.bin 99 1 $NP $ND > /dev/null 2>&1 || echo "$name failed"; }
    for cfg in "20 8" "600 12" "1100 6"; do
      set -- $cfg; NP=$1; ND=$2
      run base B200_FUSE=0; run fused B200_FUSE=1; run nogemv B200_FUSE=1 B200_FUSE_GEMV=0 B200_FUSE_ROPE=0 B200_FUSE_HOIST=0
      echo "== n_prompt=$NP n_decode=$ND"
      for v in fused nogemv; do cmp -s /tmp/base.bin /tmp/$v.bin && echo "  $v == unfused" || { echo "  $v DIFFERS from unfused"; python tools/cmp_bins.py /tmp/base.bin /tmp/$v.bin; }; done
    done ;;
  ktrace)  M=$(model); export GGML_BACKEND_PATH=$PLUGIN
           B200_KTRACE=/tmp/ktrace.txt timeout 600 $B/llama-bench -m $M -ngl 99 -fa 1 -p 0 -n 64 -r 1 -t 4 > gpurun_out/kt_lb.log 2>&1
           python tools/ktrace_report.py /tmp/ktrace.txt 0 ${KT_N:-20} | tee gpurun_out/ktrace_report.txt ;;
  split)   MODELS="${*:-llama3-8b llama3-70b}" bash tools/gpu_split.sh ;;
  refcuda) MODELS="${*:-llama3-8b}" bash tools/gpu_ref_cuda.sh ;;
  loadtime) bash tools/gpu_load_time.sh ;;
  config3)
    export GGML_BACKEND_PATH=$PLUGIN; mkdir -p /tmp/b200_models; f=/tmp/b200_models/llama3-8b-q8_0.gguf
    [ -f $f ] || python tools/make_gguf.py llama3-8b q8_0 $f > /dev/null
    for kv in f16 q8_0; do
      echo "=== Llama-3-8B Q8_0, $kv KV cache"
      timeout 600 $B/llama-bench -m $f -ngl 99 -fa 1 -ctk $kv -ctv $kv -p 512 -n 128 -d 0,28672 -r 4 -t 4 2>&1 | grep -E "pp512|tg128|error"
    done ;;
  cpu)
    unset GGML_BACKEND_PATH; mkdir -p /tmp/b200_models; T=$(nproc)
    f=/tmp/b200_models/llama3.2-1b-q4_0.gguf; [ -f $f ] || python tools/make_gguf.py llama3.2-1b q4_0 $f > /dev/null
    echo "=== reference CPU backend, $T threads"; timeout 900 $B/llama-bench -m $f -ngl 0 -p 512 -n 128 -r 2 -t $T 2>&1 | grep -E "pp512|tg128|error"
    f=$(model); timeout 900 $B/llama-bench -m $f -ngl 0 -p 64 -n 16 -r 1 -t $T 2>&1 | grep -E "pp64|tg16|error" ;;
  moe)
    export GGML_BACKEND_PATH=$PLUGIN; mkdir -p /tmp/b200_models; f=/tmp/b200_models/mixtral-8x7b-q4_k_m.gguf
    [ -f $f ] || python tools/make_gguf.py mixtral-8x7b q4_k_m $f > /dev/null
    timeout 600 $B/llama-bench -m $f -ngl 99 -fa 1 -p 512 -n 128 -r 2 -t 4 2>&1 | grep -E "pp512|tg128|error"
    timeout 600 $B/llama-batched-bench -m $f -ngl 99 -fa -npp 128 -ntg 32 -npl 1,2,4,8,16,32,64 -c 16384 2>&1 | grep -E "^\|" ;;
  ab)      bash tools/gpu_ab.sh "$@" ;;
  moedebug) bash tools/gpu_moe_debug.sh ;;
  check70b) bash tools/gpu_70b_check.sh ;;
  final)   bash tools/gpu_final.sh ;;
  *) echo "unknown task $task"; exit 2 ;;
esac
