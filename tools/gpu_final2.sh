# final evidence: pp512 launch list of the final tree, ncu --set full of the tensor-core attention kernel at pp512 / pp2048 shapes
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
mkdir -p gpurun_out
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
B=baseline/_ref/bin
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_pp.csv $B/llama-bench -m $M -ngl 99 -fa 1 -p 512 -n 0 -r 1 > gpurun_out/ncu_pp.log 2>&1
wc -l gpurun_out/launches_pp.csv
cat > /tmp/fa_drv.py <<'PY'
import math, sys, torch
sys.path.insert(0, ".")
from llamacog_b200 import ops
for N in (512, 2048):
    q = torch.randn(1, 32, N, 128, device="cuda"); k = torch.randn(1, 8, N, 128, device="cuda").half(); v = torch.randn(1, 8, N, 128, device="cuda").half()
    mask = torch.full(((N + 63) // 64 * 64, (N + 255) // 256 * 256), float("-inf"), dtype=torch.float16, device="cuda")
    for i in range(N): mask[i, : i + 1] = 0
    for _ in range(3): ops.flash_attn_ext(q, k, v, mask, 1.0 / math.sqrt(128))
torch.cuda.synchronize(); print("done")
PY
timeout 300 python /tmp/fa_drv.py > gpurun_out/fa_plain.log 2>&1; tail -1 gpurun_out/fa_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fattn_tc_kernel -c 6 -f -o gpurun_out/prof_fattn_tc python /tmp/fa_drv.py > gpurun_out/prof_fattn_ncu.log 2>&1; tail -2 gpurun_out/prof_fattn_ncu.log
ncu -i gpurun_out/prof_fattn_tc.ncu-rep --page raw --csv > gpurun_out/prof_fattn_tc_raw.csv 2>/dev/null
ls -la gpurun_out/prof_fattn_tc.ncu-rep
