#!/bin/bash
# gpu_final.sh -- the round's evidence in one box session (run through gpurun; everything lands in gpurun_out/, copy what is to be
# judged into profiles/): GPU test suite, bench.py (N = 1), launch lists of the bench workload, test-backend-ops summary, configs 1 / 3 / 4.
set -u
mkdir -p gpurun_out
R=${ROUND:-r02}
( timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -n 5 ) > gpurun_out/${R}_gputests.log; tail -n 2 gpurun_out/${R}_gputests.log
timeout 900 python bench.py > gpurun_out/${R}_bench_n1.json 2> gpurun_out/${R}_bench_n1.err; cut -c1-400 gpurun_out/${R}_bench_n1.json
timeout 600 python bench.py --impl reference > gpurun_out/${R}_bench_n1_reference.json 2>> gpurun_out/${R}_bench_n1.err; cut -c1-300 gpurun_out/${R}_bench_n1_reference.json
bash tools/gpu_run.sh launches tg > gpurun_out/${R}_launches_tg.txt 2>&1; head -n 8 gpurun_out/${R}_launches_tg.txt
bash tools/gpu_run.sh launches pp > gpurun_out/${R}_launches_pp.txt 2>&1; head -n 8 gpurun_out/${R}_launches_pp.txt
bash tools/gpu_run.sh tbo > gpurun_out/${R}_tbo_summary.txt 2>&1; tail -n 6 gpurun_out/${R}_tbo_summary.txt
bash tools/gpu_run.sh config3 > gpurun_out/${R}_config3_q8_0.log 2>&1; cat gpurun_out/${R}_config3_q8_0.log
bash tools/gpu_run.sh moe > gpurun_out/${R}_mixtral.log 2>&1; cat gpurun_out/${R}_mixtral.log
MODEL_PRESET=mixtral-8x7b bash tools/gpu_run.sh launches tg > gpurun_out/${R}_launches_tg_mixtral_final.txt 2>&1; head -n 8 gpurun_out/${R}_launches_tg_mixtral_final.txt
# one full ncu capture of the dominant launch as the plugin issues it (gate / up pair mode): DRAM traffic per launch
bash tools/gpu_run.sh ncu mmvqr python tools/prof_mmvq.py pair4k; python tools/ncu_traffic.py gpurun_out/ncu_mmvqr.ncu-rep > gpurun_out/${R}_mmvqr_pair_traffic.json 2>&1; cat gpurun_out/${R}_mmvqr_pair_traffic.json
if [ "${FULL:-0}" = 1 ]; then
bash tools/gpu_run.sh cpu > gpurun_out/${R}_cpu_baseline.log 2>&1; cat gpurun_out/${R}_cpu_baseline.log
bash tools/gpu_run.sh loadtime > gpurun_out/${R}_load_time_8b.log 2>&1; tail -n 4 gpurun_out/${R}_load_time_8b.log
fi
