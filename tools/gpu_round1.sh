set -x
lscpu | grep -E "Model name|^CPU\(s\)|Flags" | cut -c1-2500 > gpurun_out/box_cpu.txt
nvidia-smi > gpurun_out/nvidia-smi.txt
(baseline/_ref/bin/test-quantize-fns > gpurun_out/tqf.log 2>&1; echo "tqf exit $?" >> gpurun_out/tqf.log)
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python tools/mmvq_micro.py 1 > gpurun_out/micro.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mmvq_kernel -s 4 -c 3 -o gpurun_out/prof_mmvq -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
