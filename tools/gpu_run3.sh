set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
MICRO_XQ=1 python tools/prof_mmvq.py down4k up4k > gpurun_out/prof_plain.log 2>&1 &&
MICRO_XQ=1 ncu --set full --clock-control none --import-source on -k regex:mmvq_kernel -o gpurun_out/prof_mmvq_v2b -f python tools/prof_mmvq.py down4k up4k > gpurun_out/ncu_v2.log 2>&1
