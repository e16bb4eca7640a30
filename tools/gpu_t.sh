export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
M=$(python -c "import bench; print(bench.ensure_model())")
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
L=tests/host/build/logits_check
run() { name=$1; shift; env "$@" timeout 600 $L $M /tmp/$name.bin 99 1 $NP $ND > /dev/null 2>&1 || echo "$name failed"; }
for cfg in "20 8" "600 12" "1100 6"; do
  set -- $cfg; NP=$1; ND=$2
  run base B200_FUSE=0
  run fused B200_FUSE=1
  run nogemv B200_FUSE=1 B200_FUSE_GEMV=0 B200_FUSE_ROPE=0 B200_FUSE_HOIST=0
  echo "== n_prompt=$NP n_decode=$ND"
  for v in fused nogemv; do cmp -s /tmp/base.bin /tmp/$v.bin && echo "  $v == unfused" || { echo "  $v DIFFERS from unfused"; python tools/cmp_bins.py /tmp/base.bin /tmp/$v.bin; }; done
done
timeout 900 python -m pytest tests/test_gpu_fused.py -x -q 2>&1 | tail -2
