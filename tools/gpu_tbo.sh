export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
timeout 900 baseline/_ref/bin/test-backend-ops test -b B200_0 > gpurun_out/tbo_all.log 2>&1
echo "exit $?" >> gpurun_out/tbo_all.log
grep -c "OK" gpurun_out/tbo_all.log; grep -E "FAIL|not supported|exit|backends passed" gpurun_out/tbo_all.log | sed 's/\x1b\[[0-9;]*m//g' | awk '{print $1, $NF}' | sort | uniq -c | sort -rn | head -20
