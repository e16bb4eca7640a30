mkdir -p gpurun_out
timeout 1500 python tools/logits_diag.py llama3.2-1b > gpurun_out/logits_diag.log 2>&1; echo "exit $?" >> gpurun_out/logits_diag.log
cat gpurun_out/logits_diag.log
