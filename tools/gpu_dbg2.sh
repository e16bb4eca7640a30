export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
for preset in tiny tiny-moe; do python tools/make_gguf.py $preset q4_k_m /tmp/$preset.gguf > /dev/null 2>&1
for fa in 0 1; do echo "== $preset fa=$fa"; B200_DEBUG=1 timeout 120 baseline/_ref/bin/llama-bench -m /tmp/$preset.gguf -ngl 99 -fa $fa -p 48 -n 8 -r 1 2>&1 | grep -E "refused|failed|tg8|pp48" | head -5; done; done
