export LD_LIBRARY_PATH=$PWD/baseline/_ref/lib:$PWD/llamacog_b200:$LD_LIBRARY_PATH
export GGML_BACKEND_PATH=$PWD/llamacog_b200/libggml-b200.so
python tools/make_gguf.py llama3-8b q4_k_m /tmp/l8b-q4km.gguf > /dev/null 2>&1
B200_DEBUG_FUSE=1 timeout 300 tests/host/build/logits_check /tmp/l8b-q4km.gguf /tmp/x.bin 99 1 4 2 2>&1 | grep "ggml-b200: node" | head -70
