// logits_check -- end-to-end parity harness (test infrastructure): runs a fixed token sequence through
// libllama on whatever backend the environment selects (-ngl 0: reference CPU backend; -ngl 99 with
// GGML_BACKEND_PATH=libggml-b200.so: the B200 plugin) and writes the logits of the last prompt token and of
// every teacher-forced decode step to a file.  tests/test_gpu_plugin.py runs it once per backend and compares.
//
// usage: logits_check <model.gguf> <out.bin> <ngl> <flash_attn 0|1> <n_prompt> <n_decode> [n_ubatch] [none|layer|row]
#include "llama.h"
#include "ggml-backend.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

int main(int argc, char ** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s model out ngl fa n_prompt n_decode [n_ubatch] [none|layer|row]\n", argv[0]); return 2; }
    const char * path = argv[1]; const char * out = argv[2];
    const int ngl = atoi(argv[3]), fa = atoi(argv[4]), n_prompt = atoi(argv[5]), n_decode = atoi(argv[6]);
    const int n_ubatch = argc > 7 ? atoi(argv[7]) : 512;
    llama_backend_init();
    ggml_backend_load_all();
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl;
    if (argc > 8) {      // how the model is spread over several GPUs (src/llama-model.cpp:337-366, :1505-1600)
        if (strcmp(argv[8], "none") == 0) mp.split_mode = LLAMA_SPLIT_MODE_NONE;
        else if (strcmp(argv[8], "layer") == 0) mp.split_mode = LLAMA_SPLIT_MODE_LAYER;
        else if (strcmp(argv[8], "row") == 0) mp.split_mode = LLAMA_SPLIT_MODE_ROW;
        else { fprintf(stderr, "unknown split mode %s\n", argv[8]); return 2; }
    }
    llama_model * model = llama_model_load_from_file(path, mp);
    if (!model) { fprintf(stderr, "cannot load %s\n", path); return 1; }
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = n_prompt + n_decode + 8; cp.n_batch = n_prompt > n_ubatch ? n_prompt : n_ubatch; cp.n_ubatch = n_ubatch;
    cp.flash_attn = fa != 0; cp.n_threads = 16; cp.n_threads_batch = 16;
    llama_context * ctx = llama_init_from_model(model, cp);
    if (!ctx) { fprintf(stderr, "cannot create the context\n"); return 1; }
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(model));
    std::vector<llama_token> toks(n_prompt + n_decode);
    for (size_t i = 0; i < toks.size(); ++i) toks[i] = (llama_token)((i * 7919u + 13u) % (unsigned)n_vocab);
    FILE * f = fopen(out, "wb");
    if (!f) return 1;
    fwrite(&n_vocab, sizeof(int), 1, f);
    if (llama_decode(ctx, llama_batch_get_one(toks.data(), n_prompt)) != 0) { fprintf(stderr, "prefill failed\n"); return 1; }
    fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, f);
    for (int d = 0; d < n_decode; ++d) {
        if (llama_decode(ctx, llama_batch_get_one(&toks[n_prompt + d], 1)) != 0) { fprintf(stderr, "decode failed\n"); return 1; }
        fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, f);
    }
    fclose(f);
    llama_free(ctx);
    llama_model_free(model);
    llama_backend_free();
    return 0;
}
