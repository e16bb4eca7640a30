// logits_check -- end-to-end parity harness (test infrastructure): runs a fixed token sequence through
// libllama on whatever backend the environment selects (-ngl 0: reference CPU backend; -ngl 99 with
// GGML_BACKEND_PATH=libggml-b200.so: the B200 plugin) and writes the logits of the last prompt token and of
// every teacher-forced decode step to a file.  tests/test_gpu_plugin.py runs it once per backend and compares.
//
// With a tenth argument every MUL_MAT / MUL_MAT_ID / FLASH_ATTN_EXT node (handed out by the scheduler's eval callback,
// ggml-backend.cpp:1415-1447) is re-evaluated on the reference CPU backend from the same input bytes and a record
// {char name[64]; int op; int ne1; int64 n; double nmse} is written per node: the per-node parity twin
// (tests/test_gpu_plugin.py::test_per_node_parity_with_reference_cpu_backend).
//
// An eleventh argument sets the tokens per decode step (default 1): steps of 9 .. 64 tokens repeat their shapes like single-token
// decode does, which is what the plugin's CUDA-graph replay of batched decode relies on (the logits of a step's last token are written).
//
// usage: logits_check <model.gguf> <out.bin> <ngl> <flash_attn 0|1> <n_prompt> <n_decode> [n_ubatch] [none|layer|row|-] [nodes.bin|-] [step]
#include "llama.h"
#include "ggml-backend.h"
#include "ggml.h"
#include "ggml-alloc.h"
#include <cstdint>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static FILE * g_nodes = nullptr;
static ggml_backend_t g_cpu = nullptr;

// The per-node twin: when the scheduler hands out node t (computed by whatever backend runs the model), the same op is evaluated
// on the reference CPU backend FROM THE SAME INPUT BYTES (the sources are read back from the device, strides kept), and the
// normalised squared error of t against that is recorded.  This is ggml_backend_compare_graph_backend (ggml-backend.cpp:1829-1865)
// with the inputs re-synchronised at every node, so no drift accumulates: the bar can be the reference's per-op bar.
static bool node_cb(struct ggml_tensor * t, bool ask, void *) {
    const bool want = (t->op == GGML_OP_MUL_MAT || t->op == GGML_OP_MUL_MAT_ID || t->op == GGML_OP_FLASH_ATTN_EXT) && t->type == GGML_TYPE_F32;
    if (ask) return want;
    if (!want) return true;
    for (int i = 0; i < GGML_MAX_SRC; ++i) {      // weights repacked by a CPU extra buffer type (AMX ..) cannot be read back: not a node of the GPU run
        const ggml_tensor * s = t->src[i];
        if (!s) continue;
        ggml_backend_buffer_t b = s->view_src ? s->view_src->buffer : s->buffer;
        const char * bn = b ? ggml_backend_buffer_name(b) : "";
        if (!b || strstr(bn, "AMX") || strstr(bn, "REPACK") || strstr(bn, "KLEIDIAI")) return true;
    }
    ggml_init_params ip = { ggml_tensor_overhead() * (GGML_MAX_SRC + 2), nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_tensor * twin = ggml_new_tensor(ctx, t->type, GGML_MAX_DIMS, t->ne);
    twin->op = t->op;
    memcpy(twin->op_params, t->op_params, sizeof(t->op_params));
    for (int i = 0; i < GGML_MAX_SRC; ++i) {
        const ggml_tensor * s = t->src[i];
        if (!s) continue;
        ggml_tensor * c = ggml_new_tensor(ctx, s->type, GGML_MAX_DIMS, s->ne);
        for (int d = 0; d < GGML_MAX_DIMS; ++d) c->nb[d] = s->nb[d];      // views into the KV cache, permuted queries: same strides
        twin->src[i] = c;
    }
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, g_cpu);
    std::vector<uint8_t> tmp;
    for (int i = 0; i < GGML_MAX_SRC; ++i) {
        if (!t->src[i]) continue;
        const size_t nb = ggml_nbytes(t->src[i]);
        tmp.resize(nb);
        ggml_backend_tensor_get(t->src[i], tmp.data(), 0, nb);
        ggml_backend_tensor_set(twin->src[i], tmp.data(), 0, nb);
    }
    ggml_cgraph * gf = nullptr;
    ggml_init_params gp = { ggml_graph_overhead() + ggml_tensor_overhead(), nullptr, true };
    ggml_context * gctx = ggml_init(gp);
    gf = ggml_new_graph(gctx);
    ggml_build_forward_expand(gf, twin);
    ggml_backend_graph_compute(g_cpu, gf);
    const int64_t n = ggml_nelements(t);
    std::vector<float> got((size_t)n), ref((size_t)n);
    ggml_backend_tensor_get(t, got.data(), 0, (size_t)n * sizeof(float));
    ggml_backend_tensor_get(twin, ref.data(), 0, (size_t)n * sizeof(float));
    double num = 0.0, den = 0.0;
    for (int64_t i = 0; i < n; ++i) { const double e = (double)got[i] - ref[i]; num += e * e; den += (double)ref[i] * ref[i]; }
    const double nmse = den > 0.0 ? num / den : num;
    char name[64] = {0};
    strncpy(name, t->name, 63);
    const int op = (int)t->op;
    const int m = (int)t->ne[1];
    fwrite(name, 1, 64, g_nodes); fwrite(&op, sizeof(int), 1, g_nodes); fwrite(&m, sizeof(int), 1, g_nodes); fwrite(&n, sizeof(int64_t), 1, g_nodes);
    fwrite(&nmse, sizeof(double), 1, g_nodes);
    ggml_free(gctx);
    ggml_backend_buffer_free(buf);
    ggml_free(ctx);
    return true;
}

int main(int argc, char ** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s model out ngl fa n_prompt n_decode [n_ubatch] [none|layer|row]\n", argv[0]); return 2; }
    const char * path = argv[1]; const char * out = argv[2];
    const int ngl = atoi(argv[3]), fa = atoi(argv[4]), n_prompt = atoi(argv[5]), n_decode = atoi(argv[6]);
    const int n_ubatch = argc > 7 ? atoi(argv[7]) : 512;
    llama_backend_init();
    ggml_backend_load_all();
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl;
    if (argc > 8 && strcmp(argv[8], "-") != 0) {      // how the model is spread over several GPUs (src/llama-model.cpp:337-366, :1505-1600)
        if (strcmp(argv[8], "none") == 0) mp.split_mode = LLAMA_SPLIT_MODE_NONE;
        else if (strcmp(argv[8], "layer") == 0) mp.split_mode = LLAMA_SPLIT_MODE_LAYER;
        else if (strcmp(argv[8], "row") == 0) mp.split_mode = LLAMA_SPLIT_MODE_ROW;
        else { fprintf(stderr, "unknown split mode %s\n", argv[8]); return 2; }
    }
    llama_model * model = llama_model_load_from_file(path, mp);
    if (!model) { fprintf(stderr, "cannot load %s\n", path); return 1; }
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = n_prompt + n_decode * (argc > 10 ? atoi(argv[10]) : 1) + 8; cp.n_batch = n_prompt > n_ubatch ? n_prompt : n_ubatch; cp.n_ubatch = n_ubatch;
    cp.flash_attn = fa != 0; cp.n_threads = 16; cp.n_threads_batch = 16;
    const int step = argc > 10 ? atoi(argv[10]) : 1;
    if (argc > 9 && strcmp(argv[9], "-") != 0) {
        g_nodes = fopen(argv[9], "wb");
        if (!g_nodes) return 1;
        g_cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
        if (!g_cpu) { fprintf(stderr, "no CPU backend for the per-node twin\n"); return 1; }
        cp.cb_eval = node_cb; cp.cb_eval_user_data = nullptr;
    }
    llama_context * ctx = llama_init_from_model(model, cp);
    if (!ctx) { fprintf(stderr, "cannot create the context\n"); return 1; }
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(model));
    std::vector<llama_token> toks(n_prompt + n_decode * step);
    for (size_t i = 0; i < toks.size(); ++i) toks[i] = (llama_token)((i * 7919u + 13u) % (unsigned)n_vocab);
    FILE * f = fopen(out, "wb");
    if (!f) return 1;
    fwrite(&n_vocab, sizeof(int), 1, f);
    if (llama_decode(ctx, llama_batch_get_one(toks.data(), n_prompt)) != 0) { fprintf(stderr, "prefill failed\n"); return 1; }
    fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, f);
    for (int d = 0; d < n_decode; ++d) {
        if (llama_decode(ctx, llama_batch_get_one(&toks[n_prompt + d * step], step)) != 0) { fprintf(stderr, "decode failed\n"); return 1; }
        fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, f);
    }
    fclose(f);
    if (g_nodes) fclose(g_nodes);
    llama_free(ctx);
    llama_model_free(model);
    llama_backend_free();
    return 0;
}
