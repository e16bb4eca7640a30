/*
 * ggml-b200.h -- entry points of libggml-b200.so, the B200 (sm_100a) ggml backend plugin.
 *
 * The plugin is loaded by the unmodified host (libggml's registry) with dlopen; the two symbols below
 * are exactly what the registry binds (reference ggml/src/ggml-backend-reg.cpp:232-276, typedefs
 * ggml_backend_init_t / ggml_backend_score_t in ggml/src/ggml-backend-impl.h:209-214).  Everything
 * else is reached through the vtables of the returned ggml_backend_reg (ggml-backend-impl.h:17-207):
 * devices "B200_0".., their buffer types, buffers and backends (streams).
 *
 * ggml_backend_reg_get_proc_address(reg, name) answers:
 *   "ggml_backend_get_features"            -> ggml_backend_get_features_t  (ARCHS=1000a, ...)
 *   "ggml_backend_b200_device_time_ms"     -> double (*)(long * n_graphs): device time spent inside
 *                                             graph_compute since load when B200_TIMING=1 (bench.py)
 *   "ggml_backend_b200_kernel_launches"    -> int (*)(void): kernels launched so far
 *   "ggml_backend_split_buffer_type"       -> ggml_backend_buffer_type_t (*)(int main_device, const float * tensor_split):
 *                                             the row-split buffer type of `--split-mode row` (reference
 *                                             ggml-cuda.cu:753-1084, :3476-3478; consumer src/llama-model.cpp:342-360).
 *                                             Weight rows are sharded over the devices in proportion to tensor_split,
 *                                             boundaries rounded down to 128 rows; MUL_MAT is the only op accepted on it.
 *                                             Matrices too small for a cross-device hand-over to pay (below ~36 MB * n/(n-1),
 *                                             B200_SPLIT_MIN_MB) stay whole on the main device.
 *   "ggml_backend_register_host_buffer"    -> bool (*)(void * buffer, size_t size): pin a host range for DMA uploads
 *   "ggml_backend_unregister_host_buffer"  -> void (*)(void * buffer)      (reference ggml-cuda.cu:2894-2927, :3480-3486; same
 *                                             opt-in: only when GGML_CUDA_REGISTER_HOST or B200_REGISTER_HOST is set)
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#ifdef __cplusplus
extern "C" {
#endif

struct ggml_backend_reg;

/* required by the loader: returns the registry object (api_version == GGML_BACKEND_API_VERSION) */
struct ggml_backend_reg * ggml_backend_init(void);
/* optional: 0 means "do not load" (no sm_100 device visible) */
int ggml_backend_score(void);
/* row boundaries the split buffer type uses for an n_rows matrix over n_dev devices (row0: n_dev + 1 entries);
 * needs no GPU.  Mirrors get_row_split of the reference (ggml-cuda.cu:768-781). */
void ggml_backend_b200_split_rows(long long n_rows, int n_dev, const float * tensor_split, long long * row0);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
