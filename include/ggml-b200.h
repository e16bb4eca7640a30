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
 *   "ggml_backend_split_buffer_type"       -> NULL in this round (row split: SURVEY 8e, next round)
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

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
