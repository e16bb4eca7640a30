// ktrace.cu -- host side of the kernel timeline trace (common.cuh: ktrace_event).  Debug tooling: off unless
// b200_ktrace_enable() is called (the plugin does so when B200_KTRACE=<file> is set and dumps on backend free).
#include "common.cuh"
#include "../../include/b200_ops.h"

#include <cstdio>
#include <vector>

#ifndef B200_KTRACE
extern "C" int b200_ktrace_enable(void) { return -1; }            // not compiled in (see common.cuh)
extern "C" int b200_ktrace_dump(const char *) { return -1; }
#else
namespace b200 {
static std::vector<int (*)(unsigned long long *)> & setters() { static std::vector<int (*)(unsigned long long *)> v; return v; }
void ktrace_register(int (*setter)(unsigned long long *)) { setters().push_back(setter); }
static unsigned long long * g_ktrace_buf = nullptr;
}  // namespace b200

using namespace b200;

extern "C" int b200_ktrace_enable(void) {
    if (g_ktrace_buf) return 0;
    const size_t bytes = (1 + 2 * KTRACE_CAP) * sizeof(unsigned long long);
    B200_CUDA_CHECK(cudaMalloc(&g_ktrace_buf, bytes));
    B200_CUDA_CHECK(cudaMemset(g_ktrace_buf, 0, bytes));
    for (auto f : setters()) { const int rc = f(g_ktrace_buf); if (rc) return rc; }
    return 0;
}

// text, one event per line: "<ns> <gridid> <event 0=entry 1=pdl-wait-passed 2..15=kernel phase marks> <ctas> <threads>"
extern "C" int b200_ktrace_dump(const char * path) {
    if (!g_ktrace_buf || !path) return -1;
    B200_CUDA_CHECK(cudaDeviceSynchronize());
    unsigned long long n = 0;
    B200_CUDA_CHECK(cudaMemcpy(&n, g_ktrace_buf, sizeof(n), cudaMemcpyDeviceToHost));
    if (n > KTRACE_CAP) n = KTRACE_CAP;
    std::vector<unsigned long long> ev(2 * n);
    if (n) B200_CUDA_CHECK(cudaMemcpy(ev.data(), g_ktrace_buf + 1, 2 * n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    FILE * f = fopen(path, "w");
    if (!f) return -2;
    for (unsigned long long i = 0; i < n; ++i) {
        const unsigned long long w = ev[2 * i + 1];
        fprintf(f, "%llu %llu %d %d %d\n", ev[2 * i], w >> 32, (int)(w & 15), (int)((w >> 15) & 0x1FFFF), (int)((w >> 4) & 0x7FF));
    }
    fclose(f);
    return 0;
}
#endif
