// ggml-b200.cu -- the ggml backend plugin: libggml-b200.so.
//
// Implements the ggml_backend_reg / ggml_backend_device_i / ggml_backend_buffer_type_i /
// ggml_backend_buffer_i / ggml_backend_i C ABI (reference ggml/src/ggml-backend-impl.h:17-207) so
// that unmodified libllama, ggml_backend_sched, llama-bench and test-backend-ops load it through
// GGML_BACKEND_PATH (ggml/src/ggml-backend-reg.cpp:586-590).  It replaces the boundary objects of
// the reference's ggml-cuda.cu (:513-751 buffers, :2405-2867 backend, :2871-3395 device, :3399-3538
// registry); every compute node is forwarded to the plain-C operator ABI of include/b200_ops.h.
// No CPU fallback and no dispatch to other backends: supports_op is an honest whitelist.
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"

#include "../../../include/b200_ops.h"
#include "../../../include/ggml-b200.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#define B200_MAX_DEVICES 16
#define B200_CHECK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    GGML_LOG_ERROR("ggml-b200: %s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    GGML_ABORT("ggml-b200: CUDA error"); } } while (0)

namespace {

struct dev_ctx {
    int device = 0;
    std::string name, description;
    size_t total = 0;
    ggml_backend_buffer_type buft;       // device memory
    std::string buft_name;
};

struct reg_ctx {
    std::vector<ggml_backend_device> devices;      // stable addresses after init
    std::vector<dev_ctx *> dctx;
    ggml_backend_buffer_type host_buft;
};

reg_ctx * g_reg = nullptr;
ggml_backend_reg g_reg_obj;

void set_device(int device) {
    int cur = -1;
    B200_CHECK(cudaGetDevice(&cur));
    if (cur != device) B200_CHECK(cudaSetDevice(device));
}

// ------------------------------------------------------------------------------------------------
// device buffer
// ------------------------------------------------------------------------------------------------
struct buf_ctx { int device; void * base; };

void upload_fence(int device, cudaStream_t stream, bool host_wait);
void buf_free(ggml_backend_buffer_t buffer) {
    buf_ctx * c = (buf_ctx *)buffer->context;
    set_device(c->device);
    upload_fence(c->device, nullptr, true);
    cudaFree(c->base);
    delete c;
}
void * buf_get_base(ggml_backend_buffer_t buffer) { return ((buf_ctx *)buffer->context)->base; }
enum ggml_status buf_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    (void)buffer; (void)tensor;      // no extras: kernels read GGUF blocks in place; rows are never over-read
    return GGML_STATUS_SUCCESS;
}
void upload_fence(int device, cudaStream_t stream, bool host_wait);
void buf_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    buf_ctx * c = (buf_ctx *)buffer->context;
    set_device(c->device);
    upload_fence(c->device, nullptr, true);
    B200_CHECK(cudaMemsetAsync((char *)tensor->data + offset, value, size, cudaStreamPerThread));
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}
// ---- weight upload (reference set_tensor: one synchronous cudaMemcpyAsync from pageable memory per tensor,
// ggml-cuda.cu:604-611) ----
// The loader hands us pointers into the mmap'd GGUF (pageable, often not yet faulted in).  A pageable cudaMemcpy makes the driver
// copy through its own small bounce buffers with the CPU copy and the DMA taking turns.  Here: two pinned staging buffers per device
// and an upload stream; while the DMA engine drains one buffer the calling thread fills the other (that memcpy is also what faults
// the file pages in), so disk/page-cache reads and PCIe transfers overlap.  Still synchronous at the interface: the call returns
// when the last chunk has landed.  B200_STAGED_UPLOAD=0 restores the plain copy; B200_UPLOAD_CHUNK_MB sets the chunk (default 16).
struct upload_lane {
    std::mutex mu;
    void * pin[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr};
    cudaStream_t stream = nullptr;
    size_t chunk = 0;
    bool failed = false;
};
upload_lane g_upload[B200_MAX_DEVICES];
std::atomic<long long> g_upload_bytes{0}, g_upload_us{0};

bool staged_upload(int device, void * dst, const void * src, size_t size) {
    static const bool on = [] { const char * e = getenv("B200_STAGED_UPLOAD"); return !e || atoi(e) != 0; }();
    static const size_t chunk_mb = [] { const char * e = getenv("B200_UPLOAD_CHUNK_MB"); const long v = e ? atol(e) : 16; return (size_t)(v < 1 ? 1 : v > 256 ? 256 : v); }();
    if (!on || size < (size_t)4 << 20 || device < 0 || device >= B200_MAX_DEVICES) return false;
    upload_lane & L = g_upload[device];
    std::lock_guard<std::mutex> lock(L.mu);
    if (L.failed) return false;
    if (!L.stream) {
        L.chunk = chunk_mb << 20;
        if (cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking) != cudaSuccess || cudaMallocHost(&L.pin[0], L.chunk) != cudaSuccess ||
            cudaMallocHost(&L.pin[1], L.chunk) != cudaSuccess || cudaEventCreateWithFlags(&L.done[0], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&L.done[1], cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            L.failed = true;      // no pinned memory to be had: the plain path still works
            return false;
        }
    }
    size_t off = 0;
    int b = 0;
    bool used[2] = {false, false};
    while (off < size) {
        const size_t n = size - off < L.chunk ? size - off : L.chunk;
        if (used[b]) B200_CHECK(cudaEventSynchronize(L.done[b]));      // the DMA out of this buffer has finished
        memcpy(L.pin[b], (const char *)src + off, n);
        B200_CHECK(cudaMemcpyAsync((char *)dst + off, L.pin[b], n, cudaMemcpyHostToDevice, L.stream));
        B200_CHECK(cudaEventRecord(L.done[b], L.stream));
        used[b] = true;
        off += n; b ^= 1;
    }
    B200_CHECK(cudaStreamSynchronize(L.stream));
    return true;
}
// Small uploads (the token ids, positions and the mask of every decode step: 3-5 set_tensor calls per token) need not wait for
// their DMA.  A copy from pageable memory returns once the driver has taken the bytes into its own staging area -- the caller's
// buffer is free, which is all "synchronous" means to the caller -- so the stream synchronisation that followed it (10-20 us each,
// with the GPU idle) is replaced by ordering: the copies go to one upload stream per device, and whatever touches the device next
// on a backend stream or reads a tensor back first waits for that stream's last event (upload_fence).  B200_ASYNC_SMALL_UPLOAD=0
// restores the synchronising copy.
struct small_lane { std::mutex mu; cudaStream_t stream = nullptr; cudaEvent_t ev = nullptr; std::atomic<bool> pending{false}; bool failed = false; };
small_lane g_small[B200_MAX_DEVICES];
bool small_upload(int device, void * dst, const void * src, size_t size) {
    static const bool on = [] { const char * e = getenv("B200_ASYNC_SMALL_UPLOAD"); return !e || atoi(e) != 0; }();
    if (!on || size > ((size_t)256 << 10) || device < 0 || device >= B200_MAX_DEVICES) return false;
    small_lane & L = g_small[device];
    std::lock_guard<std::mutex> lock(L.mu);
    if (L.failed) return false;
    if (!L.stream) {
        if (cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&L.ev, cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError(); L.failed = true; return false;
        }
    }
    B200_CHECK(cudaMemcpyAsync(dst, src, size, cudaMemcpyHostToDevice, L.stream));      // pageable source: staged by the driver before this returns
    B200_CHECK(cudaEventRecord(L.ev, L.stream));
    L.pending.store(true, std::memory_order_release);
    return true;
}
// make `stream` (nullptr: the calling host thread) wait for the uploads still in flight to this device
void upload_fence(int device, cudaStream_t stream, bool host_wait) {
    if (device < 0 || device >= B200_MAX_DEVICES) return;
    small_lane & L = g_small[device];
    if (!L.pending.load(std::memory_order_acquire)) return;
    std::lock_guard<std::mutex> lock(L.mu);
    if (host_wait) { B200_CHECK(cudaEventSynchronize(L.ev)); L.pending.store(false, std::memory_order_release); }
    else B200_CHECK(cudaStreamWaitEvent(stream, L.ev, 0));
}
void buf_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    buf_ctx * c = (buf_ctx *)buffer->context;
    set_device(c->device);
    if (small_upload(c->device, (char *)tensor->data + offset, data, size)) return;
    upload_fence(c->device, nullptr, true);
    if (staged_upload(c->device, (char *)tensor->data + offset, data, size)) return;
    B200_CHECK(cudaMemcpyAsync((char *)tensor->data + offset, data, size, cudaMemcpyHostToDevice, cudaStreamPerThread));
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}
void buf_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    buf_ctx * c = (buf_ctx *)buffer->context;
    set_device(c->device);
    upload_fence(c->device, nullptr, true);
    B200_CHECK(cudaMemcpyAsync(data, (const char *)tensor->data + offset, size, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}
bool buffer_is_b200(ggml_backend_buffer_t buffer);
bool buf_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    if (!src->buffer || !buffer_is_b200(src->buffer)) return false;
    buf_ctx * sc = (buf_ctx *)src->buffer->context;
    buf_ctx * dc = (buf_ctx *)buffer->context;
    upload_fence(sc->device, nullptr, true);
    set_device(dc->device);
    upload_fence(dc->device, nullptr, true);
    if (sc->device == dc->device) {
        B200_CHECK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    } else {
        B200_CHECK(cudaMemcpyPeerAsync(dst->data, dc->device, src->data, sc->device, ggml_nbytes(src), cudaStreamPerThread));
    }
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}
void buf_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    buf_ctx * c = (buf_ctx *)buffer->context;
    set_device(c->device);
    upload_fence(c->device, nullptr, true);
    B200_CHECK(cudaMemsetAsync(c->base, value, buffer->size, cudaStreamPerThread));
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}
const ggml_backend_buffer_i buf_iface = {
    /* .free_buffer   = */ buf_free,
    /* .get_base      = */ buf_get_base,
    /* .init_tensor   = */ buf_init_tensor,
    /* .memset_tensor = */ buf_memset_tensor,
    /* .set_tensor    = */ buf_set_tensor,
    /* .get_tensor    = */ buf_get_tensor,
    /* .cpy_tensor    = */ buf_cpy_tensor,
    /* .clear         = */ buf_clear,
    /* .reset         = */ nullptr,
};
bool buffer_is_b200(ggml_backend_buffer_t buffer) { return buffer->iface.free_buffer == buf_free; }

// ------------------------------------------------------------------------------------------------
// device buffer type
// ------------------------------------------------------------------------------------------------
const char * buft_get_name(ggml_backend_buffer_type_t buft) { return ((dev_ctx *)buft->context)->buft_name.c_str(); }
ggml_backend_buffer_t buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    dev_ctx * d = (dev_ctx *)buft->context;
    set_device(d->device);
    void * p = nullptr;
    const size_t sz = size > 0 ? size : 1;
    cudaError_t e = cudaMalloc(&p, sz);
    if (e != cudaSuccess) {
        cudaGetLastError();
        GGML_LOG_ERROR("ggml-b200: allocating %.2f MiB on device %d failed: %s\n", sz / 1048576.0, d->device, cudaGetErrorString(e));
        return nullptr;
    }
    return ggml_backend_buffer_init(buft, buf_iface, new buf_ctx{d->device, p}, size);
}
size_t buft_get_alignment(ggml_backend_buffer_type_t) { return 128; }
size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    size_t size = ggml_nbytes(tensor);
    // GEMV bulk copies of rows that are not 16-byte aligned (views) fetch the enclosing 16-byte
    // windows: keep a little slack behind quantised tensors so the window never leaves the buffer
    if (ggml_is_quantized(tensor->type)) size += 128;
    return size;
}
bool buft_is_host(ggml_backend_buffer_type_t) { return false; }
const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};
bool buft_is_b200(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == buft_get_name; }

// ------------------------------------------------------------------------------------------------
// pinned host buffer type (reference ggml-cuda.cu:1090-1170): a CPU buffer over cudaMallocHost memory
// ------------------------------------------------------------------------------------------------
const char * host_buft_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
void host_buf_free(ggml_backend_buffer_t buffer) { cudaFreeHost(buffer->context); }
ggml_backend_buffer_t host_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    void * p = nullptr;
    cudaError_t e = cudaMallocHost(&p, size > 0 ? size : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        GGML_LOG_WARN("ggml-b200: pinned allocation of %.2f MiB failed (%s), using pageable memory\n", size / 1048576.0, cudaGetErrorString(e));
        return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);
    }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(p, size);
    buffer->buft = buft;
    buffer->iface.free_buffer = host_buf_free;
    return buffer;
}
size_t host_buft_alignment(ggml_backend_buffer_type_t) { return 64; }
bool host_buft_is_host(ggml_backend_buffer_type_t) { return true; }

}  // namespace

#include "ggml-b200-split.inc"
#include "ggml-b200-compute.inc"
#include "ggml-b200-device.inc"
