// common.cuh — host-side plumbing shared by the C-ABI translation units: error text, launch counting,
// a grow-only device buffer cache and the lazily dlopen'ed NCCL entry points.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <atomic>
#include <map>
#include <string>
#include <vector>
#include "../../include/combatseq.h"

namespace cbs {

extern thread_local char g_err[1024];
extern std::atomic<uint64_t> g_launches;

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

struct CbsError { int code; };   // thrown inside the library only; every extern "C" entry catches it

#define CBS_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            cbs::fail(CBS_ECUDA, "CUDA error %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__,   \
                      __LINE__, #call);                                                             \
            throw cbs::CbsError{CBS_ECUDA};                                                         \
        }                                                                                           \
    } while (0)

// every kernel launch goes through this macro so that cbs_kernel_launches() is the library's own count
#define CBS_LAUNCH(kernel, grid, block, smem, stream, ...)                                          \
    do {                                                                                            \
        cudaEvent_t pe_ = cbs::prof_begin(#kernel, (stream));                                       \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                                 \
        cbs::prof_end(pe_, (stream));                                                               \
        cbs::g_launches.fetch_add(1, std::memory_order_relaxed);                                    \
        CBS_CUDA(cudaGetLastError());                                                               \
    } while (0)

// optional per-kernel device timing (cbs_profile_enable): events on the launching stream around each launch
cudaEvent_t prof_begin(const char* name, cudaStream_t st);   // nullptr when profiling is off
void prof_end(cudaEvent_t begin, cudaStream_t st);

// Grow-only cache of named device buffers (one per device context).  Released by cbs_shutdown().
class DevCache {
  public:
    void* get(const std::string& name, size_t bytes) {
        Slot& s = slots_[name];
        if (s.cap < bytes) {
            if (s.p) cudaFree(s.p);
            s.p = nullptr; s.cap = 0;
            const size_t want = bytes + bytes / 8 + 256;
            cudaError_t e = cudaMalloc(&s.p, want);
            if (e != cudaSuccess) {
                s.p = nullptr;
                fail(CBS_ENOMEM, "cudaMalloc of %zu bytes for '%s' failed: %s", want, name.c_str(), cudaGetErrorString(e));
                throw CbsError{CBS_ENOMEM};
            }
            s.cap = want;
        }
        return s.p;
    }
    template <class T> T* get(const std::string& name, size_t count) { return (T*)get(name, count * sizeof(T)); }
    void release() {
        for (auto& kv : slots_) if (kv.second.p) cudaFree(kv.second.p);
        slots_.clear();
    }
  private:
    struct Slot { void* p = nullptr; size_t cap = 0; };
    std::map<std::string, Slot> slots_;
};

// pinned host staging buffers (same policy)
class PinCache {
  public:
    void* get(const std::string& name, size_t bytes) {
        Slot& s = slots_[name];
        if (s.cap < bytes) {
            if (s.p) cudaFreeHost(s.p);
            s.p = nullptr; s.cap = 0;
            cudaError_t e = cudaMallocHost(&s.p, bytes + 256);
            if (e != cudaSuccess) { s.p = nullptr; fail(CBS_ENOMEM, "cudaMallocHost(%zu) failed", bytes); throw CbsError{CBS_ENOMEM}; }
            s.cap = bytes + 256;
        }
        return s.p;
    }
    void release() {
        for (auto& kv : slots_) if (kv.second.p) cudaFreeHost(kv.second.p);
        slots_.clear();
    }
  private:
    struct Slot { void* p = nullptr; size_t cap = 0; };
    std::map<std::string, Slot> slots_;
};

// ---- NCCL through dlopen: single-GPU use never touches it -----------------------------------------
struct Id128 { char b[128]; };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, /*ncclUniqueId by value, 128 bytes*/ Id128, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi* nccl_api();   // nullptr (and error text set) if libnccl cannot be loaded

}  // namespace cbs

struct cbs_comm {
    void* nccl;        // ncclComm_t
    int rank, world, device;
};

namespace cbs {

// Everything the library keeps per CUDA device: the buffer cache, the stream the stages run on and the set of
// kernels whose function attributes have been set on THIS device (function attributes are per device, so a
// process-wide "done" flag would leave the second device of an in-process multi-GPU call unconfigured).
struct DevCtx {
    int device = -1;
    cudaStream_t st = 0;
    DevCache cache;
    PinCache pin;
    std::map<std::string, cudaEvent_t> events;               // named events (no timing), created on first use
    cudaEvent_t event(const std::string& name) {
        auto it = events.find(name);
        if (it != events.end()) return it->second;
        cudaEvent_t e = nullptr;
        CBS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        events[name] = e;
        return e;
    }
    cudaStream_t st_copy = nullptr;                          // second stream: D2H of finished output columns while kernels run
    cudaStream_t copy_stream() {
        if (!st_copy) CBS_CUDA(cudaStreamCreateWithFlags(&st_copy, cudaStreamNonBlocking));
        return st_copy;
    }
    std::map<std::pair<const void*, int>, int> func_attr;    // (kernel, attribute) -> value set on this device
    void set_func_attr(const void* func, cudaFuncAttribute attr, int value) {
        auto key = std::make_pair(func, (int)attr);
        auto it = func_attr.find(key);
        if (it != func_attr.end() && it->second >= value) return;
        CBS_CUDA(cudaFuncSetAttribute(func, attr, value));
        func_attr[key] = value;
    }
};
extern thread_local DevCtx* t_ctx;        // context of the device this host thread is driving (set by the C-ABI entry)
DevCtx* ctx_for_device(int device);       // creates the context on first use; caller holds the library mutex

}  // namespace cbs
#define g_dev (cbs::t_ctx->cache)
#define g_pin (cbs::t_ctx->pin)
