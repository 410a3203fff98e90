// combatseq.cu — the C ABI of libcombatseq.so (include/combatseq.h): host orchestration of the sm_100a kernels.
//
// Drop-in numeric body of ComBat_seq() (/root/reference/ComBat_seq.R:33-203) and of its helpers
// monte_carlo_int_NB / match_quantiles (/root/reference/helper_seq.R:8-64).  There is no CPU path: without a
// CUDA device every compute entry returns CBS_ENOGPU.
#include "common.cuh"
#include "host_design.hpp"
#include "kernels.cuh"
#include "mc.cuh"
#include "mc_gemm.cuh"
#include <dlfcn.h>
#include <chrono>
#include <mutex>
#include <numeric>

namespace cbs {

thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launches{0};
static std::mutex g_mutex;          // one call at a time per process (the R caller is single-threaded anyway)
static DevCache g_dev;
static PinCache g_pin;

// ---- tuning knobs (cbs_set_tuning) ----------------------------------------------------------------------------
static int g_mc_algo = 0;            // 0 = auto (tensor-core contraction when gene.subset.n is NULL), 1 = direct, 2 = contraction
static int g_glm_scr = 1;            // 1 = Levenberg work arrays of the lane-group GLM kernel in shared memory (one copy per gene)
static int g_mq_rows = 3;            // match_quantiles: 3 = warp per gene, out-of-line continued fraction (fastest, same bits); 2 = same, inlined;
                                     // 1 = 8x32 tiles via shared memory, inlined; 4 = tiles, out-of-line
static int g_apl_fast = 1;           // 1 = lgamma(y+1) table + hoisted log(r)/lgamma(r) in the APL kernel (same bits), 0 = per-sample calls
static double g_mc_tail_tol = 1e-8;  // largest accepted Chebyshev tail estimate (log-likelihood units)
static double g_last_mc_tail = 0;    // diagnostics: largest tail estimate of the last Monte-Carlo stage
static int g_last_mc_kc = 0;         // diagnostics: Chebyshev terms used by the last contraction launch


static double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------------
// NCCL (dlopen).  Under torchrun the torch-bundled libnccl.so.2 is already mapped and is the one resolved.
NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (api.handle) return &api;
    if (tried) return nullptr;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (api.handle) break; }
    if (!api.handle) { fail(CBS_ENCCL, "cannot dlopen libnccl.so.2: %s", dlerror()); return nullptr; }
    auto sym = [&](const char* s) { return dlsym(api.handle, s); };
    api.GetUniqueId = (int (*)(void*))sym("ncclGetUniqueId");
    api.CommInitRank = (int (*)(void**, int, Id128, int))sym("ncclCommInitRank");
    api.CommDestroy = (int (*)(void*))sym("ncclCommDestroy");
    api.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))sym("ncclAllGather");
    api.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))sym("ncclAllReduce");
    api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllGather || !api.AllReduce) {
        fail(CBS_ENCCL, "libnccl is missing required symbols");
        dlclose(api.handle); api.handle = nullptr;
        return nullptr;
    }
    return &api;
}
#define CBS_NCCL(call)                                                                              \
    do {                                                                                            \
        int r_ = (call);                                                                            \
        if (r_ != 0) {                                                                              \
            NcclApi* a_ = nccl_api();                                                               \
            fail(CBS_ENCCL, "NCCL error %d (%s) at %s:%d", r_,                                      \
                 (a_ && a_->GetErrorString) ? a_->GetErrorString(r_) : "?", __FILE__, __LINE__);    \
            throw CbsError{CBS_ENCCL};                                                              \
        }                                                                                           \
    } while (0)
constexpr int NCCL_INT8 = 0, NCCL_UINT64 = 5, NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2;   // ncclDataType_t / ncclRedOp_t values (nccl.h)

static bool have_gpu() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
    return n > 0;
}
#define CBS_NEED_GPU() do { if (!have_gpu()) return fail(CBS_ENOGPU, "no CUDA device visible: libcombatseq has no CPU path"); } while (0)

// ---------------------------------------------------------------------------------------------------
// Device copy of a HostDesign: one packed allocation per design (a few KB), tracked for release after the call.
struct DesignSet {
    std::vector<void*> allocs;
    DesignDev* d_array = nullptr;      // device array of DesignDev
    std::vector<DesignDev> h_array;
    ~DesignSet() { for (void* p : allocs) cudaFree(p); }
    void add(const HostDesign& h) {
        auto up = [&](const void* src, size_t bytes) -> void* {
            void* p = nullptr;
            CBS_CUDA(cudaMalloc(&p, bytes ? bytes : 8));
            allocs.push_back(p);
            if (bytes) CBS_CUDA(cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice));
            return p;
        };
        DesignDev d;
        d.n = h.n; d.p = h.p; d.oneway = h.oneway; d.ncls = h.ncls;
        d.col = (const int*)up(h.col.data(), h.col.size() * sizeof(int));
        d.cell_start = (const int*)up(h.cell_start.data(), h.cell_start.size() * sizeof(int));
        d.minv = (const double*)up(h.minv.data(), h.minv.size() * sizeof(double));
        d.rowptr = (const int*)up(h.rowptr.data(), h.rowptr.size() * sizeof(int));
        d.cidx = (const int*)up(h.cidx.data(), h.cidx.size() * sizeof(int));
        d.cval = (const double*)up(h.cval.data(), h.cval.size() * sizeof(double));
        d.start = (const double*)up(h.start.data(), h.start.size() * sizeof(double));
        h_array.push_back(d);
    }
    void finish() {
        void* p = nullptr;
        CBS_CUDA(cudaMalloc(&p, h_array.size() * sizeof(DesignDev) + 8));
        allocs.push_back(p);
        CBS_CUDA(cudaMemcpy(p, h_array.data(), h_array.size() * sizeof(DesignDev), cudaMemcpyHostToDevice));
        d_array = (DesignDev*)p;
    }
};

template <class T>
static T* upload(DevCache& c, const std::string& name, const T* src, size_t count, cudaStream_t st = 0) {
    T* p = c.get<T>(name, count ? count : 1);
    if (count) CBS_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    return p;
}

static inline dim3 grid1(long n, int block) { return dim3((unsigned)((n + block - 1) / block)); }

}  // namespace cbs

using namespace cbs;

#define CBS_ENTRY_BEGIN std::lock_guard<std::mutex> lock_(g_mutex); try {
#define CBS_ENTRY_END                                                                               \
    } catch (const CbsError& e) { cudaDeviceSynchronize(); cudaGetLastError(); return e.code; }     \
      catch (const std::bad_alloc&) { return fail(CBS_ENOMEM, "host allocation failed"); }          \
      catch (const std::exception& e) { return fail(CBS_EINVAL, "internal error: %s", e.what()); }  \
    return CBS_OK;

extern "C" {

int cbs_version(void) { return CBS_VERSION; }
int cbs_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
const char* cbs_last_error(void) { return g_err; }
uint64_t cbs_kernel_launches(void) { return g_launches.load(); }
int cbs_shutdown(void) {
    std::lock_guard<std::mutex> lock_(g_mutex);
    if (have_gpu()) { cudaDeviceSynchronize(); g_dev.release(); g_pin.release(); }
    return CBS_OK;
}

int cbs_comm_unique_id(void* id128) {
    NcclApi* a = nccl_api();
    if (!a) return CBS_ENCCL;
    int r = a->GetUniqueId(id128);
    return r == 0 ? CBS_OK : fail(CBS_ENCCL, "ncclGetUniqueId failed (%d)", r);
}
int cbs_comm_create(const void* id128, int rank, int world, int device, cbs_comm** out) {
    CBS_NEED_GPU();
    NcclApi* a = nccl_api();
    if (!a) return CBS_ENCCL;
    if (!out || world < 1 || rank < 0 || rank >= world) return fail(CBS_EINVAL, "bad rank/world");
    if (cudaSetDevice(device) != cudaSuccess) return fail(CBS_ECUDA, "cudaSetDevice(%d) failed", device);
    Id128 id;
    memcpy(&id, id128, 128);
    void* comm = nullptr;
    int r = a->CommInitRank(&comm, world, id, rank);
    if (r != 0) return fail(CBS_ENCCL, "ncclCommInitRank failed (%d: %s)", r, a->GetErrorString ? a->GetErrorString(r) : "?");
    *out = new cbs_comm{comm, rank, world, device};
    return CBS_OK;
}
int cbs_comm_destroy(cbs_comm* c) {
    if (!c) return CBS_OK;
    NcclApi* a = nccl_api();
    if (a && c->nccl) a->CommDestroy(c->nccl);
    delete c;
    return CBS_OK;
}

}  // extern "C"

#include "profile.inl"
#include "api_subsystems.inl"
#include "api_pipeline.inl"
