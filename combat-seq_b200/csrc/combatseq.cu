// combatseq.cu — the C ABI of libcombatseq.so (include/combatseq.h): host orchestration of the sm_100a kernels.
//
// Drop-in numeric body of ComBat_seq() (/root/reference/ComBat_seq.R:33-203) and of its helpers
// monte_carlo_int_NB / match_quantiles (/root/reference/helper_seq.R:8-64).  There is no CPU path: without a
// CUDA device every compute entry returns CBS_ENOGPU.
#include "engine.hpp"
#include "prep_kernels.cuh"
#include "host_bridge.hpp"
#include <dlfcn.h>
#include <chrono>
#include <mutex>
#include <thread>
#include <memory>

namespace cbs {

thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launches{0};
static std::mutex g_mutex;          // one call at a time per process (the R caller is single-threaded anyway)
Tuning g_tune;
Diag g_diag;
thread_local DevCtx* t_ctx = nullptr;
static std::map<int, std::unique_ptr<DevCtx>> g_ctx;        // one context per device ordinal (guarded by g_mutex)

void release_inproc_comms();

DevCtx* ctx_for_device(int device) {
    auto& p = g_ctx[device];
    if (!p) { p.reset(new DevCtx); p->device = device; }
    return p.get();
}
// make `device` (or the current device when < 0) current and select its context for this thread
static void enter_device(int device) {
    if (device >= 0) CBS_CUDA(cudaSetDevice(device));
    CBS_CUDA(cudaGetDevice(&device));
    t_ctx = ctx_for_device(device);
}

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
    // COMBATSEQ_NCCL names the library to load first.  It matters when the process maps another libnccl.so.2 LATER (PyTorch
    // bundles its own, newer one): the dynamic loader resolves that later request to whatever libnccl.so.2 is already mapped,
    // so loading an older system NCCL here first makes a subsequent `import torch` fail on a missing symbol.
    const char* names[] = {getenv("COMBATSEQ_NCCL"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { if (!n || !*n) continue; api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (api.handle) break; }
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
static bool have_gpu() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
    return n > 0;
}
#define CBS_NEED_GPU() do { if (!have_gpu()) return fail(CBS_ENOGPU, "no CUDA device visible: libcombatseq has no CPU path"); } while (0)

// ---------------------------------------------------------------------------------------------------
void DesignSet::finish() {
    std::vector<char> blob;
    auto put = [&](const void* src, size_t bytes) -> size_t {        // returns the offset of the copy inside the blob
        const size_t off = (blob.size() + 15) / 16 * 16;
        blob.resize(off + bytes);
        if (bytes) memcpy(blob.data() + off, src, bytes);
        return off;
    };
    struct Offs { size_t col, cell_start, minv, rowptr, cidx, cval, start; };
    std::vector<Offs> offs;
    for (const HostDesign& h : host)
        offs.push_back({put(h.col.data(), h.col.size() * sizeof(int)), put(h.cell_start.data(), h.cell_start.size() * sizeof(int)),
                        put(h.minv.data(), h.minv.size() * sizeof(double)), put(h.rowptr.data(), h.rowptr.size() * sizeof(int)),
                        put(h.cidx.data(), h.cidx.size() * sizeof(int)), put(h.cval.data(), h.cval.size() * sizeof(double)),
                        put(h.start.data(), h.start.size() * sizeof(double))});
    const size_t arr_off = (blob.size() + 15) / 16 * 16;
    blob.resize(arr_off + host.size() * sizeof(DesignDev) + 16);
    char* dev = (char*)g_dev.get("designs:" + name, blob.size());
    h_array.clear();
    for (size_t i = 0; i < host.size(); ++i) {
        const HostDesign& h = host[i];
        DesignDev d;
        d.n = h.n; d.p = h.p; d.oneway = h.oneway; d.ncls = h.ncls;
        d.col = (const int*)(dev + offs[i].col);
        d.cell_start = (const int*)(dev + offs[i].cell_start);
        d.minv = (const double*)(dev + offs[i].minv);
        d.rowptr = (const int*)(dev + offs[i].rowptr);
        d.cidx = (const int*)(dev + offs[i].cidx);
        d.cval = (const double*)(dev + offs[i].cval);
        d.start = (const double*)(dev + offs[i].start);
        h_array.push_back(d);
    }
    memcpy(blob.data() + arr_off, h_array.data(), h_array.size() * sizeof(DesignDev));
    CBS_CUDA(cudaMemcpyAsync(dev, blob.data(), blob.size(), cudaMemcpyHostToDevice, t_ctx->st));
    CBS_CUDA(cudaStreamSynchronize(t_ctx->st));          // the blob is a local
    d_array = (DesignDev*)(dev + arr_off);
}

void shard_range(long G, int rank, int world, long* g0, long* g1) {
    long per = (G + world - 1) / world;
    per = (per + APL_CHUNK - 1) / APL_CHUNK * APL_CHUNK;
    *g0 = std::min(G, (long)rank * per);
    *g1 = std::min(G, (long)(rank + 1) * per);
}

}  // namespace cbs

using namespace cbs;

#define CBS_ENTRY_BEGIN std::lock_guard<std::mutex> lock_(g_mutex); try { enter_device(-1);
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
    release_inproc_comms();
    if (have_gpu()) {
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto& kv : g_ctx) {
            cudaSetDevice(kv.first);
            cudaDeviceSynchronize();
            kv.second->cache.release(); kv.second->pin.release();
            for (auto& e : kv.second->events) cudaEventDestroy(e.second);
            kv.second->events.clear();
            if (kv.second->st_copy) { cudaStreamDestroy(kv.second->st_copy); kv.second->st_copy = nullptr; }
        }
        cudaSetDevice(cur);
    }
    g_ctx.clear();
    t_ctx = nullptr;
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
