// profile.inl — measurement hooks: per-kernel CUDA-event timing and the FP64 FMA peak microbenchmark.
namespace cbs {

struct ProfRec { std::string name; cudaEvent_t e0, e1; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_pending;
static std::map<std::string, std::pair<double, long>> g_prof_acc;     // name -> (total ms, launches)
static std::vector<cudaEvent_t> g_prof_pool;
static std::mutex g_prof_mutex;       // launches come from one host thread per device in single-process multi-GPU calls

static cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
cudaEvent_t prof_begin(const char* name, cudaStream_t st) {
    if (!g_prof_on) return nullptr;
    std::lock_guard<std::mutex> lk(g_prof_mutex);
    ProfRec r{name, prof_event(), prof_event()};
    cudaEventRecord(r.e0, st);
    g_prof_pending.push_back(r);
    return r.e0;
}
void prof_end(cudaEvent_t begin, cudaStream_t st) {
    if (!begin) return;
    std::lock_guard<std::mutex> lk(g_prof_mutex);
    for (auto it = g_prof_pending.rbegin(); it != g_prof_pending.rend(); ++it)
        if (it->e0 == begin) { cudaEventRecord(it->e1, st); break; }
}
static void prof_collect() {
    if (g_prof_pending.empty()) return;
    cudaDeviceSynchronize();
    for (auto& r : g_prof_pending) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
            auto& a = g_prof_acc[r.name];
            a.first += ms; a.second += 1;
        }
        g_prof_pool.push_back(r.e0); g_prof_pool.push_back(r.e1);
    }
    g_prof_pending.clear();
    cudaGetLastError();
}

// 8 independent FMA chains per thread, 4096 FMAs per chain
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
    for (int i = 0; i < 512; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 123.456) out[0] = s;
}

}  // namespace cbs

extern "C" {

int cbs_set_tuning(const char* key, double value) {
    std::lock_guard<std::mutex> lock_(g_mutex);
    if (!key) return fail(CBS_EINVAL, "null key");
    if (!strcmp(key, "mc_algo")) { if (value < 0 || value > 2) return fail(CBS_EINVAL, "mc_algo must be 0, 1 or 2"); g_tune.mc_algo = (int)value; return CBS_OK; }
    if (!strcmp(key, "mc_band")) { g_tune.mc_band = value != 0; return CBS_OK; }
    if (!strcmp(key, "mc_tail_tol")) { g_tune.mc_tail_tol = value; return CBS_OK; }
    if (!strcmp(key, "apl_fast")) { g_tune.apl_fast = value != 0; return CBS_OK; }
    if (!strcmp(key, "mq_bins")) { g_tune.mq_bins = value != 0; return CBS_OK; }
    if (!strcmp(key, "apl_lanes")) { g_tune.apl_lanes = value != 0; return CBS_OK; }
    if (!strcmp(key, "glm_hist")) {
        g_tune.glm_hist = value != 0;
        if (g_tune.glm_hist && have_gpu()) {
            try { enter_device(-1); CBS_CUDA(cudaMemset(g_dev.get<int>("glm_iter_hist", 256), 0, 256 * sizeof(int))); }
            catch (const CbsError& e) { return e.code; }
        }
        return CBS_OK;
    }
    if (!strcmp(key, "glm_scr")) { g_tune.glm_scr = value != 0; return CBS_OK; }
    if (!strcmp(key, "glm_two_pass")) { g_tune.glm_two_pass = value != 0; return CBS_OK; }
    return fail(CBS_EINVAL, "unknown tuning key '%s'", key);
}
int cbs_get_diag(const char* key, double* value) {
    std::lock_guard<std::mutex> lock_(g_mutex);
    if (!key || !value) return fail(CBS_EINVAL, "null pointer");
    if (!strcmp(key, "mc_tail")) { *value = g_diag.mc_tail; return CBS_OK; }
    if (!strcmp(key, "mc_terms")) { *value = g_diag.mc_kc; return CBS_OK; }
    if (!strcmp(key, "mc_tiles_done")) { *value = g_diag.mc_tiles_done; return CBS_OK; }
    if (!strcmp(key, "mc_tiles_all")) { *value = g_diag.mc_tiles_all; return CBS_OK; }
    if (!strncmp(key, "glm_iter_", 9)) {        // "glm_iter_<k>": genes whose Levenberg fit took k iterations (k = 255: 255 or more), summed over the launches
        const int k = atoi(key + 9);
        if (k < 0 || k > 255 || !g_tune.glm_hist || !have_gpu()) return fail(CBS_EINVAL, "glm_iter_<0..255> needs the glm_hist knob");
        try {
            enter_device(-1);
            int h = 0;
            CBS_CUDA(cudaMemcpy(&h, g_dev.get<int>("glm_iter_hist", 256) + k, sizeof(int), cudaMemcpyDeviceToHost));
            *value = h;
        } catch (const CbsError& e) { return e.code; }
        return CBS_OK;
    }
    if (!strcmp(key, "reset")) { g_diag = Diag(); *value = 0; return CBS_OK; }
    return fail(CBS_EINVAL, "unknown diagnostics key '%s'", key);
}

int cbs_fp64_mma_peak(double* tflops) {
    CBS_NEED_GPU();
    if (!tflops) return fail(CBS_EINVAL, "null pointer");
    CBS_ENTRY_BEGIN
    *tflops = fp64_mma_peak_tflops();
    CBS_ENTRY_END
}

int cbs_profile_enable(int on) {
    std::lock_guard<std::mutex> lock_(g_mutex);
    if (on) { prof_collect(); g_prof_acc.clear(); }
    else prof_collect();
    g_prof_on = on != 0;
    return CBS_OK;
}
int cbs_profile_dump(char* buf, size_t len) {
    std::lock_guard<std::mutex> lock_(g_mutex);
    if (!buf || len == 0) return fail(CBS_EINVAL, "null buffer");
    prof_collect();
    std::string s;
    char line[256];
    for (auto& kv : g_prof_acc) {
        snprintf(line, sizeof line, "%s %.6f %ld\n", kv.first.c_str(), kv.second.first, kv.second.second);
        s += line;
    }
    snprintf(buf, len, "%s", s.c_str());
    return CBS_OK;
}
int cbs_fp64_peak(double* tflops) {
    CBS_NEED_GPU();
    if (!tflops) return fail(CBS_EINVAL, "null pointer");
    CBS_ENTRY_BEGIN
    int dev = 0, sms = 0;
    CBS_CUDA(cudaGetDevice(&dev));
    CBS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    double* d = g_dev.get<double>("fp64_peak", 8);
    const int blocks = sms * 32;
    cudaEvent_t e0, e1;
    CBS_CUDA(cudaEventCreate(&e0)); CBS_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        CBS_CUDA(cudaEventRecord(e0, 0));
        CBS_LAUNCH(k_fp64_peak, dim3(blocks), 256, 0, 0, d, 0.999999, 1e-6);
        CBS_CUDA(cudaEventRecord(e1, 0));
        CBS_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        CBS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flop = 2.0 * 8 * 8 * 512 * 256.0 * blocks;
        if (rep > 0) best = std::max(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *tflops = best;
    CBS_ENTRY_END
}

}  // extern "C"
