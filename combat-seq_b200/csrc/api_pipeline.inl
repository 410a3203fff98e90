// api_pipeline.inl — cbs_combat_seq(): the whole numeric body of ComBat_seq() (ComBat_seq.R:33-203) on the device.
// Included by combatseq.cu.

namespace cbs {

// rows [g0,g1) x ncols of a G x ncols column-major matrix <-> a contiguous [ncols][per] block (per = shard size)
__global__ void k_pack_rows(const double* __restrict__ m, long G, long g0, long g1, long per, int ncols, double* __restrict__ buf) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    if (g0 + r < g1) buf[(long)c * per + r] = m[g0 + r + (long)c * G];
}
__global__ void k_unpack_rows(const double* __restrict__ buf, long G, long per, int ncols, int world, double* __restrict__ m) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    if (g >= G) return;
    const long rk = g / per, r = g - rk * per;
    m[g + (long)c * G] = buf[((long)rk * ncols + c) * per + r];
}

// all-gather the row shards of a G x ncols matrix (every rank ends up with the whole matrix)
static void allgather_rows(double* d_m, long G, int ncols, long g0, long g1, cbs_comm* comm, cudaStream_t st) {
    if (!comm || comm->world <= 1) return;
    long per = (G + comm->world - 1) / comm->world;
    per = (per + APL_CHUNK - 1) / APL_CHUNK * APL_CHUNK;
    const size_t blk = (size_t)per * ncols;
    double* send = g_dev.get<double>("ag_send", blk);
    double* recv = g_dev.get<double>("ag_recv", blk * comm->world);
    if (g1 > g0) CBS_LAUNCH(k_pack_rows, dim3((unsigned)((g1 - g0 + 255) / 256), (unsigned)ncols), 256, 0, st, d_m, G, g0, g1, per, ncols, send);
    CBS_NCCL(nccl_api()->AllGather(send, recv, blk, NCCL_FLOAT64, comm->nccl, st));
    CBS_LAUNCH(k_unpack_rows, dim3((unsigned)((G + 255) / 256), (unsigned)ncols), 256, 0, st, recv, G, per, ncols, comm->world, d_m);
}

// scatter a per-gene G x ncols matrix into G0-row layout with NaN rows for dropped genes
__global__ void k_scatter_rows(const double* __restrict__ src, long G, const int* __restrict__ keep_idx, long G0, int ncols,
                               double* __restrict__ dst) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    if (i < G) dst[keep_idx[i] + (long)c * G0] = src[i + (long)c * G];
}

// Host count matrix (G0 x N column-major) -> every rank's device copy, when the ranks share the job: each rank uploads only
// its 1/world of the rows (N strided pieces) and the shards are all-gathered over NVLink, instead of every rank pulling
// the whole matrix through its own PCIe link (240 MB per rank at C5).
template <class T>
__global__ void k_unshard_counts(const T* __restrict__ recv, long G0, int N, long per, T* __restrict__ dst) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (g >= G0) return;
    const long rk = g / per;
    dst[g + (long)s * G0] = recv[((long)rk * N + s) * per + (g - rk * per)];
}
// a big PAGEABLE caller buffer goes through the pinned staging of host_bridge.hpp
static bool bridge_wanted(const void* host, size_t bytes) { return bytes >= ((size_t)16 << 20) && is_pageable(host); }

template <class T>
static const T* upload_counts_sharded(const T* h_counts, long G0, int N, cbs_comm* comm, cudaStream_t st) {
    const int world = comm->world, rank = comm->rank;
    const long per = (G0 + world - 1) / world;
    const long a = std::min(G0, (long)rank * per), b = std::min(G0, a + per);
    T* send = g_dev.get<T>("counts_send", (size_t)per * N);
    T* recv = g_dev.get<T>("counts_recv", (size_t)per * N * world);
    T* dst = g_dev.get<T>("counts_in", (size_t)G0 * N);
    if (b > a)
        if (bridge_wanted(h_counts, (size_t)(b - a) * N * sizeof(T)))
            bridge_run(t_ctx, {BridgeJob{(char*)(h_counts + a), (size_t)G0 * sizeof(T), (char*)send, (size_t)per * sizeof(T),
                                         (size_t)(b - a) * sizeof(T), (size_t)N, nullptr}}, true, comm->world);
        else
            CBS_CUDA(cudaMemcpy2DAsync(send, (size_t)per * sizeof(T), h_counts + a, (size_t)G0 * sizeof(T), (size_t)(b - a) * sizeof(T),
                                       (size_t)N, cudaMemcpyHostToDevice, st));
    CBS_NCCL(nccl_api()->AllGather(send, recv, (size_t)per * N * sizeof(T), NCCL_INT8, comm->nccl, st));
    CBS_LAUNCH(k_unshard_counts<T>, dim3((unsigned)((G0 + 255) / 256), (unsigned)N), 256, 0, st, (const T*)recv, G0, N, per, dst);
    return dst;
}

static const char* bad_counts_message(int bad) {
    if (bad & 1) return "NA counts not allowed";                   // edgeR DGEList's wording
    if (bad & 2) return "Negative counts not allowed";             // edgeR DGEList's wording
    if (bad & 4) return "non-integer counts are not supported: this engine works on int32 counts (round them first)";
    return "counts above 2^31 - 1 are not supported";
}

template <class T>
static void pipeline(const cbs_input* in, cbs_output* out, const T* h_or_d_counts) {
    const long G0 = in->G0;
    const int N = in->N, B = in->B, p_mod = in->p_mod;
    const int p = B + p_mod - 1;
    cbs_comm* comm = in->comm;
    const int rank = comm ? comm->rank : 0, world = comm ? comm->world : 1;
    double stage[CBS_N_STAGES] = {0};
    const double t_start = now_ms();
    double t0 = t_start;
    cudaStream_t st = t_ctx->st;
    auto lap = [&](int which) { CBS_CUDA(cudaStreamSynchronize(st)); const double t = now_ms(); stage[which] += t - t0; t0 = t; };

    // ---- H2D ----
    const T* d_in;
    if (in->counts_on_device) d_in = h_or_d_counts;
    else if (world > 1) d_in = upload_counts_sharded<T>(h_or_d_counts, G0, N, comm, st);
    else if (bridge_wanted(h_or_d_counts, (size_t)G0 * N * sizeof(T))) {      // pageable caller buffer (an R matrix): staged by worker threads
        T* d = g_dev.get<T>("counts_in", (size_t)G0 * N);
        bridge_run(t_ctx, {BridgeJob{(char*)h_or_d_counts, (size_t)G0 * sizeof(T), (char*)d, (size_t)G0 * sizeof(T), (size_t)G0 * sizeof(T),
                                     (size_t)N, nullptr}}, true);
        d_in = d;
    } else d_in = upload(g_dev, "counts_in", h_or_d_counts, (size_t)G0 * N, st);
    int* d_bidx = upload(g_dev, "batch_idx", in->batch_idx, (size_t)N, st);
    lap(CBS_ST_H2D);

    // ---- a1: zero filter (ComBat_seq.R:33-40) + input validation ----
    unsigned char* d_keep = g_dev.get<unsigned char>("keep", (size_t)G0);
    int* d_bad = g_dev.get<int>("bad_counts", 1);
    CBS_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
    unsigned long long* d_seen = g_dev.get<unsigned long long>("keep_seen", (size_t)2 * G0);
    CBS_CUDA(cudaMemsetAsync(d_seen, 0, (size_t)2 * G0 * sizeof(unsigned long long), st));
    CBS_LAUNCH(k_keep_mask<T>, dim3((unsigned)((G0 + 127) / 128), (unsigned)std::max(1, std::min(16, N / 32))), 128, 0, st, d_in, G0, N,
               d_bidx, d_seen, d_bad);
    CBS_LAUNCH(k_keep_finish, grid1(G0, 256), 256, 0, st, (const unsigned long long*)d_seen, G0, B, d_keep);
    std::vector<unsigned char> keep(G0);
    int bad = 0;
    CBS_CUDA(cudaMemcpyAsync(keep.data(), d_keep, (size_t)G0, cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
    if (bad) { fail(CBS_EINVAL, "%s", bad_counts_message(bad)); throw CbsError{CBS_EINVAL}; }
    std::vector<int> keep_idx;
    keep_idx.reserve(G0);
    for (long g = 0; g < G0; ++g) if (keep[g]) keep_idx.push_back((int)g);
    const long G = (long)keep_idx.size();
    if (out->keep) memcpy(out->keep, keep.data(), (size_t)G0);
    if (G == 0) { fail(CBS_EINVAL, "no genes left after the zero filter"); throw CbsError{CBS_EINVAL}; }
    if (in->shrink && in->gene_subset_n > 0 && in->gene_subset_n > G - 1) {     // sample(1:(nrow(dat)-1), n)  helper_seq.R:19
        fail(CBS_EINVAL, "cannot take a sample larger than the population when 'replace = FALSE'");
        throw CbsError{CBS_EINVAL};
    }
    int* d_keep_idx = upload(g_dev, "keep_idx", keep_idx.data(), (size_t)G, st);

    // ---- a1/a2: int32 compaction + library sizes (DGEList, ComBat_seq.R:43) ----
    int32_t* d_cnt = g_dev.get<int32_t>("cnt", (size_t)G * N);
    unsigned long long* d_cs = g_dev.get<unsigned long long>("colsum", (size_t)N);
    CBS_CUDA(cudaMemsetAsync(d_cs, 0, (size_t)N * sizeof(unsigned long long), st));
    CBS_LAUNCH(k_compact_colsum<T>, dim3((unsigned)((G + 255) / 256), (unsigned)N), 256, 0, st, d_in, G0, d_keep_idx, G, d_cnt, d_cs);
    std::vector<unsigned long long> cs(N);
    CBS_CUDA(cudaMemcpyAsync(cs.data(), d_cs, (size_t)N * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
    std::vector<double> lib(N), logoff(N);
    for (int s = 0; s < N; ++s) { lib[s] = (double)cs[s]; logoff[s] = std::log(lib[s]); }
    if (out->lib_size) memcpy(out->lib_size, lib.data(), (size_t)N * sizeof(double));
    std::vector<double> offpairs((size_t)2 * N);                  // (log(lib), lib) pairs: see glm.cuh RowSrc::logoff
    for (int s = 0; s < N; ++s) { offpairs[2 * s] = logoff[s]; offpairs[2 * s + 1] = lib[s]; }
    double* d_logoff = upload(g_dev, "logoff", offpairs.data(), (size_t)2 * N, st);

    long g0, g1;
    shard_range(G, rank, world, &g0, &g1);

    // ---- a3/a4: designs (ComBat_seq.R:53-81, fallback predicate :101,:111) ----
    std::vector<std::vector<int>> bcols(B);
    for (int s = 0; s < N; ++s) bcols[in->batch_idx[s]].push_back(s);
    std::vector<double> design((size_t)N * p, 0.0);
    for (int s = 0; s < N; ++s) design[(size_t)in->batch_idx[s] * N + s] = 1.0;
    for (int k = 1; k < p_mod; ++k) for (int s = 0; s < N; ++s) design[(size_t)(B + k - 1) * N + s] = in->mod[(size_t)k * N + s];
    DesignSet full("full");
    full.add(make_design(design.data(), N, p, nullptr));
    full.finish();
    DesignSet bds("batch");                    // per-batch dispersion designs
    std::vector<char> fallback(B, 0), do_tagwise(B, 0);
    for (int b = 0; b < B; ++b) {
        const int nb = (int)bcols[b].size();
        std::vector<double> modb((size_t)nb * p_mod);
        for (int t = 0; t < nb; ++t) for (int k = 0; k < p_mod; ++k) modb[(size_t)k * nb + t] = in->mod[(size_t)k * N + bcols[b][t]];
        if (in->use_fallback) fallback[b] = in->use_fallback[b] != 0;
        else fallback[b] = (nb <= p - B + 1) || (qr_rank(modb.data(), nb, p_mod) < p_mod);
        if (fallback[b]) {
            std::vector<double> ones(nb, 1.0);
            bds.add(make_design(ones.data(), nb, 1, bcols[b].data()));
        } else {
            bds.add(make_design(modb.data(), nb, p_mod, bcols[b].data()));
            do_tagwise[b] = 1;
        }
    }
    bds.finish();
    unsigned char* d_sel = g_dev.get<unsigned char>("sel", (size_t)B * G);
    CBS_LAUNCH(k_rowsum_sel, dim3((unsigned)((G + 127) / 128), (unsigned)B), 128, 0, st, d_cnt, G, N, d_bidx, B, d_sel);
    lap(CBS_ST_PREP);

    // ---- a5: common dispersion per batch (ComBat_seq.R:100-107) ----
    std::vector<double> common;
    run_common_disp(d_cnt, G, g0, g1, d_logoff, bds, d_sel, comm, st, common, nullptr);
    if (out->disp_common) memcpy(out->disp_common, common.data(), (size_t)B * sizeof(double));
    lap(CBS_ST_COMMON_DISP);

    // ---- a6: gene-wise dispersion per batch (ComBat_seq.R:110-119) ----
    double* d_phi_hat = g_dev.get<double>("phi_hat", (size_t)G * B);
    run_tagwise_disp(d_cnt, G, g0, g1, d_logoff, bds, d_sel, common, do_tagwise, st, d_phi_hat);
    lap(CBS_ST_TAGWISE_DISP);

    // ---- a8-a10: the two glmFit calls (ComBat_seq.R:130-138) ----
    std::vector<double> aw(B);
    for (int b = 0; b < B; ++b) aw[b] = (double)bcols[b].size() / N;             // n_batches / n_sample
    double* d_aw = upload(g_dev, "alpha_w", aw.data(), (size_t)B, st);
    double* d_alpha = g_dev.get<double>("alpha", (size_t)G);
    double* d_coef = g_dev.get<double>("coef", (size_t)G * p);
    double* d_mu_hat = g_dev.get<double>("mu_hat", (size_t)G * N);
    const int glm_flags = glm_class_flags(full.host[0], in->batch_idx);
    // glm_f: only its prior-count coefficients are used, through alpha_g (ComBat_seq.R:131)
    run_glm_fit(d_cnt, G, g0, g1, N, d_logoff, lib, d_phi_hat, d_bidx, full.d_array, p, nullptr, 1e-4, nullptr, nullptr, d_aw, B, d_alpha, st, glm_flags);
    // glm_f2: offset = log(lib) + alpha_g (ComBat_seq.R:132-134); gamma_hat = coef[,1:B], mu_hat = fitted.values
    run_glm_fit(d_cnt, G, g0, g1, N, d_logoff, lib, d_phi_hat, d_bidx, full.d_array, p, d_alpha, 1e-4, d_coef, d_mu_hat, nullptr, 0, nullptr, st, glm_flags);
    double* d_gamma_hat = d_coef;        // first B columns
    lap(CBS_ST_GLM);

    // ---- a11: Monte-Carlo EB (ComBat_seq.R:141-170) ----
    double* d_gamma_star = d_gamma_hat;
    double* d_phi_star_mat = d_phi_hat;
    std::vector<unsigned long long> nfb(B, 0);
    if (in->shrink) {
        allgather_rows(d_mu_hat, G, N, g0, g1, comm, st);
        allgather_rows(d_gamma_hat, G, B, g0, g1, comm, st);
        allgather_rows(d_phi_hat, G, B, g0, g1, comm, st);
        d_gamma_star = g_dev.get<double>("gamma_star", (size_t)G * B);
        double* d_ps = g_dev.get<double>("phi_star_mat", (size_t)G * B);
        run_mc_stage(d_cnt, G, B, bcols, g0, g1, d_mu_hat, d_gamma_hat, d_phi_hat, in->mc_mode,
                     in->gene_subset_n > 0 ? in->gene_subset_n : 0, in->seed, d_gamma_star, d_ps, nfb.data(), st, comm);
        if (in->shrink_disp) d_phi_star_mat = d_ps;                                // ComBat_seq.R:162-165
    }
    if (out->n_mc_fallback) for (int b = 0; b < B; ++b) out->n_mc_fallback[b] = (int64_t)nfb[b];
    lap(CBS_ST_MC);

    // ---- a12-a14: batch-free distribution + match_quantiles + re-insertion (ComBat_seq.R:173-203) ----
    double* d_phi_star = g_dev.get<double>("phi_star", (size_t)G);
    CBS_LAUNCH(k_phi_star_mean, grid1(G, 256), 256, 0, st, d_phi_star_mat, G, B, d_phi_star);
    double* d_adj;
    if (out->adjusted_on_device) d_adj = out->adjusted;
    else d_adj = out->adjusted ? g_dev.get<double>("adjusted", (size_t)G0 * N) : nullptr;
    // G0 rows written by this rank: rows of its kept genes plus the dropped genes in between
    const long r0 = (rank == 0 || g0 >= G) ? (rank == 0 ? 0 : G0) : keep_idx[g0];
    const long r1 = (g1 >= G) ? G0 : keep_idx[g1];
    out->row_begin = r0; out->row_end = r1;
    // match_quantiles runs in up to 8 launches over the samples (in batch order); when the result goes to a host buffer, the
    // columns of a finished launch are copied out on a second stream while the next launch computes (the D2H of the adjusted
    // matrix, 480 MB at C5, hides behind the kernel instead of following it).
    const bool to_host = out->adjusted && !out->adjusted_on_device;
    cudaStream_t st2 = to_host ? t_ctx->copy_stream() : nullptr;
    bool copied_by_columns = false;
    if (d_adj) {
        CBS_LAUNCH(k_counts_to_f64<T>, grid1(G0 * N, 256), 256, 0, st, d_in, G0 * N, d_adj);
        if (g1 > g0) {
            std::vector<int> order;                                   // samples sorted by batch (stable)
            order.reserve(N);
            for (int b = 0; b < B; ++b) order.insert(order.end(), bcols[b].begin(), bcols[b].end());
            int* d_order = upload(g_dev, "mq_order", order.data(), (size_t)N, st);
            const int nblk = (N + 31) / 32, nchunk = (to_host && r1 > r0) ? std::min(8, nblk) : 1;
            std::vector<int> edge(nchunk + 1);
            for (int c = 0; c <= nchunk; ++c) edge[c] = std::min(N, (int)((long)nblk * c / nchunk) * 32);
            for (int c = 0; c < nchunk; ++c) {                        // all launches are enqueued before the first copy: a copy into
                launch_match_quantiles(d_cnt, G, g0, g1, d_mu_hat, d_bidx, d_order, N, d_phi_hat, d_gamma_star, d_phi_star,   // pageable memory blocks the host
                                       d_keep_idx, G0, d_adj, st, edge[c], edge[c + 1]);
                if (nchunk > 1) CBS_CUDA(cudaEventRecord(t_ctx->event("mq_chunk" + std::to_string(c)), st));
            }
            if (nchunk > 1) {
                copied_by_columns = true;
                const bool bridge = bridge_wanted(out->adjusted, (size_t)(r1 - r0) * N * sizeof(double));
                std::vector<BridgeJob> jobs;
                for (int c = 0; c < nchunk; ++c) {
                    cudaEvent_t done = t_ctx->event("mq_chunk" + std::to_string(c));
                    if (!bridge) CBS_CUDA(cudaStreamWaitEvent(st2, done, 0));
                    for (int si = edge[c]; si < edge[c + 1];) {       // runs of consecutive columns go out as one 2-D copy
                        int sj = si + 1;
                        while (sj < edge[c + 1] && order[sj] == order[sj - 1] + 1) ++sj;
                        const size_t off = (size_t)r0 + (size_t)order[si] * G0;
                        if (bridge)
                            jobs.push_back({(char*)(out->adjusted + off), (size_t)G0 * sizeof(double), (char*)(d_adj + off), (size_t)G0 * sizeof(double),
                                            (size_t)(r1 - r0) * sizeof(double), (size_t)(sj - si), done});
                        else
                            CBS_CUDA(cudaMemcpy2DAsync(out->adjusted + off, (size_t)G0 * sizeof(double), d_adj + off, (size_t)G0 * sizeof(double),
                                                       (size_t)(r1 - r0) * sizeof(double), (size_t)(sj - si), cudaMemcpyDeviceToHost, st2));
                        si = sj;
                    }
                }
                if (bridge) bridge_run(t_ctx, jobs, false, world);           // pageable caller buffer: staged by worker threads
            }
        }
    }
    lap(CBS_ST_MATCH_QUANTILES);

    // ---- D2H ----
    auto fetch_rows = [&](double* host, const double* dev_G0, int ncols) {   // rows [r0,r1) of a G0 x ncols matrix
        if (!host || r1 <= r0) return;
        CBS_CUDA(cudaMemcpy2DAsync(host + r0, (size_t)G0 * sizeof(double), dev_G0 + r0, (size_t)G0 * sizeof(double),
                                   (size_t)(r1 - r0) * sizeof(double), (size_t)ncols, cudaMemcpyDeviceToHost, st));
    };
    if (to_host && !copied_by_columns) {
        if (r0 == 0 && r1 == G0) CBS_CUDA(cudaMemcpyAsync(out->adjusted, d_adj, (size_t)G0 * N * sizeof(double), cudaMemcpyDeviceToHost, st));
        else fetch_rows(out->adjusted, d_adj, N);
    }
    auto fetch_scattered = [&](double* host, const double* dev_G, int ncols, const char* name) {
        if (!host) return;
        double* tmp = g_dev.get<double>(name, (size_t)G0 * ncols);
        const double qnan = std::nan("");
        CBS_LAUNCH(k_fill, grid1(G0 * ncols, 256), 256, 0, st, tmp, G0 * ncols, qnan);
        CBS_LAUNCH(k_scatter_rows, dim3((unsigned)((G + 255) / 256), (unsigned)ncols), 256, 0, st, dev_G, G, d_keep_idx, G0, ncols, tmp);
        if (world == 1) CBS_CUDA(cudaMemcpyAsync(host, tmp, (size_t)G0 * ncols * sizeof(double), cudaMemcpyDeviceToHost, st));
        else fetch_rows(host, tmp, ncols);
    };
    fetch_scattered(out->phi_hat, d_phi_hat, B, "out_phi_hat");
    fetch_scattered(out->gamma_hat, d_gamma_hat, B, "out_gamma_hat");
    fetch_scattered(out->gamma_star, d_gamma_star, B, "out_gamma_star");
    fetch_scattered(out->phi_star, d_phi_star_mat, B, "out_phi_star");
    fetch_scattered(out->mu_hat, d_mu_hat, N, "out_mu_hat");
    if (copied_by_columns) CBS_CUDA(cudaStreamSynchronize(st2));
    lap(CBS_ST_D2H);
    stage[CBS_ST_TOTAL] = now_ms() - t_start;
    if (out->stage_ms) memcpy(out->stage_ms, stage, sizeof stage);
}

// ---- single-process multi-GPU (cbs_input.n_gpus): one host thread per device, an NCCL communicator kept per process ----
struct InprocComms { int n = 0; std::vector<cbs_comm> ranks; };
static InprocComms g_inproc;

void release_inproc_comms() {
    NcclApi* a = g_inproc.n ? nccl_api() : nullptr;
    for (auto& c : g_inproc.ranks) if (a && c.nccl) { cudaSetDevice(c.device); a->CommDestroy(c.nccl); }
    g_inproc.ranks.clear();
    g_inproc.n = 0;
}

static void ensure_inproc_comms(int n) {
    if (g_inproc.n == n) return;
    release_inproc_comms();
    NcclApi* a = nccl_api();
    if (!a) throw CbsError{CBS_ENCCL};
    Id128 id;
    CBS_NCCL(a->GetUniqueId(&id));
    g_inproc.ranks.assign(n, cbs_comm{nullptr, 0, n, 0});
    std::vector<int> rc(n, 0);
    std::vector<std::thread> th;
    for (int r = 0; r < n; ++r)
        th.emplace_back([&, r] {
            if (cudaSetDevice(r) != cudaSuccess) { rc[r] = -1; return; }
            void* c = nullptr;
            rc[r] = a->CommInitRank(&c, n, id, r);
            g_inproc.ranks[r] = cbs_comm{c, r, n, r};
        });
    for (auto& t : th) t.join();
    for (int r = 0; r < n; ++r)
        if (rc[r] != 0) {
            fail(CBS_ENCCL, "ncclCommInitRank failed on device %d (%d)", r, rc[r]);
            g_inproc.n = n; release_inproc_comms();
            throw CbsError{CBS_ENCCL};
        }
    g_inproc.n = n;
}

template <class T>
static int pipeline_inproc(const cbs_input* in, cbs_output* out, const T* counts, int n) {
    ensure_inproc_comms(n);
    for (int r = 0; r < n; ++r) ctx_for_device(r);              // contexts are created under the library mutex
    std::vector<int> code(n, CBS_OK);
    std::vector<std::string> msg(n);
    std::vector<cbs_output> outs(n, *out);
    std::vector<std::thread> th;
    for (int r = 0; r < n; ++r)
        th.emplace_back([&, r] {
            try {
                enter_device(r);
                cbs_input ir = *in;
                ir.comm = &g_inproc.ranks[r];
                ir.device = r;
                if (r != 0) {                                   // per-call vectors come from rank 0; matrices are written by row range
                    outs[r].keep = nullptr; outs[r].lib_size = nullptr; outs[r].disp_common = nullptr;
                    outs[r].stage_ms = nullptr; outs[r].n_mc_fallback = nullptr;
                }
                pipeline<T>(&ir, &outs[r], counts);
                CBS_CUDA(cudaStreamSynchronize(t_ctx->st));
            } catch (const CbsError& e) { code[r] = e.code; msg[r] = g_err; cudaDeviceSynchronize(); cudaGetLastError(); }
              catch (const std::exception& e) { code[r] = CBS_EINVAL; msg[r] = e.what(); }
        });
    for (auto& t : th) t.join();
    for (int r = 0; r < n; ++r) if (code[r] != CBS_OK) return fail(code[r], "%s", msg[r].c_str());
    out->row_begin = 0; out->row_end = in->G0;
    return CBS_OK;
}

}  // namespace cbs

extern "C" int cbs_shard_range(int64_t G, int32_t rank, int32_t world, int64_t* g0, int64_t* g1) {
    if (G < 0 || world < 1 || rank < 0 || rank >= world || !g0 || !g1) return fail(CBS_EINVAL, "bad shard arguments");
    long a, b;
    shard_range((long)G, rank, world, &a, &b);
    *g0 = a; *g1 = b;
    return CBS_OK;
}

extern "C" int cbs_combat_seq(const cbs_input* in, cbs_output* out) {
    // argument validation comes first (it needs no device): the reference's stop() conditions keep their wording
    if (!in || !out) return fail(CBS_EINVAL, "null input/output struct");
    if (in->G0 <= 0 || in->N <= 0 || in->B <= 0 || in->p_mod < 1) return fail(CBS_EINVAL, "bad dimensions");
    if ((in->counts_f64 != nullptr) == (in->counts_i32 != nullptr)) return fail(CBS_EINVAL, "exactly one of counts_f64 / counts_i32 must be set");
    if (in->counts_on_device && !in->counts_i32) return fail(CBS_EINVAL, "counts_on_device needs counts_i32");
    if (!in->batch_idx || !in->mod) return fail(CBS_EINVAL, "batch_idx and mod are required");
    if (in->B < 2) return fail(CBS_EINVAL, "ComBat-seq needs at least 2 batches");
    if (in->B > 128) return fail(CBS_EINVAL, "at most 128 batches are supported");
    if (in->n_gpus < 0) return fail(CBS_EINVAL, "n_gpus must be >= 0");
    if (in->n_gpus != 1 && (in->comm || in->counts_on_device || out->adjusted_on_device))
        return fail(CBS_EINVAL, "n_gpus != 1 (single-process multi-GPU) needs comm == NULL and host buffers");
    const int p = in->B + in->p_mod - 1;
    if (p > PMAX) return fail(CBS_EINVAL, "design has %d columns; this build supports up to %d", p, PMAX);
    std::vector<int> nb(in->B, 0);
    for (int s = 0; s < in->N; ++s) {
        if (in->batch_idx[s] < 0 || in->batch_idx[s] >= in->B) return fail(CBS_EINVAL, "batch_idx out of range");
        nb[in->batch_idx[s]]++;
    }
    for (int b = 0; b < in->B; ++b)
        if (nb[b] <= 1) return fail(CBS_ESINGLE_SAMPLE_BATCH, "ComBat-seq doesn't support 1 sample per batch yet");   // ComBat_seq.R:29-31
    for (int s = 0; s < in->N; ++s) if (in->mod[s] != 1.0) return fail(CBS_EINVAL, "column 0 of mod must be the intercept");
    // confounding checks (ComBat_seq.R:84-90)
    {
        const int N = in->N, B = in->B;
        std::vector<double> design((size_t)N * p, 0.0);
        for (int s = 0; s < N; ++s) design[(size_t)in->batch_idx[s] * N + s] = 1.0;
        for (int k = 1; k < in->p_mod; ++k) for (int s = 0; s < N; ++s) design[(size_t)(B + k - 1) * N + s] = in->mod[(size_t)k * N + s];
        if (qr_rank(design.data(), N, p) < p) {
            if (p == B + 1) return fail(CBS_ECONF_COVAR_BATCH, "The covariate is confounded with batch! Remove the covariate and rerun ComBat-Seq");
            if (p > B + 1) {
                if (qr_rank(design.data() + (size_t)B * N, N, p - B) < p - B)
                    return fail(CBS_ECONF_COVARS, "The covariates are confounded! Please remove one or more of the covariates so the design is not confounded");
                return fail(CBS_ECONF_COVAR_WITH_BATCH, "At least one covariate is confounded with batch! Please remove confounded covariates and rerun ComBat-Seq");
            }
        }
    }
    CBS_NEED_GPU();
    std::lock_guard<std::mutex> lock_(g_mutex);
    try {
        int n = in->n_gpus;
        if (n == 0) n = cbs_device_count();
        if (n > cbs_device_count()) return fail(CBS_EINVAL, "n_gpus = %d but only %d CUDA devices are visible", n, cbs_device_count());
        if (n > 1) {
            int cur = 0;
            cudaGetDevice(&cur);
            const int rc = in->counts_f64 ? pipeline_inproc<double>(in, out, in->counts_f64, n) : pipeline_inproc<int32_t>(in, out, in->counts_i32, n);
            cudaSetDevice(cur);
            return rc;
        }
        enter_device(in->comm ? in->comm->device : in->device);
        if (in->counts_f64) pipeline<double>(in, out, in->counts_f64);
        else pipeline<int32_t>(in, out, in->counts_i32);
    CBS_ENTRY_END
}
