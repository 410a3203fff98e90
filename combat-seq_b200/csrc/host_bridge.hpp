// host_bridge.hpp — copies between PAGEABLE host buffers and the device, staged through pinned memory by worker threads.
//
// An R numeric matrix (the caller the drop-in is for, /root/reference/ComBat_seq.R:22) lives in pageable memory.  A plain
// cudaMemcpy from or to such a buffer is staged by the driver on the calling thread at 10-13 GB/s; for C5 that is 480 MB in
// and 480 MB out, a quarter of the whole call.  Here up to BRIDGE_WORKERS threads (half the host's cores at most) each copy their share of the rows through two
// pinned staging buffers and their own stream, so the host memcpy of one piece overlaps the DMA of another and several cores
// share the memcpy.  Pinned (or registered) caller buffers do not come here: they are copied directly and asynchronously.
#pragma once
#include "common.cuh"
#include <cstring>
#include <thread>
#include <vector>

namespace cbs {

#ifndef CBS_BRIDGE_WORKERS
#define CBS_BRIDGE_WORKERS 8      // measured on a 16-core B200 host, C5 from pageable buffers: 2 -> 201.6, 4 -> 194.5, 8 -> 181.0, 16 -> 181.0 ms per call
#endif
constexpr int BRIDGE_WORKERS = CBS_BRIDGE_WORKERS;
constexpr size_t BRIDGE_PIECE = 8u << 20;      // bytes per staging buffer

inline bool is_pageable(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

// One 2-D copy job: `height` lines of `width` bytes; host lines `hpitch` apart, device lines `dpitch` apart.
// wait (D2H only, may be null): event after which the device lines are final.
struct BridgeJob {
    char* host; size_t hpitch;
    char* dev; size_t dpitch;
    size_t width, height;
    cudaEvent_t wait;
};

// Runs the jobs (all host-to-device or all device-to-host) and returns when every byte has arrived.  Must be called by the
// thread that owns the device context `ctx` (its pinned cache provides the staging buffers).
inline void bridge_run(DevCtx* ctx, const std::vector<BridgeJob>& jobs, bool to_device, int share = 1) {
    struct Unit { const BridgeJob* j; size_t line0, lines; };
    std::vector<BridgeJob> cut;                                   // lines longer than a staging buffer are cut into column strips
    for (const BridgeJob& j : jobs)
        for (size_t off = 0; off < j.width && j.height; off += BRIDGE_PIECE)
            cut.push_back({j.host + off, j.hpitch, j.dev + off, j.dpitch, std::min(BRIDGE_PIECE, j.width - off), j.height, j.wait});
    std::vector<Unit> units;
    for (const BridgeJob& j : cut) {
        const size_t per = std::max<size_t>(1, BRIDGE_PIECE / j.width);
        for (size_t l = 0; l < j.height; l += per) units.push_back({&j, l, std::min(per, j.height - l)});
    }
    if (units.empty()) return;
    const unsigned hw = std::thread::hardware_concurrency();
    // `share` = devices (ranks or threads of one call) staging through this host at the same time: they divide the cores
    const unsigned budget = (hw ? hw / 2 : 2u) / (unsigned)std::max(1, share);
    const int nw = (int)std::min<size_t>(std::max(1u, std::min((unsigned)BRIDGE_WORKERS, budget)), units.size());
    std::vector<char*> stage(2 * nw);
    for (int i = 0; i < 2 * nw; ++i) stage[i] = (char*)ctx->pin.get("bridge" + std::to_string(i), BRIDGE_PIECE);
    const int device = ctx->device;
    std::vector<cudaError_t> err(nw, cudaSuccess);
    auto worker = [&](int w) {
        cudaError_t e = cudaSetDevice(device);
        cudaStream_t s = nullptr;
        cudaEvent_t ev[2] = {nullptr, nullptr};
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
        for (int k = 0; k < 2 && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
        auto lines_to_stage = [&](const Unit& u, char* buf) {
            const BridgeJob& j = *u.j;
            if (j.hpitch == j.width) memcpy(buf, j.host + u.line0 * j.hpitch, u.lines * j.width);
            else for (size_t l = 0; l < u.lines; ++l) memcpy(buf + l * j.width, j.host + (u.line0 + l) * j.hpitch, j.width);
        };
        auto stage_to_lines = [&](const Unit& u, const char* buf) {
            const BridgeJob& j = *u.j;
            if (j.hpitch == j.width) memcpy(j.host + u.line0 * j.hpitch, buf, u.lines * j.width);
            else for (size_t l = 0; l < u.lines; ++l) memcpy(j.host + (u.line0 + l) * j.hpitch, buf + l * j.width, j.width);
        };
        int k = 0;
        long prev = -1;                                           // D2H: unit whose bytes are on their way into stage[k ^ 1]
        for (size_t i = w; i < units.size() && e == cudaSuccess; i += nw, k ^= 1) {
            const Unit& u = units[i];
            const BridgeJob& j = *u.j;
            char* buf = stage[2 * w + k];
            if (to_device) {
                e = cudaEventSynchronize(ev[k]);                  // the copy that last read this buffer is done
                if (e != cudaSuccess) break;
                lines_to_stage(u, buf);
                e = cudaMemcpy2DAsync(j.dev + u.line0 * j.dpitch, j.dpitch, buf, j.width, j.width, u.lines, cudaMemcpyHostToDevice, s);
                if (e == cudaSuccess) e = cudaEventRecord(ev[k], s);
            } else {
                if (j.wait) e = cudaStreamWaitEvent(s, j.wait, 0);
                if (e == cudaSuccess)
                    e = cudaMemcpy2DAsync(buf, j.width, j.dev + u.line0 * j.dpitch, j.dpitch, j.width, u.lines, cudaMemcpyDeviceToHost, s);
                if (e == cudaSuccess) e = cudaEventRecord(ev[k], s);
                if (prev >= 0 && e == cudaSuccess) {              // meanwhile: the previous unit, from the other buffer, to the caller
                    e = cudaEventSynchronize(ev[k ^ 1]);
                    if (e == cudaSuccess) stage_to_lines(units[prev], stage[2 * w + (k ^ 1)]);
                }
                prev = (long)i;
            }
        }
        if (e == cudaSuccess && s) e = cudaStreamSynchronize(s);
        if (!to_device && prev >= 0 && e == cudaSuccess) stage_to_lines(units[prev], stage[2 * w + (k ^ 1)]);
        for (int q = 0; q < 2; ++q) if (ev[q]) cudaEventDestroy(ev[q]);
        if (s) cudaStreamDestroy(s);
        err[w] = e;
    };
    std::vector<std::thread> th;
    for (int w = 1; w < nw; ++w) th.emplace_back(worker, w);
    worker(0);
    for (auto& t : th) t.join();
    for (int w = 0; w < nw; ++w) CBS_CUDA(err[w]);
}

}  // namespace cbs
