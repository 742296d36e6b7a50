// orbx_api.cu — C ABI of liborbx.so: handle life cycle and the extractor entry points
// (declared in include/orbx.h; each cites the reference function it stands in for).

#include "orbx_internal.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <thread>

static thread_local char g_err[512] = "";

void orbx_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static inline int cv_roundf(float v) { return (int)lrintf(v); }
static inline int cv_round(double v) { return (int)lrint(v); }

// ORBextractor::ORBextractor tables (reference src/ORBextractor.cc:410-470)
static void compute_tables(orbx_handle* h)
{
    const int nl = h->params.nlevels;
    const double scaleFactor = (double)h->params.scale_factor;   // member is a double holding a float
    h->scale.assign(nl, 1.f); h->sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        h->scale[i] = (float)(h->scale[i - 1] * scaleFactor);
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    h->invScale.resize(nl); h->invSigma2.resize(nl);
    for (int i = 0; i < nl; ++i) { h->invScale[i] = 1.0f / h->scale[i]; h->invSigma2[i] = 1.0f / h->sigma2[i]; }
    h->perLevel.assign(nl, 0);
    const float factor = (float)(1.0f / scaleFactor);
    float nDesired = h->params.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        h->perLevel[l] = cv_roundf(nDesired);
        sum += h->perLevel[l];
        nDesired *= factor;
    }
    h->perLevel[nl - 1] = std::max(h->params.nfeatures - sum, 0);
    // umax: end of each row of the circular patch (src/ORBextractor.cc:452-469)
    const int HALF = 15;
    int v, v0;
    const int vmax = (int)floorf(HALF * sqrtf(2.f) / 2 + 1);
    const int vmin = (int)ceilf(HALF * sqrtf(2.f) / 2);
    const double hp2 = HALF * HALF;
    for (v = 0; v < 16; ++v) h->umax[v] = 0;
    for (v = 0; v <= vmax; ++v) h->umax[v] = cv_round(sqrt(hp2 - v * v));
    for (v = HALF, v0 = 0; v >= vmin; --v) {
        while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
        h->umax[v] = v0;
        ++v0;
    }
}

namespace {
struct ScratchBlock { void* p; size_t size; int dev; bool used; };
struct ScratchPool {
    std::vector<ScratchBlock> blocks;
    void* host = nullptr;
    size_t hostBytes = 0;
    ~ScratchPool() { for (ScratchBlock& b : blocks) cudaFree(b.p); if (host) cudaFreeHost(host); }   // thread exit; errors at process teardown are ignored
};
thread_local ScratchPool g_scratch;
}

cudaStream_t orbx_call_stream()
{
    thread_local cudaStream_t st[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    if (!st[dev] && cudaStreamCreateWithFlags(&st[dev], cudaStreamNonBlocking) != cudaSuccess) { st[dev] = 0; cudaGetLastError(); }
    return st[dev];                              // 0 (the legacy stream) only if no stream could be made
}

int orbx_scratch_alloc(void** p, size_t bytes)
{
    static const bool pooled = !getenv("ORBX_NO_SCRATCH_POOL");
    if (!pooled) { ORBX_CUDA(cudaMalloc(p, bytes ? bytes : 4)); return ORBX_OK; }
    size_t sz = 512;
    while (sz < bytes) sz <<= 1;
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    for (ScratchBlock& b : g_scratch.blocks)
        if (!b.used && b.dev == dev && b.size == sz) { b.used = true; *p = b.p; return ORBX_OK; }
    ORBX_CUDA(cudaMalloc(p, sz));
    g_scratch.blocks.push_back({*p, sz, dev, true});
    return ORBX_OK;
}

int orbx_scratch_host(void** p, size_t bytes)
{
    if (g_scratch.hostBytes < bytes) {
        // the buffer belongs to this thread's calls, which all end with a synchronisation of its stream: nothing else can be using it
        if (g_scratch.host) { cudaStreamSynchronize(orbx_call_stream()); cudaFreeHost(g_scratch.host); g_scratch.host = nullptr; g_scratch.hostBytes = 0; }
        size_t sz = 1 << 16;
        while (sz < bytes) sz <<= 1;
        ORBX_CUDA(cudaMallocHost(&g_scratch.host, sz));
        g_scratch.hostBytes = sz;
    }
    *p = g_scratch.host;
    return ORBX_OK;
}

void orbx_scratch_free(void* p)
{
    if (!p) return;
    for (ScratchBlock& b : g_scratch.blocks)
        if (b.p == p) { b.used = false; return; }
    cudaFree(p);                              // not from this thread's pool (pool disabled, or released on another thread)
}

void orbx_drop_graph(orbx_handle* h)
{
    if (h->graph1Valid) cudaGraphExecDestroy(h->graph1);
    h->graph1 = nullptr;
    h->graph1Valid = false;
}

void orbx_prof_mark(orbx_handle* h, int stage)
{
    if (!h->profiling) return;
    if (h->evUsed == h->evPool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        h->evPool.push_back(e);
        h->evStage.push_back(-1);
    }
    cudaEventRecord(h->evPool[h->evUsed], h->stream);
    h->evStage[h->evUsed] = stage;
    ++h->evUsed;
}

// Frames from the caller (host or device) into the pitched workspace.  A 2-D host->device copy
// with different source and destination pitches runs at a fraction of the link rate (one DMA
// descriptor per 1-2 KB row), so contiguous host frames are copied LINEARLY into a device landing
// zone and re-pitched there by a device-to-device 2-D copy.
int orbx_copy_frames(orbx_handle* h, uint8_t* dst, long long dstFrameBytes, uint8_t* stage, size_t stageBytes, const uint8_t* src, size_t frame_stride,
                     size_t row_stride, int width, int height, int nf, bool srcOnDevice, cudaStream_t st, long long* frameStrideOut)
{
    const cudaMemcpyKind kind = srcOnDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (frame_stride == row_stride * (size_t)height) {
        // row-contiguous frames: one tall image of nf*height rows
        if (!srcOnDevice && (int)row_stride != h->l0Pitch && (size_t)nf * frame_stride <= stageBytes) {     // stageBytes: what the caller reserved at `stage`
            ORBX_CUDA(cudaMemcpyAsync(stage, src, (size_t)nf * frame_stride, cudaMemcpyHostToDevice, st));
            ORBX_CUDA(cudaMemcpy2DAsync(dst, h->l0Pitch, stage, row_stride, width, (size_t)height * nf, cudaMemcpyDeviceToDevice, st));
        } else
            ORBX_CUDA(cudaMemcpy2DAsync(dst, h->l0Pitch, src, row_stride, width, (size_t)height * nf, kind, st));
        *frameStrideOut = (long long)h->l0Pitch * height;
    } else {
        for (int f = 0; f < nf; ++f)
            ORBX_CUDA(cudaMemcpy2DAsync(dst + (size_t)f * dstFrameBytes, h->l0Pitch, src + (size_t)f * frame_stride, row_stride, width, height, kind, st));
        *frameStrideOut = dstFrameBytes;
    }
    return ORBX_OK;
}

extern "C" {

const char* orbx_last_error(void) { return g_err; }

int orbx_set_profiling(orbx_handle* h, int enable)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    h->profiling = enable != 0;
    return ORBX_OK;
}

int orbx_get_profile(orbx_handle* h, double* stage_ms, int64_t* stage_launches)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    for (size_t i = 1; i < h->evUsed; ++i) {
        const int st = h->evStage[i];
        if (st < 0) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, h->evPool[i - 1], h->evPool[i]) == cudaSuccess) h->stageMs[st] += ms;
    }
    h->evUsed = 0;
    for (int s = 0; s < ORBX_NSTAGES; ++s) {
        if (stage_ms) stage_ms[s] = h->stageMs[s];
        if (stage_launches) stage_launches[s] = h->stageLaunches[s];
        h->stageMs[s] = 0.0;
        h->stageLaunches[s] = 0;
    }
    return ORBX_OK;
}

int orbx_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { orbx_set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return ORBX_E_CUDA; }
    return n;
}

static int create_impl(orbx_handle* h, const orbx_params* params, int result_slots, void* stream);

int orbx_create(const orbx_params* params, int max_width, int max_height, int chunk_frames,
                int result_slots, int device, void* stream, orbx_handle** out)
{
    if (!params || !out) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (params->nlevels < 1 || params->nlevels > ORBX_MAX_LEVELS || params->nfeatures < 1 ||
        params->scale_factor <= 1.0f || params->min_th_fast < 1 || params->min_th_fast > 254 ||
        params->ini_th_fast < 1 || params->ini_th_fast > 254) {
        orbx_set_error("unsupported ORBextractor parameters (nlevels 1..%d, scale>1, 1<=minTh,iniTh<255)", ORBX_MAX_LEVELS);
        return ORBX_E_INVALID;
    }
    if (max_width < 64 || max_height < 64 || max_width > 4000 || max_height > 4000 || chunk_frames < 1 || result_slots < 1) {
        orbx_set_error("bad handle dimensions");
        return ORBX_E_INVALID;
    }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (ndev <= 0 || device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    orbx_handle* h = new orbx_handle();
    h->params = *params;
    h->device = device;
    h->maxW = max_width; h->maxH = max_height;
    h->chunkFrames = chunk_frames; h->resultSlots = result_slots;
    h->curW = h->curH = 0;
    h->dCells = nullptr; h->dCoef = nullptr; h->dBlurTiles = nullptr; h->dKpMap = nullptr; h->nBlurTiles = 0;
    h->dLevel0 = h->dStageLin = h->dRaw = h->dBlur = nullptr;
    h->dCand = h->dKeysA = h->dKeysB = nullptr; h->dCellSeg = nullptr;
    h->dLevelKps = nullptr; h->dLevelKpCount = nullptr; h->dStatus = nullptr;
    h->dKpsUn = nullptr; h->storeUndistorted = false; h->dFrontStage = nullptr; h->frontStageBytes = 0; h->boundsValid = false; h->bounds[0] = h->bounds[1] = h->bounds[2] = h->bounds[3] = 0.f;
    h->dKps = nullptr; h->dDesc = nullptr; h->dCount = nullptr; h->dSlotLevelCount = nullptr;
    h->dGridStart = nullptr; h->dGridIdx = nullptr; h->dMatches12 = nullptr; h->dNMatches = nullptr;
    h->dNCompares = nullptr; h->dURight = nullptr; h->dDepth = nullptr; h->dMatchScratch = nullptr;
    h->matchScratchPerPair = 0;
    h->hStage = nullptr; h->hStageBytes = 0; h->hPack = nullptr; h->hPackBytes = 0; h->evPack = nullptr;
    h->graph1 = nullptr; h->graph1Valid = false;
    h->lastChunkFrames = 0;
    h->profiling = false; h->evUsed = 0;
    for (int s = 0; s < ORBX_NSTAGES; ++s) { h->stageMs[s] = 0.0; h->stageLaunches[s] = 0; }
    h->stream = h->copyStream = h->d2hStream = nullptr; h->ownStream = false; h->evDone = nullptr;
    for (int b = 0; b < ORBX_L0_BUFS; ++b) { h->evCopied[b] = h->evFree[b] = nullptr; h->evFreeValid[b] = false; }
    for (int b = 0; b < ORBX_L0_BUFS; ++b) { h->hInPool[b] = nullptr; h->evInPool[b] = nullptr; h->evInPoolValid[b] = false; }
    for (int b = 0; b < 2; ++b) { h->hOutPool[b] = nullptr; h->evOutPool[b] = nullptr; }
    h->hInPoolBytes = h->hOutPoolBytes = 0;
    h->kpCap = 0;
    memset(&h->geom, 0, sizeof(h->geom));
    const int rc = create_impl(h, params, result_slots, stream);
    if (rc != ORBX_OK) { orbx_destroy(h); return rc; }       // nothing of a half-built handle survives
    *out = h;
    return ORBX_OK;
}

static int create_impl(orbx_handle* h, const orbx_params* params, int result_slots, void* stream)
{
    const int max_width = h->maxW, max_height = h->maxH;
    if (stream) { h->stream = (cudaStream_t)stream; h->ownStream = false; }
    else { ORBX_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)); h->ownStream = true; }
    ORBX_CUDA(cudaStreamCreateWithFlags(&h->copyStream, cudaStreamNonBlocking));
    ORBX_CUDA(cudaStreamCreateWithFlags(&h->d2hStream, cudaStreamNonBlocking));
    ORBX_CUDA(cudaEventCreateWithFlags(&h->evDone, cudaEventDisableTiming));
    for (int b = 0; b < ORBX_L0_BUFS; ++b) {
        ORBX_CUDA(cudaEventCreateWithFlags(&h->evCopied[b], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->evFree[b], cudaEventDisableTiming));
        h->evFreeValid[b] = false;
    }
    compute_tables(h);
    // per-frame keypoint capacity for the largest shape: sum_l max(N_l+3, 4*nIni)+1, nIni from the
    // aspect ratio of the maximum image (per-level ratios differ by rounding; +8 slack per level)
    {
        int cap = 0;
        for (int l = 0; l < params->nlevels; ++l) {
            const int w = cv_roundf((float)max_width * h->invScale[l]) - 32, hh = cv_roundf((float)max_height * h->invScale[l]) - 32;
            int nIni = (w > 0 && hh > 0) ? (int)roundf((float)w / hh) : 1;
            if (nIni < 1) nIni = 1;
            cap += std::max(h->perLevel[l] + 3, 4 * (nIni + 1)) + 1;
        }
        h->kpCap = (cap + 7) & ~7;
    }
    const size_t R = (size_t)result_slots;
    ORBX_CUDA(cudaMalloc((void**)&h->dKps, R * h->kpCap * sizeof(orbx_keypoint)));
    ORBX_CUDA(cudaMalloc((void**)&h->dDesc, R * h->kpCap * 32));
    ORBX_CUDA(cudaMalloc((void**)&h->dCount, R * sizeof(int)));
    ORBX_CUDA(cudaMalloc((void**)&h->dSlotLevelCount, R * ORBX_MAX_LEVELS * sizeof(int)));
    ORBX_CUDA(cudaMalloc((void**)&h->dStatus, sizeof(int)));
    ORBX_CUDA(cudaMemset(h->dCount, 0, R * sizeof(int)));
    ORBX_CUDA(cudaMemset(h->dStatus, 0, sizeof(int)));
    return ORBX_OK;
}

} // extern "C"

// Everything sized by the per-frame keypoint capacity is released and the two extractor outputs re-allocated;
// the matcher / frame-front buffers come back lazily (orbx_ensure_match_buffers, orbx_frames_rgbd).
int orbx_grow_result_store(orbx_handle* h, int newCap)
{
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    void** ptrs[] = {(void**)&h->dKps, (void**)&h->dDesc, (void**)&h->dKpsUn, (void**)&h->dGridStart, (void**)&h->dGridIdx, (void**)&h->dMatches12,
                     (void**)&h->dNMatches, (void**)&h->dNCompares, (void**)&h->dURight, (void**)&h->dDepth, (void**)&h->dMatchScratch};
    for (void** p : ptrs) { if (*p) cudaFree(*p); *p = nullptr; }
    h->matchScratchPerPair = 0;
    h->kpCap = 0;
    const size_t R = (size_t)h->resultSlots;
    ORBX_CUDA(cudaMalloc((void**)&h->dKps, R * newCap * sizeof(orbx_keypoint)));
    ORBX_CUDA(cudaMalloc((void**)&h->dDesc, R * newCap * 32));
    ORBX_CUDA(cudaMemsetAsync(h->dCount, 0, R * sizeof(int), h->stream));
    h->kpCap = newCap;
    return ORBX_OK;
}

extern "C" {

int orbx_reserve_shape(orbx_handle* h, int width, int height)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (width < 64 || height < 64) { orbx_set_error("image smaller than 64x64 is outside the extractor's domain"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    const int rc = orbx_setup_shape(h, width, height);
    return rc ? rc : h->kpCap;
}

int orbx_destroy(orbx_handle* h)
{
    if (!h) return ORBX_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    void* ptrs[] = {h->dCells, h->dCoef, h->dBlurTiles, h->dKpMap, h->dKpsUn, h->dFrontStage, h->dLevel0, h->dStageLin, h->dRaw, h->dBlur, h->dCand, h->dKeysA,
                    h->dKeysB, h->dCellSeg, h->dLevelKps, h->dLevelKpCount, h->dStatus, h->dKps,
                    h->dDesc, h->dCount, h->dSlotLevelCount, h->dGridStart, h->dGridIdx, h->dMatches12, h->dNMatches,
                    h->dNCompares, h->dURight, h->dDepth, h->dMatchScratch};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (h->hStage) cudaFreeHost(h->hStage);
    if (h->hPack) cudaFreeHost(h->hPack);
    for (int b = 0; b < ORBX_L0_BUFS; ++b) { if (h->hInPool[b]) cudaFreeHost(h->hInPool[b]); if (h->evInPool[b]) cudaEventDestroy(h->evInPool[b]); }
    for (int b = 0; b < 2; ++b) { if (h->hOutPool[b]) cudaFreeHost(h->hOutPool[b]); if (h->evOutPool[b]) cudaEventDestroy(h->evOutPool[b]); }
    if (h->evPack) cudaEventDestroy(h->evPack);
    orbx_drop_graph(h);
    for (cudaStream_t x : h->sideStreams) cudaStreamDestroy(x);
    for (cudaEvent_t e : h->sideEvents) cudaEventDestroy(e);
    for (cudaEvent_t e : h->levelEvents) cudaEventDestroy(e);
    for (cudaEvent_t e : h->evPool) cudaEventDestroy(e);
    for (int b = 0; b < ORBX_L0_BUFS; ++b) { if (h->evCopied[b]) cudaEventDestroy(h->evCopied[b]); if (h->evFree[b]) cudaEventDestroy(h->evFree[b]); }
    if (h->copyStream) cudaStreamDestroy(h->copyStream);
    if (h->d2hStream) cudaStreamDestroy(h->d2hStream);
    if (h->evDone) cudaEventDestroy(h->evDone);
    if (h->ownStream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return ORBX_OK;
}

int orbx_sync(orbx_handle* h)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    int st = 0;
    ORBX_CUDA(cudaMemcpy(&st, h->dStatus, sizeof(int), cudaMemcpyDeviceToHost));
    if (st) {
        cudaMemset(h->dStatus, 0, sizeof(int));
        orbx_set_error("device-side capacity overflow (flags 0x%x): quadtree node/keypoint table too small", st);
        return ORBX_E_CAPACITY;
    }
    return ORBX_OK;
}

int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2,
                    float* inv_sigma2, int32_t* features_per_level)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    for (int i = 0; i < h->params.nlevels; ++i) {
        if (scale) scale[i] = h->scale[i];
        if (inv_scale) inv_scale[i] = h->invScale[i];
        if (sigma2) sigma2[i] = h->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = h->invSigma2[i];
        if (features_per_level) features_per_level[i] = h->perLevel[i];
    }
    return ORBX_OK;
}

int orbx_keypoint_capacity(const orbx_handle* h) { return h ? h->kpCap : ORBX_E_INVALID; }

int orbx_extract_batch(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                       size_t row_stride, int width, int height, int nframes,
                       int images_on_device, int first_slot)
{
    return orbx_extract_batch_to_host(h, images, frame_stride, row_stride, width, height, nframes, images_on_device,
                                      first_slot, nullptr, nullptr, nullptr, 0);
}

} // extern "C"

// ---- staging pool of the batch pipeline for pageable caller buffers -------------------------------------------------
static bool is_pageable(const void* p)
{
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

// f(begin, end) over [0, n) on a few host threads (ORBX_HOST_THREADS, default min(8, cores)); the caller's thread takes a share
template <class F>
static void host_parallel_for(int n, F f)
{
    static const int maxT = [] {
        const char* e = getenv("ORBX_HOST_THREADS");
        const int hw = (int)std::thread::hardware_concurrency();
        return std::max(1, e ? atoi(e) : std::min(8, hw > 0 ? hw : 1));
    }();
    const int T = std::min(n, maxT);
    if (T <= 1) { if (n > 0) f(0, n); return; }
    const int per = (n + T - 1) / T;
    std::vector<std::thread> th;
    for (int t = 1; t < T; ++t) { const int a = t * per, b = std::min(n, a + per); if (a < b) th.emplace_back([=] { f(a, b); }); }
    f(0, std::min(per, n));
    for (std::thread& x : th) x.join();
}

static int ensure_in_pool(orbx_handle* h, size_t bytes)
{
    if (h->hInPoolBytes >= bytes) return ORBX_OK;
    for (int b = 0; b < ORBX_L0_BUFS; ++b) {
        if (h->evInPoolValid[b]) { cudaEventSynchronize(h->evInPool[b]); h->evInPoolValid[b] = false; }
        if (h->hInPool[b]) { cudaFreeHost(h->hInPool[b]); h->hInPool[b] = nullptr; }
    }
    h->hInPoolBytes = 0;
    for (int b = 0; b < ORBX_L0_BUFS; ++b) {
        ORBX_CUDA(cudaMallocHost((void**)&h->hInPool[b], bytes));
        if (!h->evInPool[b]) ORBX_CUDA(cudaEventCreateWithFlags(&h->evInPool[b], cudaEventDisableTiming));
    }
    h->hInPoolBytes = bytes;
    return ORBX_OK;
}

static int ensure_out_pool(orbx_handle* h, size_t bytes)
{
    if (h->hOutPoolBytes >= bytes) return ORBX_OK;
    for (int b = 0; b < 2; ++b) if (h->hOutPool[b]) { cudaFreeHost(h->hOutPool[b]); h->hOutPool[b] = nullptr; }
    h->hOutPoolBytes = 0;
    for (int b = 0; b < 2; ++b) {
        ORBX_CUDA(cudaMallocHost((void**)&h->hOutPool[b], bytes));
        if (!h->evOutPool[b]) ORBX_CUDA(cudaEventCreateWithFlags(&h->evOutPool[b], cudaEventDisableTiming));
    }
    h->hOutPoolBytes = bytes;
    return ORBX_OK;
}

// consecutive-frame SearchForInitialization folded into the chunk pipeline (orbx_extract_match_to_host)
struct MatchOpts { int window; float nnratio; int checkOri; int32_t* matches12; int32_t* nmatches; };

static int batch_to_host_impl(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                              size_t row_stride, int width, int height, int nframes,
                              int images_on_device, int first_slot, orbx_keypoint* kps, uint8_t* desc,
                              int32_t* counts, int capacity, const MatchOpts* mo)
{
    ORBX_NVTX("orbx_extract_batch");
    if (!h || !images) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    const bool toHost = kps || desc || counts || (mo && (mo->matches12 || mo->nmatches));
    if ((kps || desc) && capacity < h->kpCap) { orbx_set_error("output capacity %d < %d", capacity, h->kpCap); return ORBX_E_CAPACITY; }
    if (width <= 0 || height <= 0 || nframes <= 0) return ORBX_OK;   // empty image: silent return
    if (width < 64 || height < 64) { orbx_set_error("image smaller than 64x64 is outside the extractor's domain"); return ORBX_E_INVALID; }
    if (row_stride < (size_t)width) { orbx_set_error("row_stride < width"); return ORBX_E_INVALID; }
    if (first_slot < 0 || first_slot + nframes > h->resultSlots) { orbx_set_error("slots [%d,%d) exceed result store (%d)", first_slot, first_slot + nframes, h->resultSlots); return ORBX_E_CAPACITY; }
    ORBX_CUDA(cudaSetDevice(h->device));
    int rc = orbx_setup_shape(h, width, height);
    if (rc) return rc;
    h->storeUndistorted = false;            // plain extraction: mvKeysUn = mvKeys, bounds = the image (orbx_frames_rgbd sets them after)
    if (mo) { rc = orbx_ensure_match_buffers(h); if (rc) return rc; }
    const bool direct = images_on_device && (row_stride % 4 == 0) && (((uintptr_t)images) % 4 == 0) && (frame_stride % 4 == 0);
    // Host (or unaligned device) images: chunk i+1 is copied into the other half of dLevel0 on the
    // copy stream while the kernels of chunk i run; events order copy -> compute -> buffer reuse.
    if (!direct) {
        cudaEvent_t evStart;
        ORBX_CUDA(cudaEventCreateWithFlags(&evStart, cudaEventDisableTiming));
        ORBX_CUDA(cudaEventRecord(evStart, h->stream));                 // copies must not overtake earlier work on the handle
        ORBX_CUDA(cudaStreamWaitEvent(h->copyStream, evStart, 0));
        ORBX_CUDA(cudaEventDestroy(evStart));
    }
    static const int hostChunkEnv = getenv("ORBX_HOST_CHUNK") ? std::max(1, atoi(getenv("ORBX_HOST_CHUNK"))) : 256;
    const int full = direct ? h->chunkFrames : std::min(h->chunkFrames, hostChunkEnv);
    // Pageable caller buffers: a driver-staged copy blocks the host until it is done, which serialises copy / kernels / copy.
    // They go through the handle's page-locked pools instead: inputs are copied there by a few host threads and uploaded from
    // there, results land there and are handed over one chunk later (25 k -> 82 k frames/s at 640x480).
    static const bool noPool = getenv("ORBX_NO_HOST_POOL") != nullptr;
    const bool stageIn = !noPool && !images_on_device && is_pageable(images);
    const bool stageOut = !noPool && toHost && (is_pageable(kps) || is_pageable(desc) || is_pageable(counts) ||
                                                (mo && (is_pageable(mo->matches12) || is_pageable(mo->nmatches))));
    const size_t tightFrame = (size_t)width * height;
    if (stageIn) { rc = ensure_in_pool(h, (size_t)full * tightFrame); if (rc) return rc; }
    // result pool layout of one chunk: [mvKeys][descriptors][counts][matches12][nmatches], each sized for `full` frames
    const size_t oK = 0, oD = oK + (size_t)full * h->kpCap * sizeof(orbx_keypoint), oC = oD + (size_t)full * h->kpCap * 32,
                 oM = oC + (((size_t)full * 4 + 15) & ~(size_t)15), oN = oM + (size_t)full * h->kpCap * 4, outBytes = oN + (size_t)full * 4;
    if (stageOut) { rc = ensure_out_pool(h, outBytes); if (rc) return rc; }
    struct Pending { bool live; int f0, nf, p0, np, ob; } pend = {false, 0, 0, 0, 0, 0};
    auto flush = [&](const Pending& P) -> int {       // hand a finished chunk's results from the pool to the caller
        ORBX_CUDA(cudaEventSynchronize(h->evOutPool[P.ob]));
        const uint8_t* S = h->hOutPool[P.ob];
        const int32_t* cnt = reinterpret_cast<const int32_t*>(S + oC);
        host_parallel_for(P.nf, [&](int a, int b) {
            for (int f = a; f < b; ++f) {
                const size_t c = (size_t)std::min(std::max(cnt[f], 0), h->kpCap);     // only the records that exist
                if (kps) memcpy(kps + (size_t)(P.f0 + f) * capacity, S + oK + (size_t)f * h->kpCap * sizeof(orbx_keypoint), c * sizeof(orbx_keypoint));
                if (desc) memcpy(desc + (size_t)(P.f0 + f) * capacity * 32, S + oD + (size_t)f * h->kpCap * 32, c * 32);
            }
        });
        if (counts) memcpy(counts + P.f0, cnt, (size_t)P.nf * 4);
        if (mo && P.np > 0) {
            if (mo->matches12)
                host_parallel_for(P.np, [&](int a, int b) {
                    for (int q = a; q < b; ++q) memcpy(mo->matches12 + (size_t)(P.p0 + q) * capacity, S + oM + (size_t)q * h->kpCap * 4, (size_t)h->kpCap * 4);
                });
            if (mo->nmatches) memcpy(mo->nmatches + P.p0, S + oN, (size_t)P.np * 4);
        }
        return ORBX_OK;
    };
    // ORBX_PIPE_TRACE=1: print the chunk timeline (copy / kernels / result copy, ms from the first copy) to stderr
    const bool trace = !direct && getenv("ORBX_PIPE_TRACE") != nullptr;
    std::vector<cudaEvent_t> tev;
    std::vector<int> tnf;
    auto mark = [&](cudaStream_t st) { if (!trace) return; cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); tev.push_back(e); };
    // the chunk schedule (ramped head, tapered tail: see below), known up front so that a copy thread can run one chunk ahead
    std::vector<std::pair<int, int>> sched;                     // (first frame, frames)
    for (int f0 = 0, nf = 0, c = 0; f0 < nframes; f0 += nf, ++c) {
        int cf = full;
        if (!direct && nframes > full) cf = c == 0 ? std::max(1, cf / 4) : (c == 1 ? std::max(1, cf / 2) : cf);
        nf = std::min(cf, nframes - f0);
        if (!direct && toHost && nframes > full && nf == nframes - f0 && nf > std::max(1, full / 4))
            nf = std::max(std::max(1, full / 4), (nf + 1) / 2);
        sched.push_back({f0, nf});
    }
    // pageable inputs: chunk c is copied into pool buffer c % ORBX_L0_BUFS (last read by the upload of chunk c - ORBX_L0_BUFS) by
    // a copy thread that runs while the caller's thread queues chunk c - 1 and hands over the results of chunk c - 2
    auto fill_pool = [&](int c) {
        cudaSetDevice(h->device);
        const int b = c % ORBX_L0_BUFS, f0 = sched[c].first, nf = sched[c].second;
        if (h->evInPoolValid[b]) cudaEventSynchronize(h->evInPool[b]);
        uint8_t* pool = h->hInPool[b];
        const uint8_t* src = images + (size_t)f0 * frame_stride;
        if (row_stride == (size_t)width)
            host_parallel_for(nf, [&](int a, int e) { for (int f = a; f < e; ++f) memcpy(pool + (size_t)f * tightFrame, src + (size_t)f * frame_stride, tightFrame); });
        else
            host_parallel_for(nf, [&](int a, int e) {
                for (int f = a; f < e; ++f)
                    for (int y = 0; y < height; ++y) memcpy(pool + (size_t)f * tightFrame + (size_t)y * width, src + (size_t)f * frame_stride + (size_t)y * row_stride, (size_t)width);
            });
    };
    struct Joiner { std::thread t; ~Joiner() { if (t.joinable()) t.join(); } } ahead;      // joins on every exit path
    if (stageIn) fill_pool(0);
    for (int chunkNo = 0; chunkNo < (int)sched.size(); ++chunkNo) {
        const int f0 = sched[chunkNo].first, nf = sched[chunkNo].second;
        // host images: ramp the first two chunks (1/4, 1/2 of a chunk) so that less of the first
        // host->device copy is exposed before the kernels start
        // The pipeline from host images is bound by the copy (55 GB/s = 5.9 us per 640x480 frame against 4.8 us of kernels), and
        // what it exposes is the first copy and the kernels + result copy of the last chunk: it runs on smaller chunks
        // (hostChunk, 256 frames unless ORBX_HOST_CHUNK says otherwise) than device-resident batches, which only pay the
        // per-chunk launch ramps (measured: 256 -> 156 k frames/s end to end, 512 -> 151 k, 1024 -> 140 k).
        // ... and the tail is tapered the same way: the kernels and the device->host copy of the last chunk are exposed
        const uint8_t* src = images + (size_t)f0 * frame_stride;
        Level0 l0;
        l0.base1 = nullptr; l0.stereo = 0;
        if (direct) { l0.base = src; l0.frameStride = (long long)frame_stride; l0.pitch = (int)row_stride; }
        else {
            const int b = chunkNo % ORBX_L0_BUFS;
            uint8_t* dstBuf = h->dLevel0 + (size_t)b * h->chunkFrames * h->l0FrameBytes;
            if (h->evFreeValid[b]) ORBX_CUDA(cudaStreamWaitEvent(h->copyStream, h->evFree[b], 0));
            const uint8_t* csrc = src;
            size_t cFrameStride = frame_stride, cRowStride = row_stride;
            if (stageIn) {
                if (ahead.t.joinable()) ahead.t.join();              // this chunk's images are in pool buffer b
                csrc = h->hInPool[b]; cFrameStride = tightFrame; cRowStride = (size_t)width;
            }
            mark(h->copyStream);
            rc = orbx_copy_frames(h, dstBuf, h->l0FrameBytes, h->dStageLin + (size_t)b * h->chunkFrames * h->l0FrameBytes, (size_t)h->chunkFrames * h->l0FrameBytes, csrc, cFrameStride,
                                  cRowStride, width, height, nf, images_on_device != 0, h->copyStream, &l0.frameStride);
            if (rc) return rc;
            if (stageIn) {
                ORBX_CUDA(cudaEventRecord(h->evInPool[b], h->copyStream)); h->evInPoolValid[b] = true;
                if (chunkNo + 1 < (int)sched.size()) {
                    if (ORBX_L0_BUFS >= 2) ahead.t = std::thread(fill_pool, chunkNo + 1);      // the other buffer: it waits for that buffer's last upload itself
                    else fill_pool(chunkNo + 1);
                }
            }
            ORBX_CUDA(cudaEventRecord(h->evCopied[b], h->copyStream));
            mark(h->copyStream);
            ORBX_CUDA(cudaStreamWaitEvent(h->stream, h->evCopied[b], 0));
            l0.base = dstBuf; l0.pitch = h->l0Pitch;
        }
        mark(h->stream);
        rc = orbx_run_chunk(h, l0, nf, first_slot + f0);
        if (rc) return rc;
        if (!direct) { ORBX_CUDA(cudaEventRecord(h->evFree[chunkNo % ORBX_L0_BUFS], h->stream)); h->evFreeValid[chunkNo % ORBX_L0_BUFS] = true; }
        // pairs (f-1, f) for the frames of this chunk; the grid of frame f0-1 was built with the previous chunk
        const int p0 = std::max(f0 - 1, 0), np = f0 + nf - 1 - p0;
        if (mo) {
            rc = orbx_match_chunk(h, first_slot + f0, nf, first_slot + p0, np, mo->window, mo->nnratio, mo->checkOri);
            if (rc) return rc;
        }
        mark(h->stream);
        tnf.push_back(nf);
        if (toHost) {
            // results of this chunk go back on the copy stream while the next chunk computes
            const int s0 = first_slot + f0;
            ORBX_CUDA(cudaEventRecord(h->evDone, h->stream));
            ORBX_CUDA(cudaStreamWaitEvent(h->d2hStream, h->evDone, 0));
            if (stageOut) {
                // into the pool (contiguous copies), handed to the caller after the NEXT chunk has been queued
                const int ob = chunkNo & 1;
                uint8_t* S = h->hOutPool[ob];               // free: its previous chunk (chunkNo - 2) was flushed one iteration ago
                if (kps) ORBX_CUDA(cudaMemcpyAsync(S + oK, h->dKps + (size_t)s0 * h->kpCap, (size_t)nf * h->kpCap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, h->d2hStream));
                if (desc) ORBX_CUDA(cudaMemcpyAsync(S + oD, h->dDesc + (size_t)s0 * h->kpCap * 32, (size_t)nf * h->kpCap * 32, cudaMemcpyDeviceToHost, h->d2hStream));
                ORBX_CUDA(cudaMemcpyAsync(S + oC, h->dCount + s0, (size_t)nf * 4, cudaMemcpyDeviceToHost, h->d2hStream));
                if (mo && np > 0) {
                    const int ps = first_slot + p0;
                    if (mo->matches12) ORBX_CUDA(cudaMemcpyAsync(S + oM, h->dMatches12 + (size_t)ps * h->kpCap, (size_t)np * h->kpCap * 4, cudaMemcpyDeviceToHost, h->d2hStream));
                    if (mo->nmatches) ORBX_CUDA(cudaMemcpyAsync(S + oN, h->dNMatches + ps, (size_t)np * 4, cudaMemcpyDeviceToHost, h->d2hStream));
                }
                ORBX_CUDA(cudaEventRecord(h->evOutPool[ob], h->d2hStream));
                mark(h->d2hStream);
                if (pend.live) { rc = flush(pend); if (rc) return rc; }
                pend = Pending{true, f0, nf, p0, np, ob};
                continue;
            }
            if (kps)
                ORBX_CUDA(cudaMemcpy2DAsync(kps + (size_t)f0 * capacity, (size_t)capacity * sizeof(orbx_keypoint), h->dKps + (size_t)s0 * h->kpCap,
                                            (size_t)h->kpCap * sizeof(orbx_keypoint), (size_t)h->kpCap * sizeof(orbx_keypoint), nf,
                                            cudaMemcpyDeviceToHost, h->d2hStream));
            if (desc)
                ORBX_CUDA(cudaMemcpy2DAsync(desc + (size_t)f0 * capacity * 32, (size_t)capacity * 32, h->dDesc + (size_t)s0 * h->kpCap * 32,
                                            (size_t)h->kpCap * 32, (size_t)h->kpCap * 32, nf, cudaMemcpyDeviceToHost, h->d2hStream));
            if (counts) ORBX_CUDA(cudaMemcpyAsync(counts + f0, h->dCount + s0, nf * sizeof(int), cudaMemcpyDeviceToHost, h->d2hStream));
            if (mo && np > 0) {
                const int ps = first_slot + p0;
                if (mo->matches12)
                    ORBX_CUDA(cudaMemcpy2DAsync(mo->matches12 + (size_t)p0 * capacity, (size_t)capacity * 4, h->dMatches12 + (size_t)ps * h->kpCap,
                                                (size_t)h->kpCap * 4, (size_t)h->kpCap * 4, np, cudaMemcpyDeviceToHost, h->d2hStream));
                if (mo->nmatches) ORBX_CUDA(cudaMemcpyAsync(mo->nmatches + p0, h->dNMatches + ps, np * sizeof(int), cudaMemcpyDeviceToHost, h->d2hStream));
            }
            mark(h->d2hStream);
        }
    }
    if (toHost) {
        if (pend.live) { rc = flush(pend); if (rc) return rc; }
        ORBX_CUDA(cudaStreamSynchronize(h->d2hStream));
        rc = orbx_sync(h);
        if (trace && rc == ORBX_OK) {
            cudaStreamSynchronize(h->copyStream);
            const size_t per = 5;
            for (size_t c = 0; c * per + per <= tev.size(); ++c) {
                float t[5];
                for (int k = 0; k < 5; ++k) cudaEventElapsedTime(&t[k], tev[0], tev[c * per + k]);
                fprintf(stderr, "chunk %2zu nf %4d  copy %7.3f-%7.3f  kernels %7.3f-%7.3f  d2h done %7.3f\n", c, tnf[c], t[0], t[1], t[2], t[3], t[4]);
            }
        }
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
        return rc;
    }
    return ORBX_OK;
}

extern "C" {

int orbx_extract_batch_to_host(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                               size_t row_stride, int width, int height, int nframes,
                               int images_on_device, int first_slot, orbx_keypoint* kps, uint8_t* desc,
                               int32_t* counts, int capacity)
{
    return batch_to_host_impl(h, images, frame_stride, row_stride, width, height, nframes, images_on_device, first_slot, kps, desc,
                              counts, capacity, nullptr);
}

int orbx_extract_match_to_host(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                               size_t row_stride, int width, int height, int nframes,
                               int images_on_device, int first_slot, orbx_keypoint* kps, uint8_t* desc,
                               int32_t* counts, int capacity, int window_size, float nnratio,
                               int check_orientation, int32_t* matches12, int32_t* nmatches)
{
    if (matches12 && capacity < (h ? h->kpCap : 0)) { orbx_set_error("matches12 capacity %d < %d", capacity, h->kpCap); return ORBX_E_CAPACITY; }
    const MatchOpts mo = {window_size, nnratio, check_orientation, matches12, nmatches};
    return batch_to_host_impl(h, images, frame_stride, row_stride, width, height, nframes, images_on_device, first_slot, kps, desc,
                              counts, capacity, &mo);
}

int orbx_download(orbx_handle* h, int first_slot, int nframes, orbx_keypoint* kps, uint8_t* desc,
                  int32_t* counts, int capacity)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (nframes <= 0) return ORBX_OK;
    if (first_slot < 0 || first_slot + nframes > h->resultSlots) { orbx_set_error("slot range out of bounds"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    std::vector<int> cnt(nframes);
    ORBX_CUDA(cudaMemcpyAsync(cnt.data(), h->dCount + first_slot, nframes * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    int rc = orbx_sync(h);
    if (rc) return rc;
    int maxc = 0;
    for (int i = 0; i < nframes; ++i) { maxc = std::max(maxc, cnt[i]); if (counts) counts[i] = cnt[i]; }
    if ((kps || desc) && maxc > capacity) { orbx_set_error("capacity %d < %d keypoints", capacity, maxc); return ORBX_E_CAPACITY; }
    const int rows = std::min(capacity, h->kpCap);
    if (kps)
        ORBX_CUDA(cudaMemcpy2DAsync(kps, (size_t)capacity * sizeof(orbx_keypoint), h->dKps + (size_t)first_slot * h->kpCap,
                                    (size_t)h->kpCap * sizeof(orbx_keypoint), (size_t)rows * sizeof(orbx_keypoint), nframes,
                                    cudaMemcpyDeviceToHost, h->stream));
    if (desc)
        ORBX_CUDA(cudaMemcpy2DAsync(desc, (size_t)capacity * 32, h->dDesc + (size_t)first_slot * h->kpCap * 32,
                                    (size_t)h->kpCap * 32, (size_t)rows * 32, nframes, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

} // extern "C"

int orbx_ensure_host_stage(orbx_handle* h, size_t need)
{
    if (h->hStageBytes >= need) return ORBX_OK;
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (h->hStage) cudaFreeHost(h->hStage);
    h->hStage = nullptr; h->hStageBytes = 0;
    ORBX_CUDA(cudaHostAlloc(&h->hStage, need, cudaHostAllocMapped));       // kernels may read it in place (UVA: same pointer)
    h->hStageBytes = need;
    return ORBX_OK;
}

OrbxPackLayout orbx_pack_layout(int kpCap)
{
    const size_t cap = (size_t)kpCap;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    OrbxPackLayout L;
    L.kps = 256;
    L.desc = L.kps + up(cap * sizeof(orbx_keypoint));
    L.kun = L.desc + up(cap * 32);
    L.ur = L.kun + up(cap * sizeof(orbx_keypoint));
    L.z = L.ur + up(cap * 4);
    L.total = L.z + up(cap * 4);
    return L;
}

int orbx_ensure_pack(orbx_handle* h)
{
    const size_t need = orbx_pack_layout(h->kpCap).total;
    if (!h->evPack) ORBX_CUDA(cudaEventCreateWithFlags(&h->evPack, cudaEventDisableTiming));
    if (h->hPackBytes >= need) return ORBX_OK;
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    orbx_drop_graph(h);                                    // the captured graph writes through the old mapping
    if (h->hPack) cudaFreeHost(h->hPack);
    h->hPack = nullptr; h->hPackBytes = 0;
    ORBX_CUDA(cudaHostAlloc((void**)&h->hPack, need, cudaHostAllocMapped));
    h->hPackBytes = need;
    return ORBX_OK;
}

// Results of slot 0 written straight into the caller-visible (mapped, pinned) pack: count and status word, then the `count`
// keypoint records and descriptors.  Last node of the single-frame graph: the host only synchronises and reads.
__global__ void __launch_bounds__(128)
pack_single_kernel(const orbx_keypoint* __restrict__ kps, const uint8_t* __restrict__ desc, const int* __restrict__ count,
                   const int* __restrict__ status, int cap, uint8_t* __restrict__ pack, size_t offKps, size_t offDesc)
{
    const int k = blockIdx.x * 128 + threadIdx.x, n = min(count[0], cap);
    if (k == 0) { reinterpret_cast<int*>(pack)[0] = count[0]; reinterpret_cast<int*>(pack)[1] = status[0]; }
    if (k >= n) return;
    const uint32_t* s = reinterpret_cast<const uint32_t*>(kps + k);
    uint32_t* d = reinterpret_cast<uint32_t*>(pack + offKps) + (size_t)k * 7;
#pragma unroll
    for (int i = 0; i < 7; ++i) d[i] = s[i];
    const uint4* ds = reinterpret_cast<const uint4*>(desc + (size_t)k * 32);
    uint4* dd = reinterpret_cast<uint4*>(pack + offDesc + (size_t)k * 32);
    dd[0] = ds[0]; dd[1] = ds[1];
}

static int launch_pack(orbx_handle* h)
{
    const OrbxPackLayout L = orbx_pack_layout(h->kpCap);
    pack_single_kernel<<<(h->kpCap + 127) / 128, 128, 0, h->stream>>>(h->dKps, h->dDesc, h->dCount, h->dStatus, h->kpCap, h->hPack, L.kps, L.desc);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// The kernels of ONE frame whose level 0 already sits in h->dLevel0 (ordered on the handle's stream): replayed from a CUDA
// graph captured on the first call of an image shape; plain launches when stage timing is on, the stream is the caller's
// (it may be one that cannot be captured) or a test forces kernel parameters through the environment.  Results: slot 0.
int orbx_launch_single_frame(orbx_handle* h)
{
    cudaStream_t st = h->stream;
    int rc = orbx_ensure_pack(h);
    if (rc) return rc;
    h->evFreeValid[0] = false;                // buffer 0 is now ordered by the handle's stream alone
    Level0 l0;
    l0.base = h->dLevel0; l0.base1 = nullptr; l0.stereo = 0; l0.frameStride = h->l0FrameBytes; l0.pitch = h->l0Pitch;
    const bool useGraph = !h->profiling && h->ownStream && !getenv("ORBX_NO_GRAPH") && !getenv("ORBX_OCTREE_KEYCAP") && !getenv("ORBX_RS_TARGET");
    if (useGraph && !h->graph1Valid) {
        // first frame of this shape: run once plainly (sets function attributes, uploads the per-device tables),
        // then capture the same launches
        static const bool forkLevels = !getenv("ORBX_LAT_NOFORK");
        rc = orbx_run_chunk(h, l0, 1, 0);
        if (rc) return rc;
        if (forkLevels) { rc = orbx_run_frame_forked(h, l0, 0); if (rc) return rc; }    // creates the side streams outside the capture
        rc = launch_pack(h); if (rc) return rc;
        long long before[ORBX_NSTAGES];
        for (int s = 0; s < ORBX_NSTAGES; ++s) before[s] = h->stageLaunches[s];
        cudaGraph_t graph = nullptr;
        ORBX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
        rc = forkLevels ? orbx_run_frame_forked(h, l0, 0) : orbx_run_chunk(h, l0, 1, 0);
        if (!rc) rc = launch_pack(h);
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        for (int s = 0; s < ORBX_NSTAGES; ++s) { h->graph1Launches[s] = h->stageLaunches[s] - before[s]; h->stageLaunches[s] = before[s]; }
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (ce != cudaSuccess) { orbx_set_error("CUDA graph capture: %s", cudaGetErrorString(ce)); return ORBX_E_CUDA; }
        const cudaError_t ie = cudaGraphInstantiate(&h->graph1, graph, 0);
        cudaGraphDestroy(graph);
        if (ie != cudaSuccess) { orbx_set_error("cudaGraphInstantiate: %s", cudaGetErrorString(ie)); return ORBX_E_CUDA; }
        h->graph1Valid = true;
    } else if (useGraph) {
        ORBX_CUDA(cudaGraphLaunch(h->graph1, st));
        for (int s = 0; s < ORBX_NSTAGES; ++s) h->stageLaunches[s] += h->graph1Launches[s];
        h->lastL0 = l0;
        h->lastChunkFrames = 1;
    } else {
        rc = orbx_run_chunk(h, l0, 1, 0);
        if (rc) return rc;
        rc = launch_pack(h); if (rc) return rc;
    }
    return ORBX_OK;
}

extern "C" {

// Single frame, host buffers: the call Tracking makes once per image (Frame::ExtractORB, src/Frame.cc:247-253).
// Latency path: the image goes through pinned staging in ONE host->device copy, the kernels of the frame are
// replayed from a CUDA graph captured once per image shape, and the graph's last node writes count, status, keypoints
// and descriptors straight into a pinned, device-mapped buffer: no device->host copy commands (four of them cost ~30 us),
// one stream synchronisation.
int orbx_extract(orbx_handle* h, const uint8_t* image, int width, int height, size_t row_stride,
                 orbx_keypoint* kps, uint8_t* desc, int capacity, int* n)
{
    ORBX_NVTX("orbx_extract");
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || !image) return ORBX_OK;     // _image.empty(): return, outputs untouched
    if (width < 64 || height < 64) { orbx_set_error("image smaller than 64x64 is outside the extractor's domain"); return ORBX_E_INVALID; }
    if (row_stride < (size_t)width) { orbx_set_error("row_stride < width"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    int rc = orbx_setup_shape(h, width, height);
    if (rc) return rc;
    h->storeUndistorted = false;
    cudaStream_t st = h->stream;
    // pinned staging for a pageable image (at the workspace pitch); results come back through the mapped pack
    const size_t imgBytes = (size_t)h->l0Pitch * height;
    rc = orbx_ensure_host_stage(h, imgBytes);
    if (rc) return rc;
    uint8_t* stage = (uint8_t*)h->hStage;
    cudaPointerAttributes pa;
    const bool srcPinned = cudaPointerGetAttributes(&pa, image) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();                       // an unregistered host pointer is not an error here
    if (srcPinned) {
        // the caller's image is page-locked: the DMA engine reads it in place
        ORBX_CUDA(cudaMemcpy2DAsync(h->dLevel0, h->l0Pitch, image, row_stride, width, height, cudaMemcpyHostToDevice, st));
    } else {
        // pageable image (a cv::Mat): one memcpy into the pinned staging at the workspace pitch, one DMA (row slices
        // that overlap the two were measured slower: four copy launches cost more than the overlap saves)
        if ((size_t)h->l0Pitch == row_stride) memcpy(stage, image, (size_t)row_stride * (height - 1) + width);
        else for (int y = 0; y < height; ++y) memcpy(stage + (size_t)y * h->l0Pitch, image + (size_t)y * row_stride, width);
        ORBX_CUDA(cudaMemcpyAsync(h->dLevel0, stage, imgBytes, cudaMemcpyHostToDevice, st));
    }
    rc = orbx_launch_single_frame(h);
    if (rc) return rc;
    ORBX_CUDA(cudaStreamSynchronize(st));                 // the graph's last node wrote count, status, keypoints and descriptors into hPack
    const OrbxPackLayout PL = orbx_pack_layout(h->kpCap);
    const int32_t* hdr = (const int32_t*)h->hPack;
    if (hdr[1]) {
        ORBX_CUDA(cudaMemsetAsync(h->dStatus, 0, sizeof(int), st));
        orbx_set_error("device-side capacity overflow (flags 0x%x): quadtree node/keypoint table too small", hdr[1]);
        return ORBX_E_CAPACITY;
    }
    const int cnt = hdr[0];
    if ((kps || desc) && cnt > capacity) { orbx_set_error("capacity %d < %d keypoints", capacity, cnt); return ORBX_E_CAPACITY; }
    if (kps && cnt > 0) memcpy(kps, h->hPack + PL.kps, (size_t)cnt * sizeof(orbx_keypoint));
    if (desc && cnt > 0) memcpy(desc, h->hPack + PL.desc, (size_t)cnt * 32);
    if (n) *n = cnt;
    return ORBX_OK;
}

int orbx_get_level(orbx_handle* h, int frame, int level, int blurred, int border, uint8_t* dst,
                   size_t dst_stride, int* width, int* height)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (h->lastChunkFrames <= 0 || frame < 0 || frame >= h->lastChunkFrames || level < 0 || level >= h->geom.nlevels) {
        orbx_set_error("no such frame/level in the last chunk"); return ORBX_E_STATE;
    }
    if (border < 0 || border > 64) { orbx_set_error("border out of range"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    const LevelGeom& L = h->geom.L[level];
    if (width) *width = L.w;
    if (height) *height = L.h;
    if (!dst) return ORBX_OK;
    if (dst_stride < (size_t)(L.w + 2 * border)) { orbx_set_error("dst_stride too small"); return ORBX_E_INVALID; }
    const uint8_t* src; size_t sp;
    if (blurred) { src = h->dBlur + (size_t)frame * h->geom.pyrFrameBytes + L.off; sp = L.pitch; }
    else if (level == 0) { src = l0_frame(h->lastL0, frame); sp = h->lastL0.pitch; }
    else { src = h->dRaw + (size_t)frame * h->geom.pyrFrameBytes + L.off; sp = L.pitch; }
    ORBX_CUDA(cudaMemcpy2DAsync(dst + (size_t)border * dst_stride + border, dst_stride, src, sp, L.w, L.h, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (border > 0) {
        // copyMakeBorder(..., BORDER_REFLECT_101) of the level (src/ORBextractor.cc:1122-1128), host side
        auto refl = [](int i, int n) { if (n == 1) return 0; while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * (n - 1) - i; return i; };
        uint8_t* core = dst + (size_t)border * dst_stride + border;
        for (int y = 0; y < L.h; ++y) {
            uint8_t* row = core + (size_t)y * dst_stride;
            for (int x = 1; x <= border; ++x) { row[-x] = row[refl(-x, L.w)]; row[L.w - 1 + x] = row[refl(L.w - 1 + x, L.w)]; }
        }
        for (int y = 1; y <= border; ++y) {
            memcpy(dst + (size_t)(border - y) * dst_stride, dst + (size_t)(border + refl(-y, L.h)) * dst_stride, L.w + 2 * border);
            memcpy(dst + (size_t)(border + L.h - 1 + y) * dst_stride, dst + (size_t)(border + refl(L.h - 1 + y, L.h)) * dst_stride, L.w + 2 * border);
        }
    }
    return ORBX_OK;
}

int orbx_get_candidates(orbx_handle* h, int frame, int level, float* xyr, int capacity, int* n)
{
    if (!h || !n) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (h->lastChunkFrames <= 0 || frame < 0 || frame >= h->lastChunkFrames || level < 0 || level >= h->geom.nlevels) {
        orbx_set_error("no such frame/level in the last chunk"); return ORBX_E_STATE;
    }
    ORBX_CUDA(cudaSetDevice(h->device));
    const Geom& g = h->geom;
    const LevelGeom& L = g.L[level];
    std::vector<int2> seg(L.nCells);
    ORBX_CUDA(cudaMemcpyAsync(seg.data(), h->dCellSeg + (size_t)frame * g.ncells + L.cellBeg, L.nCells * sizeof(int2), cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    // the level's whole candidate region: every cell owns a fixed segment of it (debug / test entry point)
    std::vector<uint32_t> c(std::max(L.candCap, 1));
    if (L.candCap) ORBX_CUDA(cudaMemcpy(c.data(), h->dCand + (size_t)frame * g.candTotal + L.candOff, (size_t)L.candCap * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    int k = 0;
    for (int ci = 0; ci < L.nCells; ++ci)
        for (int i = 0; i < seg[ci].y; ++i, ++k) {
            if (k < capacity && xyr) {
                const uint32_t p = c[seg[ci].x + i];
                xyr[3 * k] = (float)(p & 0xfff); xyr[3 * k + 1] = (float)((p >> 12) & 0xfff); xyr[3 * k + 2] = (float)(p >> 24);
            }
        }
    *n = k;
    return ORBX_OK;
}

} // extern "C"
