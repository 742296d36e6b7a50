// orbx_match_common.cuh — device helpers shared by the matcher kernels (orbx_match.cu,
// orbx_stereo.cu, orbx_knn.cu).  Not part of the public ABI.
#ifndef ORBX_MATCH_COMMON_CUH
#define ORBX_MATCH_COMMON_CUH

#include "orbx_internal.h"
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>

typedef uint8_t u8;
#define FULL 0xffffffffu

__device__ __forceinline__ int hamming256(const uint32_t a[8], const uint32_t* __restrict__ b)
{
    const uint4 b0 = *reinterpret_cast<const uint4*>(b), b1 = *reinterpret_cast<const uint4*>(b + 4);
    return __popc(a[0] ^ b0.x) + __popc(a[1] ^ b0.y) + __popc(a[2] ^ b0.z) + __popc(a[3] ^ b0.w) +
           __popc(a[4] ^ b1.x) + __popc(a[5] ^ b1.y) + __popc(a[6] ^ b1.z) + __popc(a[7] ^ b1.w);
}

// both descriptors in registers
__device__ __forceinline__ int hamming256_regs(const uint32_t a[8], const uint32_t b[8])
{
    int d = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) d += __popc(a[i] ^ b[i]);
    return d;
}

__device__ __forceinline__ void load_desc(uint32_t a[8], const u8* __restrict__ p)
{
    const uint4 v0 = *reinterpret_cast<const uint4*>(p), v1 = *reinterpret_cast<const uint4*>(p + 16);
    a[0] = v0.x; a[1] = v0.y; a[2] = v0.z; a[3] = v0.w; a[4] = v1.x; a[5] = v1.y; a[6] = v1.z; a[7] = v1.w;
}

// Ascending bitonic sort of cap (a power of two, <= 4 * NT) 64-bit keys in shared memory by a block of NT threads (NT a
// multiple of 32, every thread of the block calls).  Thread t keeps elements t, t + NT, ... in registers: exchanges at
// distances >= NT stay inside the thread, distances < 32 are warp shuffles, only the ones between go through shared memory --
// 15 of the 55 steps of a 1024-key sort.  (The plain shared-memory version, a __syncthreads per step, measured 35 us for
// 1024 keys; this one is the sort of the per-frame BoW / triangulation calls.)
// NA arrays of the same cap are sorted side by side (one set of barriers, independent instruction streams)
template <int NT, int E, int NA>
__device__ __forceinline__ void block_sort64_impl(unsigned long long* const (&arr)[NA], int cap)
{
    const int tid = threadIdx.x;
    unsigned long long v[NA][E];
#pragma unroll
    for (int s = 0; s < NA; ++s)
#pragma unroll
        for (int r = 0; r < E; ++r) v[s][r] = r * NT + tid < cap ? arr[s][r * NT + tid] : ~0ull;
    // keep the smaller of (mine, other) when takeMin, else the larger (equal keys: either)
    auto pick = [](unsigned long long mine, unsigned long long other, bool takeMin) { return ((other < mine) == takeMin) ? other : mine; };
    for (int k = 2; k <= cap; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= NT) {                                   // partners live in the same thread (E > 1 only)
                auto cx = [&](unsigned long long& x, unsigned long long& y, int r) {      // x = element r * NT + tid, y its partner above
                    const bool asc = ((r * NT + tid) & k) == 0;
                    const unsigned long long lo = x < y ? x : y, hi = x < y ? y : x;
                    x = asc ? lo : hi; y = asc ? hi : lo;
                };
#pragma unroll
                for (int s = 0; s < NA; ++s) {
                    if (E >= 2 && j == NT) { cx(v[s][0], v[s][E >= 2 ? 1 : 0], 0); if (E == 4) cx(v[s][E == 4 ? 2 : 0], v[s][E == 4 ? 3 : 0], 2); }
                    else if (E == 4) { cx(v[s][0], v[s][E == 4 ? 2 : 0], 0); cx(v[s][E == 4 ? 1 : 0], v[s][E == 4 ? 3 : 0], 1); }     // j == 2 * NT
                }
            } else if (j >= 32) {                            // through shared memory
                __syncthreads();
#pragma unroll
                for (int s = 0; s < NA; ++s)
#pragma unroll
                    for (int r = 0; r < E; ++r) if (r * NT + tid < cap) arr[s][r * NT + tid] = v[s][r];
                __syncthreads();
#pragma unroll
                for (int s = 0; s < NA; ++s)
#pragma unroll
                    for (int r = 0; r < E; ++r) {
                        const int e = r * NT + tid;
                        if (e < cap) v[s][r] = pick(v[s][r], arr[s][e ^ j], ((e & j) == 0) == ((e & k) == 0));
                    }
            } else {                                         // inside the warp
#pragma unroll
                for (int s = 0; s < NA; ++s)
#pragma unroll
                    for (int r = 0; r < E; ++r) {
                        const int e = r * NT + tid;
                        v[s][r] = pick(v[s][r], __shfl_xor_sync(FULL, v[s][r], j), ((e & j) == 0) == ((e & k) == 0));
                    }
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < NA; ++s)
#pragma unroll
        for (int r = 0; r < E; ++r) if (r * NT + tid < cap) arr[s][r * NT + tid] = v[s][r];
    __syncthreads();
}

template <int NT>
__device__ __forceinline__ void block_sort64(unsigned long long* a, int cap)
{
    unsigned long long* const arr[1] = {a};
    if (cap > 2 * NT) block_sort64_impl<NT, 4, 1>(arr, cap);
    else if (cap > NT) block_sort64_impl<NT, 2, 1>(arr, cap);
    else block_sort64_impl<NT, 1, 1>(arr, cap);
}

// two arrays of the same cap (<= 2 * NT)
template <int NT>
__device__ __forceinline__ void block_sort64_pair(unsigned long long* a, unsigned long long* b, int cap)
{
    unsigned long long* const arr[2] = {a, b};
    if (cap > NT) block_sort64_impl<NT, 2, 2>(arr, cap);
    else block_sort64_impl<NT, 1, 2>(arr, cap);
}

struct GridParams { float minX, minY, invW, invH; };

// frames live in a store: kps[slot*cap + i], desc[(slot*cap + i)*32], count[slot],
// gridStart[slot*(CELLS+1) + c], gridIdx[slot*cap + j]
struct FrameStore {
    const orbx_keypoint* kps;
    const u8* desc;
    const int* count;
    unsigned short* gridStart;
    unsigned short* gridIdx;
    int cap;
};

// Frame::GetFeaturesInArea cell window (src/Frame.cc:332-346); returns false when empty.
__device__ __forceinline__ bool area_cells(const GridParams& gp, float x, float y, float r, int& cx0, int& cx1,
                                           int& cy0, int& cy1)
{
    cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
    if (cx0 >= ORBX_GRID_COLS) return false;
    cx1 = min(ORBX_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
    if (cx1 < 0) return false;
    cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
    if (cy0 >= ORBX_GRID_ROWS) return false;
    cy1 = min(ORBX_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
    if (cy1 < 0) return false;
    return true;
}

// top-2 merge across the warp: best = min (dist, key), second = second smallest distance
__device__ __forceinline__ void warp_top2(int& best, unsigned& key, int& second)
{
    for (int d = 16; d > 0; d >>= 1) {
        const int ob = __shfl_xor_sync(FULL, best, d);
        const unsigned ok = __shfl_xor_sync(FULL, key, d);
        const int os = __shfl_xor_sync(FULL, second, d);
        if (ob < best || (ob == best && ok < key)) { second = min(min(second, os), best); best = ob; key = ok; }
        else second = min(min(second, os), ob);
    }
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642)
static __device__ void three_maxima(const int* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < L; ++i) {
        const int s = histo[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

__device__ __forceinline__ int rot_bin(float a1, float a2)
{
    const float factor = 1.0f / ORBX_HISTO_LENGTH;
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, factor));
    if (bin == ORBX_HISTO_LENGTH) bin = 0;
    return bin;
}


// ---- host-side helpers shared by the stateless matcher entry points (orbx_match.cu, orbx_kfmatch.cu) ----
GridParams orbx_grid_params(float minX, float maxX, float minY, float maxY);
int orbx_launch_grid_build(const FrameStore& fs, int firstSlot, int nframes, const GridParams& gp, cudaStream_t st);

// device scratch that releases itself (per-thread block cache, orbx_scratch_alloc: no cudaMalloc after the first call)
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { orbx_scratch_free(p); }
    int alloc(size_t bytes) { return orbx_scratch_alloc(&p, bytes); }
    template <typename T> T* as() { return (T*)p; }
};

// ORBX_LAT_TRACE=1: device-side phase times of the stateless calls (events on the legacy stream), averaged over 50 calls, to stderr
struct PhaseTrace {
    static const int MAXP = 8;
    bool on; int np; cudaEvent_t ev[MAXP]; const char* tag[MAXP];
    explicit PhaseTrace(bool enable) : on(enable), np(0) {}
    void mark(const char* t, cudaStream_t st) { if (!on || np >= MAXP) return; cudaEventCreate(&ev[np]); cudaEventRecord(ev[np], st); tag[np++] = t; }
    void report(const char* name, double* acc, int& calls) {
        if (!on) return;
        for (int i = 1; i < np; ++i) { float ms = 0; cudaEventElapsedTime(&ms, ev[i - 1], ev[i]); acc[i] += ms * 1000.0; }
        for (int i = 0; i < np; ++i) cudaEventDestroy(ev[i]);
        if (++calls % 50 == 0) {
            fprintf(stderr, "[orbx lat] %s:", name);
            for (int i = 1; i < np; ++i) { fprintf(stderr, " %s %.1f us", tag[i], acc[i] / 50.0); acc[i] = 0; }
            fprintf(stderr, "\n");
        }
    }
};
inline bool orbx_lat_trace_on() { static const bool on = getenv("ORBX_LAT_TRACE") != nullptr; return on; }

// One stateless call = one device arena + one page-locked staging buffer laid out [inputs][in/out][outputs][work]: everything the
// call uploads goes up in ONE copy and everything it returns comes back in ONE copy (a dozen small synchronous copies of
// pageable memory cost 10-20 us each, more than the kernels of a per-frame call).  Segments are declared first (in / inout /
// out / work return an id), commit() lays them out, stages and uploads; dev<T>(id) is the segment on the device; finish()
// downloads, waits and hands the outputs back.  All work on the calling thread's own stream (orbx_call_stream).
struct CallArena {
    struct Seg { int kind; const void* src; void* dst; size_t bytes, off, room; };    // room: bytes laid out (>= bytes copied)
    std::vector<Seg> segs;
    DevBuf arena;
    PhaseTrace pt{orbx_lat_trace_on()};
    cudaStream_t st = orbx_call_stream();          // every copy and kernel of the call
    const char* name = "call";          // static storage: keys the ORBX_LAT_TRACE accumulators
    double* acc = nullptr; int* calls = nullptr;
    void trace(const char* n, double* a, int* c) { name = n; acc = a; calls = c; }
    u8* host = nullptr; u8* device = nullptr;
    size_t upBytes = 0, downBegin = 0, downEnd = 0;
    int add(int kind, const void* src, void* dst, size_t bytes, size_t room = 0) { segs.push_back({kind, src, dst, bytes, 0, std::max(bytes, room)}); return (int)segs.size() - 1; }
    int in(const void* src, size_t bytes) { return add(0, src, nullptr, bytes); }
    int inout(void* p, size_t bytes) { return add(1, p, p, bytes); }
    int inout_fill(void* dst, size_t bytes) { return add(1, nullptr, dst, bytes); }    // filled through hostp() before upload(); dst may be null (device-only state)
    int out(void* dst, size_t bytes, size_t room = 0) { return add(2, nullptr, dst, bytes, room); }    // room: what the kernels may write
    int work(size_t bytes) { return add(3, nullptr, nullptr, bytes); }
    int layout()
    {
        size_t off = 0;
        for (int kind = 0; kind < 4; ++kind) {
            if (kind == 1) downBegin = off;
            if (kind == 2) upBytes = off;
            if (kind == 3) downEnd = off;
            for (Seg& s : segs) if (s.kind == kind) { s.off = off; off += (s.room + 255) & ~(size_t)255; }
        }
        int rc = arena.alloc(off ? off : 256); if (rc) return rc;
        if ((rc = orbx_scratch_host((void**)&host, downEnd ? downEnd : 256))) return rc;
        device = arena.as<u8>();
        return ORBX_OK;
    }
    int upload()
    {
        for (const Seg& s : segs) if (s.kind <= 1 && s.bytes && s.src) memcpy(host + s.off, s.src, s.bytes);
        pt.mark("start", st);
        if (upBytes) ORBX_CUDA(cudaMemcpyAsync(device, host, upBytes, cudaMemcpyHostToDevice, st));
        pt.mark("h2d", st);
        return ORBX_OK;
    }
    int commit() { const int rc = layout(); return rc ? rc : upload(); }
    template <typename T> T* hostp(int id) { return reinterpret_cast<T*>(host + segs[id].off); }   // inputs without a source: fill between layout() and upload()
    template <typename T> T* dev(int id) { return reinterpret_cast<T*>(device + segs[id].off); }
    int finish()
    {
        ORBX_CUDA(cudaGetLastError());
        pt.mark("kernels", st);
        if (downEnd > downBegin) ORBX_CUDA(cudaMemcpyAsync(host + downBegin, device + downBegin, downEnd - downBegin, cudaMemcpyDeviceToHost, st));
        pt.mark("d2h", st);
        ORBX_CUDA(cudaStreamSynchronize(st));
        if (acc) pt.report(name, acc, *calls);
        for (const Seg& s : segs) if ((s.kind == 1 || s.kind == 2) && s.bytes && s.dst) memcpy(s.dst, host + s.off, s.bytes);
        return ORBX_OK;
    }
};

// temporary device frames of the stateless entry points
struct TempFrames {
    orbx_keypoint* kps = nullptr; u8* desc = nullptr; int* count = nullptr;
    unsigned short* gridStart = nullptr; unsigned short* gridIdx = nullptr;
    int cap = 0, nslots = 0;
    ~TempFrames() { orbx_scratch_free(kps); orbx_scratch_free(desc); orbx_scratch_free(count); orbx_scratch_free(gridStart); orbx_scratch_free(gridIdx); }
    int alloc(int slots, int capacity)
    {
        cap = std::max(capacity, 8); nslots = slots;
        int rc;
        if ((rc = orbx_scratch_alloc((void**)&kps, (size_t)slots * cap * sizeof(orbx_keypoint))) ||
            (rc = orbx_scratch_alloc((void**)&desc, (size_t)slots * cap * 32)) ||
            (rc = orbx_scratch_alloc((void**)&count, slots * sizeof(int))) ||
            (rc = orbx_scratch_alloc((void**)&gridStart, (size_t)slots * (ORBX_GRID_CELLS + 1) * sizeof(unsigned short))) ||
            (rc = orbx_scratch_alloc((void**)&gridIdx, (size_t)slots * cap * sizeof(unsigned short)))) return rc;
        return ORBX_OK;
    }
    int upload(int slot, const orbm_frame* f)
    {
        ORBX_CUDA(cudaMemcpy(count + slot, &f->n, sizeof(int), cudaMemcpyHostToDevice));
        if (f->n > 0) {
            ORBX_CUDA(cudaMemcpy(kps + (size_t)slot * cap, f->keys_un, (size_t)f->n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice));
            ORBX_CUDA(cudaMemcpy(desc + (size_t)slot * cap * 32, f->descriptors, (size_t)f->n * 32, cudaMemcpyHostToDevice));
        }
        return ORBX_OK;
    }
    FrameStore store() const
    {
        FrameStore fs; fs.kps = kps; fs.desc = desc; fs.count = count; fs.gridStart = gridStart; fs.gridIdx = gridIdx; fs.cap = cap;
        return fs;
    }
};


#endif
