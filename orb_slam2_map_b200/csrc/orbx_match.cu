// orbx_match.cu — sm_100a kernels for the Frame grid and the ORBmatcher Hamming searches.
//
//   K6  grid_build_kernel      Frame::AssignFeaturesToGrid / PosInGrid   (src/Frame.cc:230-245,382-392)
//   K7  search_init_kernel     ORBmatcher::SearchForInitialization       (src/ORBmatcher.cc:405-520)
//   K8  stereo_* kernels       Frame::ComputeStereoMatches               (src/Frame.cc:466-638)
//   K9  search_proj_kernel     ORBmatcher::SearchByProjection x4         (src/ORBmatcher.cc:45-129,290-403,1328-1599)
//   K10 knn2_kernel / K11 knn2_merge_kernel  brute-force 2-NN (BASELINE.json config 5)
//   hamming256 = ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) as 8x (XOR, POPC).
//
// The guided searches are sequential in the query index in the reference (a match changes what
// later queries may take).  One warp per frame (pair) replays that order; inside one query the
// candidates are spread over the lanes and "first minimum in GetFeaturesInArea order" becomes a
// min over (distance, cell, index) — the enumeration order of src/Frame.cc:350-376.

#include "orbx_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

typedef uint8_t u8;
#define FULL 0xffffffffu

__device__ __forceinline__ int hamming256(const uint32_t a[8], const uint32_t* __restrict__ b)
{
    const uint4 b0 = *reinterpret_cast<const uint4*>(b), b1 = *reinterpret_cast<const uint4*>(b + 4);
    return __popc(a[0] ^ b0.x) + __popc(a[1] ^ b0.y) + __popc(a[2] ^ b0.z) + __popc(a[3] ^ b0.w) +
           __popc(a[4] ^ b1.x) + __popc(a[5] ^ b1.y) + __popc(a[6] ^ b1.z) + __popc(a[7] ^ b1.w);
}

__device__ __forceinline__ void load_desc(uint32_t a[8], const u8* __restrict__ p)
{
    const uint4 v0 = *reinterpret_cast<const uint4*>(p), v1 = *reinterpret_cast<const uint4*>(p + 16);
    a[0] = v0.x; a[1] = v0.y; a[2] = v0.z; a[3] = v0.w; a[4] = v1.x; a[5] = v1.y; a[6] = v1.z; a[7] = v1.w;
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

// =============================================================================================
// K6: frame grid.  One warp per frame: count -> exclusive scan -> stable fill (ascending keypoint
// index inside a cell = the push_back order of AssignFeaturesToGrid).  Cell id = ix*48+iy.
// =============================================================================================
__global__ void __launch_bounds__(32)
grid_build_kernel(FrameStore fs, int firstSlot, GridParams gp)
{
    extern __shared__ unsigned short sm[];
    unsigned short* start = sm;                         // [CELLS]
    unsigned short* cellOf = sm + ORBX_GRID_CELLS;      // [cap]
    const int lane = threadIdx.x;
    const int slot = firstSlot + blockIdx.x;
    const int n = min(fs.count[slot], fs.cap);
    const orbx_keypoint* kps = fs.kps + (size_t)slot * fs.cap;
    for (int c = lane; c < ORBX_GRID_CELLS; c += 32) start[c] = 0;
    __syncwarp();
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        int cell = -1 - lane;    // distinct invalid values so invalid lanes never group
        if (i < n) {
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
            if (px >= 0 && px < ORBX_GRID_COLS && py >= 0 && py < ORBX_GRID_ROWS) cell = px * ORBX_GRID_ROWS + py;
            cellOf[i] = cell >= 0 ? (unsigned short)cell : (unsigned short)0xffff;
        }
        const unsigned m = __match_any_sync(FULL, cell);
        if (cell >= 0 && (__ffs(m) - 1) == lane) start[cell] += (unsigned short)__popc(m);
        __syncwarp();
    }
    // exclusive scan of the 3072 counts: 96 consecutive cells per lane
    const int per = ORBX_GRID_CELLS / 32;
    int sum = 0;
    for (int c = 0; c < per; ++c) sum += start[lane * per + c];
    int incl = sum;
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
    int run = incl - sum;
    unsigned short* gs = fs.gridStart + (size_t)slot * (ORBX_GRID_CELLS + 1);
    for (int c = 0; c < per; ++c) {
        const int cnt = start[lane * per + c];
        start[lane * per + c] = (unsigned short)run;
        gs[lane * per + c] = (unsigned short)run;
        run += cnt;
    }
    if (lane == 31) gs[ORBX_GRID_CELLS] = (unsigned short)run;
    __syncwarp();
    unsigned short* gi = fs.gridIdx + (size_t)slot * fs.cap;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        int cell = -1 - lane;
        if (i < n && cellOf[i] != 0xffff) cell = cellOf[i];
        const unsigned m = __match_any_sync(FULL, cell);
        int base = 0;
        if (cell >= 0) base = start[cell];
        __syncwarp();
        if (cell >= 0) {
            gi[base + __popc(m & ((1u << lane) - 1))] = (unsigned short)i;
            if ((__ffs(m) - 1) == lane) start[cell] = (unsigned short)(base + __popc(m));
        }
        __syncwarp();
    }
}

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
__device__ void three_maxima(const int* histo, int L, int& ind1, int& ind2, int& ind3)
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

// =============================================================================================
// K7: SearchForInitialization, one warp per frame pair (slotA -> slotB).
// =============================================================================================
__global__ void __launch_bounds__(32)
search_init_kernel(FrameStore fs, int firstSlot, int slotStride, GridParams gp, float* prevMatchedBase,
                   int windowSize, float nnratio, int checkOri, int* __restrict__ matches12Base,
                   int* __restrict__ nmatchesOut, unsigned long long* __restrict__ ncompares)
{
    extern __shared__ unsigned short sm[];
    const int lane = threadIdx.x;
    const int pair = blockIdx.x;
    const int sA = firstSlot + pair * slotStride, sB = sA + 1;
    const int n1 = min(fs.count[sA], fs.cap), n2 = min(fs.count[sB], fs.cap);
    unsigned short* md = sm;                   // vMatchedDistance (0xffff = INT_MAX)
    unsigned short* m21 = sm + fs.cap;         // vnMatches21 (0xffff = -1)
    u8* binOf = reinterpret_cast<u8*>(sm + 2 * fs.cap);
    int* hist = reinterpret_cast<int*>(binOf + ((fs.cap + 3) & ~3));
    for (int i = lane; i < n2; i += 32) { md[i] = 0xffff; m21[i] = 0xffff; }
    for (int i = lane; i < n1; i += 32) binOf[i] = 0xff;
    if (lane < ORBX_HISTO_LENGTH) hist[lane] = 0;
    int* m12 = matches12Base + (size_t)pair * fs.cap;
    for (int i = lane; i < n1; i += 32) m12[i] = -1;
    __syncwarp();
    const orbx_keypoint* k1 = fs.kps + (size_t)sA * fs.cap;
    const orbx_keypoint* k2 = fs.kps + (size_t)sB * fs.cap;
    const u8* d1 = fs.desc + (size_t)sA * fs.cap * 32;
    const u8* d2 = fs.desc + (size_t)sB * fs.cap * 32;
    const unsigned short* gs = fs.gridStart + (size_t)sB * (ORBX_GRID_CELLS + 1);
    const unsigned short* gi = fs.gridIdx + (size_t)sB * fs.cap;
    const float* prev = prevMatchedBase ? prevMatchedBase + (size_t)pair * fs.cap * 2 : nullptr;
    const float r = (float)windowSize;
    int nmatches = 0;
    unsigned long long ncmp = 0;
    for (int i1 = 0; i1 < n1; ++i1) {
        const orbx_keypoint kp1 = k1[i1];
        if (kp1.octave > 0) continue;
        const float x = prev ? prev[2 * i1] : kp1.x, y = prev ? prev[2 * i1 + 1] : kp1.y;
        int cx0, cx1, cy0, cy1;
        if (!area_cells(gp, x, y, r, cx0, cx1, cy0, cy1)) continue;
        uint32_t q[8];
        load_desc(q, d1 + (size_t)i1 * 32);
        const int ncy = cy1 - cy0 + 1, nc = (cx1 - cx0 + 1) * ncy;
        int best = 0x7fffffff, second = 0x7fffffff;
        unsigned key = 0xffffffffu;
        for (int c = lane; c < nc; c += 32) {
            const int ix = cx0 + c / ncy, iy = cy0 + c % ncy;
            const int cell = ix * ORBX_GRID_ROWS + iy;
            const int s = gs[cell], e = gs[cell + 1];
            for (int j = s; j < e; ++j) {
                const int idx = gi[j];
                const orbx_keypoint kp2 = k2[idx];
                if (kp2.octave > 0) continue;          // level window [0,0]: bCheckLevels, octave>maxLevel
                if (!(fabsf(__fsub_rn(kp2.x, x)) < r && fabsf(__fsub_rn(kp2.y, y)) < r)) continue;
                const int dist = hamming256(q, reinterpret_cast<const uint32_t*>(d2 + (size_t)idx * 32));
                ++ncmp;
                if ((int)md[idx] <= dist) continue;    // md = 0xffff never blocks
                if (dist < best) { second = best; best = dist; key = ((unsigned)cell << 16) | (unsigned)idx; }
                else if (dist < second) second = dist;
            }
        }
        warp_top2(best, key, second);
        if (best <= ORBX_TH_LOW && (float)best < __fmul_rn((float)second, nnratio)) {
            const int bi = (int)(key & 0xffff);
            const int old = m21[bi];
            __syncwarp();
            if (lane == 0) {
                if (old != 0xffff) m12[old] = -1;
                m12[i1] = bi;
                m21[bi] = (unsigned short)i1;
                md[bi] = (unsigned short)best;
                if (checkOri) { const int bin = rot_bin(kp1.angle, k2[bi].angle); binOf[i1] = (u8)bin; hist[bin] += 1; }
            }
            if (old != 0xffff) --nmatches;
            ++nmatches;
            __syncwarp();
        }
    }
    __syncwarp();
    if (checkOri) {
        int i1a, i2a, i3a;
        three_maxima(hist, ORBX_HISTO_LENGTH, i1a, i2a, i3a);
        __threadfence_block();
        for (int i0 = 0; i0 < n1; i0 += 32) {
            const int i = i0 + lane;
            bool cull = false;
            if (i < n1) {
                const int b = binOf[i];
                cull = b != 0xff && b != i1a && b != i2a && b != i3a && m12[i] >= 0;
                if (cull) m12[i] = -1;
            }
            nmatches -= __popc(__ballot_sync(FULL, cull));
        }
    }
    if (prevMatchedBase) {
        float* pw = prevMatchedBase + (size_t)pair * fs.cap * 2;
        for (int i = lane; i < n1; i += 32) {
            const int m = m12[i];
            if (m >= 0) { pw[2 * i] = k2[m].x; pw[2 * i + 1] = k2[m].y; }
        }
    }
    for (int d = 16; d > 0; d >>= 1) ncmp += __shfl_xor_sync(FULL, ncmp, d);
    if (lane == 0) {
        nmatchesOut[pair] = nmatches;
        if (ncompares) atomicAdd(ncompares, ncmp);
    }
}

__global__ void hamming_pairs_kernel(const u8* __restrict__ a, const u8* __restrict__ b, int n, int* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[8];
    load_desc(q, a + (size_t)i * 32);
    out[i] = hamming256(q, reinterpret_cast<const uint32_t*>(b + (size_t)i * 32));
}

// =============================================================================================
// host side
// =============================================================================================
static int ensure_match_buffers(orbx_handle* h)
{
    if (h->dGridStart) return ORBX_OK;
    const size_t R = (size_t)h->resultSlots;
    ORBX_CUDA(cudaMalloc((void**)&h->dGridStart, R * (ORBX_GRID_CELLS + 1) * sizeof(unsigned short)));
    ORBX_CUDA(cudaMalloc((void**)&h->dGridIdx, R * h->kpCap * sizeof(unsigned short)));
    ORBX_CUDA(cudaMalloc((void**)&h->dMatches12, R * h->kpCap * sizeof(int)));
    ORBX_CUDA(cudaMalloc((void**)&h->dNMatches, R * sizeof(int)));
    ORBX_CUDA(cudaMalloc((void**)&h->dNCompares, sizeof(unsigned long long)));
    ORBX_CUDA(cudaMalloc((void**)&h->dURight, R * h->kpCap * sizeof(float)));
    ORBX_CUDA(cudaMalloc((void**)&h->dDepth, R * h->kpCap * sizeof(float)));
    return ORBX_OK;
}

static FrameStore store_of(orbx_handle* h)
{
    FrameStore fs;
    fs.kps = h->dKps; fs.desc = h->dDesc; fs.count = h->dCount;
    fs.gridStart = h->dGridStart; fs.gridIdx = h->dGridIdx; fs.cap = h->kpCap;
    return fs;
}

static GridParams grid_params(float minX, float maxX, float minY, float maxY)
{
    GridParams gp;
    gp.minX = minX; gp.minY = minY;
    gp.invW = static_cast<float>(ORBX_GRID_COLS) / (maxX - minX);   // src/Frame.cc:101-102
    gp.invH = static_cast<float>(ORBX_GRID_ROWS) / (maxY - minY);
    return gp;
}

static int launch_grid_build(const FrameStore& fs, int firstSlot, int nframes, const GridParams& gp, cudaStream_t st)
{
    const int smem = (ORBX_GRID_CELLS + fs.cap) * (int)sizeof(unsigned short);
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    grid_build_kernel<<<nframes, 32, smem, st>>>(fs, firstSlot, gp);
    return ORBX_OK;
}

static int launch_search_init(const FrameStore& fs, int firstSlot, int slotStride, int npairs, const GridParams& gp,
                              float* prev, int window, float nnratio, int checkOri, int* m12, int* nm,
                              unsigned long long* ncmp, cudaStream_t st)
{
    const int smem = fs.cap * 4 + ((fs.cap + 3) & ~3) + ORBX_HISTO_LENGTH * 4 + 16;
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(search_init_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    search_init_kernel<<<npairs, 32, smem, st>>>(fs, firstSlot, slotStride, gp, prev, window, nnratio, checkOri, m12, nm, ncmp);
    return ORBX_OK;
}

int orbx_build_grids(orbx_handle* h, int firstSlot, int nframes, int width, int height)
{
    int rc = ensure_match_buffers(h);
    if (rc) return rc;
    return launch_grid_build(store_of(h), firstSlot, nframes, grid_params(0.f, (float)width, 0.f, (float)height), h->stream);
}

extern "C" {

int orbx_search_init_pairs(orbx_handle* h, int first_slot, int npairs, int window_size, float nnratio,
                           int check_orientation, int32_t* matches12, int32_t* nmatches, int capacity,
                           uint64_t* ncompares)
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (npairs <= 0) return ORBX_OK;
    if (first_slot < 0 || first_slot + npairs + 1 > h->resultSlots) { orbx_set_error("pair range exceeds result store"); return ORBX_E_INVALID; }
    if (h->curW <= 0) { orbx_set_error("no frames extracted yet"); return ORBX_E_STATE; }
    if (matches12 && capacity < h->kpCap) { orbx_set_error("matches12 capacity %d < %d", capacity, h->kpCap); return ORBX_E_CAPACITY; }
    ORBX_CUDA(cudaSetDevice(h->device));
    int rc = ensure_match_buffers(h);
    if (rc) return rc;
    const FrameStore fs = store_of(h);
    const GridParams gp = grid_params(0.f, (float)h->curW, 0.f, (float)h->curH);
    orbx_prof_mark(h, -1);
    rc = launch_grid_build(fs, first_slot, npairs + 1, gp, h->stream);
    if (rc) return rc;
    ORBX_COUNT(h, ORBX_STAGE_GRID, 1);
    orbx_prof_mark(h, ORBX_STAGE_GRID);
    if (ncompares) ORBX_CUDA(cudaMemsetAsync(h->dNCompares, 0, sizeof(unsigned long long), h->stream));
    rc = launch_search_init(fs, first_slot, 1, npairs, gp, nullptr, window_size, nnratio, check_orientation,
                            h->dMatches12 + (size_t)first_slot * h->kpCap, h->dNMatches + first_slot,
                            ncompares ? h->dNCompares : nullptr, h->stream);
    if (rc) return rc;
    ORBX_COUNT(h, ORBX_STAGE_SEARCH_INIT, 1);
    orbx_prof_mark(h, ORBX_STAGE_SEARCH_INIT);
    ORBX_CUDA(cudaGetLastError());
    if (matches12)
        ORBX_CUDA(cudaMemcpy2DAsync(matches12, (size_t)capacity * 4, h->dMatches12 + (size_t)first_slot * h->kpCap, (size_t)h->kpCap * 4,
                                    (size_t)h->kpCap * 4, npairs, cudaMemcpyDeviceToHost, h->stream));
    if (nmatches) ORBX_CUDA(cudaMemcpyAsync(nmatches, h->dNMatches + first_slot, npairs * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    if (ncompares) ORBX_CUDA(cudaMemcpyAsync(ncompares, h->dNCompares, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    if (matches12 || nmatches || ncompares) return orbx_sync(h);
    return ORBX_OK;
}

// ---- stateless entry points: temporary device frames -----------------------------------------
struct TempFrames {
    orbx_keypoint* kps = nullptr; u8* desc = nullptr; int* count = nullptr;
    unsigned short* gridStart = nullptr; unsigned short* gridIdx = nullptr;
    int cap = 0, nslots = 0;
    ~TempFrames() { cudaFree(kps); cudaFree(desc); cudaFree(count); cudaFree(gridStart); cudaFree(gridIdx); }
    int alloc(int slots, int capacity)
    {
        cap = std::max(capacity, 8); nslots = slots;
        ORBX_CUDA(cudaMalloc((void**)&kps, (size_t)slots * cap * sizeof(orbx_keypoint)));
        ORBX_CUDA(cudaMalloc((void**)&desc, (size_t)slots * cap * 32));
        ORBX_CUDA(cudaMalloc((void**)&count, slots * sizeof(int)));
        ORBX_CUDA(cudaMalloc((void**)&gridStart, (size_t)slots * (ORBX_GRID_CELLS + 1) * sizeof(unsigned short)));
        ORBX_CUDA(cudaMalloc((void**)&gridIdx, (size_t)slots * cap * sizeof(unsigned short)));
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

static int check_frame(const orbm_frame* f)
{
    if (!f || f->n < 0 || (f->n > 0 && (!f->keys_un || !f->descriptors))) { orbx_set_error("bad orbm_frame"); return ORBX_E_INVALID; }
    if (f->n > 65000) { orbx_set_error("more than 65000 keypoints per frame unsupported"); return ORBX_E_CAPACITY; }
    if (!(f->max_x > f->min_x) || !(f->max_y > f->min_y)) { orbx_set_error("bad frame bounds"); return ORBX_E_INVALID; }
    return ORBX_OK;
}

static int require_device()
{
    int n = 0;
    ORBX_CUDA(cudaGetDeviceCount(&n));
    if (n <= 0) { orbx_set_error("no CUDA device; there is no CPU fallback"); return ORBX_E_CUDA; }
    return ORBX_OK;
}

int orbm_descriptor_distance(const uint8_t* a, const uint8_t* b, int count, int32_t* dist)
{
    if (count <= 0) return ORBX_OK;
    if (!a || !b || !dist) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    int rc = require_device();
    if (rc) return rc;
    u8 *da = nullptr, *db = nullptr; int* dd = nullptr;
    ORBX_CUDA(cudaMalloc((void**)&da, (size_t)count * 32));
    ORBX_CUDA(cudaMalloc((void**)&db, (size_t)count * 32));
    ORBX_CUDA(cudaMalloc((void**)&dd, (size_t)count * 4));
    ORBX_CUDA(cudaMemcpy(da, a, (size_t)count * 32, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(db, b, (size_t)count * 32, cudaMemcpyHostToDevice));
    hamming_pairs_kernel<<<(count + 255) / 256, 256>>>(da, db, count, dd);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpy(dist, dd, (size_t)count * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(db); cudaFree(dd);
    return ORBX_OK;
}

int orbm_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_matched,
                                   int32_t* matches12, int window_size, float nnratio,
                                   int check_orientation, int32_t* nmatches)
{
    int rc = check_frame(f1); if (rc) return rc;
    rc = check_frame(f2); if (rc) return rc;
    if (!prev_matched || !matches12) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    rc = require_device(); if (rc) return rc;
    TempFrames tf;
    rc = tf.alloc(2, std::max(f1->n, f2->n)); if (rc) return rc;
    rc = tf.upload(0, f1); if (rc) return rc;
    rc = tf.upload(1, f2); if (rc) return rc;
    float* dPrev = nullptr; int* dM12 = nullptr; int* dNm = nullptr;
    ORBX_CUDA(cudaMalloc((void**)&dPrev, (size_t)tf.cap * 2 * sizeof(float)));
    ORBX_CUDA(cudaMalloc((void**)&dM12, (size_t)tf.cap * sizeof(int)));
    ORBX_CUDA(cudaMalloc((void**)&dNm, sizeof(int)));
    if (f1->n) ORBX_CUDA(cudaMemcpy(dPrev, prev_matched, (size_t)f1->n * 2 * sizeof(float), cudaMemcpyHostToDevice));
    const GridParams gp = grid_params(f2->min_x, f2->max_x, f2->min_y, f2->max_y);
    rc = launch_grid_build(tf.store(), 1, 1, gp, 0); if (rc) return rc;
    rc = launch_search_init(tf.store(), 0, 1, 1, gp, dPrev, window_size, nnratio, check_orientation, dM12, dNm, nullptr, 0);
    if (rc) return rc;
    ORBX_CUDA(cudaGetLastError());
    int nm = 0;
    ORBX_CUDA(cudaMemcpy(&nm, dNm, sizeof(int), cudaMemcpyDeviceToHost));
    if (f1->n) {
        ORBX_CUDA(cudaMemcpy(matches12, dM12, (size_t)f1->n * sizeof(int), cudaMemcpyDeviceToHost));
        ORBX_CUDA(cudaMemcpy(prev_matched, dPrev, (size_t)f1->n * 2 * sizeof(float), cudaMemcpyDeviceToHost));
    }
    if (nmatches) *nmatches = nm;
    cudaFree(dPrev); cudaFree(dM12); cudaFree(dNm);
    return ORBX_OK;
}

#define STUB { orbx_set_error("not implemented yet"); return ORBX_E_STATE; }
int orbx_extract_stereo_batch(orbx_handle*, const uint8_t*, const uint8_t*, size_t, size_t, int, int, int, int, int, float, float, float*, float*, int, uint64_t*) STUB
int orbm_search_by_projection(int, const orbm_frame*, const orbm_proj_query*, const uint8_t*, int, const uint8_t*, uint8_t*, int32_t*, int32_t*, int32_t*, int, float, int, int32_t*) STUB
int orbm_compute_stereo_matches(const orbm_stereo_input*, float*, float*, uint64_t*) STUB
int orbm_knn2_device(const uint8_t*, int64_t, const uint8_t*, int64_t, int32_t, uint16_t*, int32_t*, void*) STUB
int orbm_knn2_merge_device(const uint16_t*, const int32_t*, int, int64_t, uint16_t*, int32_t*, void*) STUB

} // extern "C"
