// orbx_match_common.cuh — device helpers shared by the matcher kernels (orbx_match.cu,
// orbx_stereo.cu, orbx_knn.cu).  Not part of the public ABI.
#ifndef ORBX_MATCH_COMMON_CUH
#define ORBX_MATCH_COMMON_CUH

#include "orbx_internal.h"

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
