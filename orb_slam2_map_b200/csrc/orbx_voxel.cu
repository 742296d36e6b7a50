// orbx_voxel.cu — the voxel-grid down-sampling of this fork's dense map (second half):
// `voxel.setLeafSize(resolution ...); voxel.setInputCloud(cloud); voxel.filter(*tmp)` at src/PointCloudMap.cc:41,
// 240-241, 259-260, 277-280, i.e. pcl::VoxelGrid<pcl::PointXYZRGBA>::applyFilter with its default parameters.
//
// PCL (find_package(PCL 1.7), CMakeLists.txt:39) is not vendored under /root/reference: the published algorithm is restated
// (oracle/frame_oracle.cpp: oracle_voxel_grid) — bounds of the finite points, min_b = floor(min * 1/leaf), voxel index
// idx = i + j*dx + k*dx*dy, one output point per occupied voxel in ascending idx order = the centroid of its points.
// PCL sums a voxel's points in the order an UNSTABLE std::sort leaves them, so its float centroids are not bit-defined;
// oracle and kernels fix the order to the input order (one legitimate execution).  Colours are integer sums: exact.
//
//   K24 voxel_bounds_kernel   min / max of the finite points (per-block partials, finished on the host: the grid extent
//                             sizes the occupancy bitmap)
//   K25 voxel_mark_kernel     voxel index per point, occupancy bitmap (atomicOr)
//   K26 scan kernels          exclusive prefix sums (bitmap popcounts -> rank of every occupied voxel = its output slot;
//                             points per voxel -> member segments)
//   K27 voxel_count / voxel_fill / voxel_reduce   members of every voxel, put back into input order (heap sort inside the
//                             segment), summed sequentially by one thread per voxel
// HBM-bound integer / float work; nothing here is a contraction.

#include "orbx_internal.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <vector>

#define VX_THREADS 256
#define VX_ITEMS 16                     // elements per thread in the scan kernels

__global__ void __launch_bounds__(VX_THREADS)
voxel_bounds_kernel(const orbx_cloud_point* __restrict__ pts, int n, float* __restrict__ partial /* [blocks][6] */)
{
    __shared__ float red[VX_THREADS / 32][6];
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * VX_THREADS + threadIdx.x; i < n; i += gridDim.x * VX_THREADS) {
        const orbx_cloud_point p = pts[i];
        if (!isfinite(p.x) || !isfinite(p.y) || !isfinite(p.z)) continue;
        mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
        mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], d));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], d));
        }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) for (int k = 0; k < 3; ++k) { red[wid][k] = mn[k]; red[wid][3 + k] = mx[k]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = red[0][threadIdx.x];
        for (int w = 1; w < VX_THREADS / 32; ++w) v = threadIdx.x < 3 ? fminf(v, red[w][threadIdx.x]) : fmaxf(v, red[w][threadIdx.x]);
        partial[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

struct VoxelGeom { float inv; int minb[3]; int mul1, mul2; };

__device__ __forceinline__ bool voxel_index(const VoxelGeom& g, const orbx_cloud_point& p, uint32_t& idx)
{
    if (!isfinite(p.x) || !isfinite(p.y) || !isfinite(p.z)) return false;
    // static_cast<int>(floor(pt.x * inverse_leaf_size[0]) - static_cast<float>(min_b[0]))
    const int i = (int)__fsub_rn(floorf(__fmul_rn(p.x, g.inv)), (float)g.minb[0]);
    const int j = (int)__fsub_rn(floorf(__fmul_rn(p.y, g.inv)), (float)g.minb[1]);
    const int k = (int)__fsub_rn(floorf(__fmul_rn(p.z, g.inv)), (float)g.minb[2]);
    idx = (uint32_t)(i + j * g.mul1 + k * g.mul2);
    return true;
}

__global__ void __launch_bounds__(VX_THREADS)
voxel_mark_kernel(const orbx_cloud_point* __restrict__ pts, int n, VoxelGeom g, uint32_t* __restrict__ idxOut, uint32_t* __restrict__ bitmap)
{
    const int i = blockIdx.x * VX_THREADS + threadIdx.x;
    if (i >= n) return;
    uint32_t idx;
    if (voxel_index(g, pts[i], idx)) { idxOut[i] = idx; atomicOr(&bitmap[idx >> 5], 1u << (idx & 31)); }
    else idxOut[i] = 0xffffffffu;
}

// ---- exclusive prefix sum over n values (POPC = 1: the value of element i is popc(in[i])) in three launches --------------
template <int POPC>
__device__ __forceinline__ uint32_t scan_val(const uint32_t* in, long long i, long long n) { return i < n ? (POPC ? (uint32_t)__popc(in[i]) : in[i]) : 0u; }

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* wsum, uint32_t& total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += u; }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < VX_THREADS / 32; ++w) { const uint32_t s = wsum[w]; if (w < wid) off += s; tot += s; }
    __syncthreads();
    total = tot;
    return off + incl - v;
}

template <int POPC>
__global__ void __launch_bounds__(VX_THREADS)
scan_sums_kernel(const uint32_t* __restrict__ in, long long n, uint32_t* __restrict__ blockSums)
{
    __shared__ uint32_t wsum[VX_THREADS / 32];
    const long long base = ((long long)blockIdx.x * VX_THREADS + threadIdx.x) * VX_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < VX_ITEMS; ++k) s += scan_val<POPC>(in, base + k, n);
    uint32_t total;
    block_excl_scan(s, wsum, total);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}

// one block: exclusive scan of the block sums in place; sums[nb] = grand total
__global__ void __launch_bounds__(VX_THREADS)
scan_top_kernel(uint32_t* __restrict__ sums, int nb)
{
    __shared__ uint32_t wsum[VX_THREADS / 32];
    uint32_t carry = 0;
    for (int b0 = 0; b0 < nb; b0 += VX_THREADS) {
        const int b = b0 + threadIdx.x;
        const uint32_t v = b < nb ? sums[b] : 0u;
        uint32_t total;
        const uint32_t ex = block_excl_scan(v, wsum, total);
        if (b < nb) sums[b] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[nb] = carry;
}

template <int POPC>
__global__ void __launch_bounds__(VX_THREADS)
scan_apply_kernel(const uint32_t* __restrict__ in, long long n, const uint32_t* __restrict__ blockSums, uint32_t* __restrict__ out)
{
    __shared__ uint32_t wsum[VX_THREADS / 32];
    const long long base = ((long long)blockIdx.x * VX_THREADS + threadIdx.x) * VX_ITEMS;
    uint32_t v[VX_ITEMS], s = 0;
#pragma unroll
    for (int k = 0; k < VX_ITEMS; ++k) { v[k] = scan_val<POPC>(in, base + k, n); s += v[k]; }
    uint32_t total;
    uint32_t run = blockSums[blockIdx.x] + block_excl_scan(s, wsum, total);
#pragma unroll
    for (int k = 0; k < VX_ITEMS; ++k) { if (base + k < n) out[base + k] = run; run += v[k]; }
}

// rank of an occupied voxel = occupied voxels with a smaller index = its output slot (PCL emits ascending idx)
__device__ __forceinline__ uint32_t voxel_rank(const uint32_t* __restrict__ bitmap, const uint32_t* __restrict__ wordRank, uint32_t idx)
{
    return wordRank[idx >> 5] + (uint32_t)__popc(bitmap[idx >> 5] & ((1u << (idx & 31)) - 1u));
}

__global__ void __launch_bounds__(VX_THREADS)
voxel_count_kernel(const uint32_t* __restrict__ idx, int n, const uint32_t* __restrict__ bitmap, const uint32_t* __restrict__ wordRank,
                   uint32_t* __restrict__ rankOut, uint32_t* __restrict__ counts)
{
    const int i = blockIdx.x * VX_THREADS + threadIdx.x;
    if (i >= n) return;
    const uint32_t v = idx[i];
    if (v == 0xffffffffu) { rankOut[i] = 0xffffffffu; return; }
    const uint32_t r = voxel_rank(bitmap, wordRank, v);
    rankOut[i] = r;
    atomicAdd(&counts[r], 1u);
}

__global__ void __launch_bounds__(VX_THREADS)
voxel_fill_kernel(const uint32_t* __restrict__ rank, int n, const uint32_t* __restrict__ segStart, uint32_t* __restrict__ cursor,
                  uint32_t* __restrict__ members)
{
    const int i = blockIdx.x * VX_THREADS + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = rank[i];
    if (r == 0xffffffffu) return;
    members[segStart[r] + atomicAdd(&cursor[r], 1u)] = (uint32_t)i;
}

// one thread per occupied voxel: its members back into input order (heap sort in place: the atomics above filled the segment in
// arbitrary order), then the centroid with the sums taken in that order
__global__ void __launch_bounds__(VX_THREADS)
voxel_reduce_kernel(const orbx_cloud_point* __restrict__ pts, const uint32_t* __restrict__ segStart, int nvox, uint32_t* __restrict__ members,
                    orbx_cloud_point* __restrict__ out)
{
    const int v = blockIdx.x * VX_THREADS + threadIdx.x;
    if (v >= nvox) return;
    uint32_t* m = members + segStart[v];
    const int cnt = (int)(segStart[v + 1] - segStart[v]);
    auto sift = [&](int root, int end) {
        for (;;) {
            int child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && m[child] < m[child + 1]) ++child;
            if (m[root] >= m[child]) break;
            const uint32_t t = m[root]; m[root] = m[child]; m[child] = t;
            root = child;
        }
    };
    if (cnt > 1) {
        bool sorted = true;
        for (int i = 1; i < cnt; ++i) if (m[i - 1] > m[i]) { sorted = false; break; }
        if (!sorted) {
            for (int i = cnt / 2 - 1; i >= 0; --i) sift(i, cnt);
            for (int e = cnt - 1; e > 0; --e) { const uint32_t t = m[0]; m[0] = m[e]; m[e] = t; sift(0, e); }
        }
    }
    float sx = 0.f, sy = 0.f, sz = 0.f;
    uint32_t sr = 0, sg = 0, sb = 0;
    for (int i = 0; i < cnt; ++i) {
        const orbx_cloud_point p = pts[m[i]];
        sx = __fadd_rn(sx, p.x); sy = __fadd_rn(sy, p.y); sz = __fadd_rn(sz, p.z);
        sr += p.r; sg += p.g; sb += p.b;
    }
    const float fc = (float)cnt;
    orbx_cloud_point o;
    o.x = __fdiv_rn(sx, fc); o.y = __fdiv_rn(sy, fc); o.z = __fdiv_rn(sz, fc);
    o.r = (uint8_t)(int)__fdiv_rn((float)sr, fc); o.g = (uint8_t)(int)__fdiv_rn((float)sg, fc); o.b = (uint8_t)(int)__fdiv_rn((float)sb, fc);
    o.a = 255;
    out[v] = o;
}

// exclusive scan of n values into out (n entries) ; *totalOut (device) = sums[nb] afterwards
template <int POPC>
static int device_scan(const uint32_t* in, long long n, uint32_t* out, uint32_t* sums /* nb + 1 */, int nb, cudaStream_t st)
{
    scan_sums_kernel<POPC><<<nb, VX_THREADS, 0, st>>>(in, n, sums);
    scan_top_kernel<<<1, VX_THREADS, 0, st>>>(sums, nb);
    scan_apply_kernel<POPC><<<nb, VX_THREADS, 0, st>>>(in, n, sums, out);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

extern "C" {

int orbm_voxel_grid(const orbx_cloud_point* points, int n, float leaf, int device, orbx_cloud_point* out, int capacity, int32_t* nout)
{
    ORBX_NVTX("orbm_voxel_grid");
    if (!nout || (n > 0 && !points)) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    *nout = 0;
    if (n <= 0) return ORBX_OK;
    if (!(leaf > 0.f)) { orbx_set_error("leaf size must be positive"); return ORBX_E_INVALID; }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (ndev <= 0 || device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available; there is no CPU fallback", device); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    cudaStream_t st = 0;
    struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } } dPts, dPart, dIdx, dBitmap, dWordRank, dSums, dRank, dCounts, dSeg, dCursor, dMembers, dOut;
    ORBX_CUDA(cudaMalloc(&dPts.p, (size_t)n * sizeof(orbx_cloud_point)));
    ORBX_CUDA(cudaMemcpyAsync(dPts.p, points, (size_t)n * sizeof(orbx_cloud_point), cudaMemcpyHostToDevice, st));
    // ---- bounds of the finite points (getMinMax3D) ----------------------------------------------------------------------
    const int pb = std::min(1024, (n + VX_THREADS - 1) / VX_THREADS);
    ORBX_CUDA(cudaMalloc(&dPart.p, (size_t)pb * 6 * sizeof(float)));
    voxel_bounds_kernel<<<pb, VX_THREADS, 0, st>>>((const orbx_cloud_point*)dPts.p, n, (float*)dPart.p);
    std::vector<float> part((size_t)pb * 6);
    ORBX_CUDA(cudaMemcpyAsync(part.data(), dPart.p, part.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int b = 0; b < pb; ++b) for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], part[b * 6 + k]); mx[k] = std::max(mx[k], part[b * 6 + 3 + k]); }
    if (mn[0] > mx[0]) return ORBX_OK;                                   // no finite point
    VoxelGeom g;
    g.inv = 1.0f / leaf;
    long long div[3];
    for (int k = 0; k < 3; ++k) {
        // PCL refuses grids whose index would overflow an int: (max - min) * inverse_leaf + 1 per axis, product > INT_MAX
        const long long d = (long long)((mx[k] - mn[k]) * g.inv) + 1;
        g.minb[k] = (int)floorf(mn[k] * g.inv);
        div[k] = (long long)(int)floorf(mx[k] * g.inv) - g.minb[k] + 1;
        if (d > 0x7fffffffLL || div[k] > 0x7fffffffLL) { orbx_set_error("leaf size too small for the extent of the cloud (index overflow)"); return ORBX_E_CAPACITY; }
    }
    const long long dcheck = ((long long)((mx[0] - mn[0]) * g.inv) + 1) * ((long long)((mx[1] - mn[1]) * g.inv) + 1);
    if (dcheck > 0x7fffffffLL || dcheck * ((long long)((mx[2] - mn[2]) * g.inv) + 1) > 0x7fffffffLL || div[0] * div[1] * div[2] > 0x7fffffffLL) {
        orbx_set_error("leaf size too small for the extent of the cloud (index overflow): PCL returns the cloud unfiltered here");
        return ORBX_E_CAPACITY;
    }
    g.mul1 = (int)div[0]; g.mul2 = (int)(div[0] * div[1]);
    const long long nvoxAll = div[0] * div[1] * div[2], words = (nvoxAll + 31) / 32;
    // ---- occupancy and ranks -------------------------------------------------------------------------------------------------
    const int nbPts = (n + VX_THREADS - 1) / VX_THREADS;
    const int nbWords = (int)((words + (long long)VX_THREADS * VX_ITEMS - 1) / ((long long)VX_THREADS * VX_ITEMS));
    ORBX_CUDA(cudaMalloc(&dIdx.p, (size_t)n * 4));
    ORBX_CUDA(cudaMalloc(&dBitmap.p, (size_t)words * 4));
    ORBX_CUDA(cudaMalloc(&dWordRank.p, (size_t)words * 4));
    ORBX_CUDA(cudaMalloc(&dSums.p, ((size_t)std::max(nbWords, (n + VX_THREADS * VX_ITEMS - 1) / (VX_THREADS * VX_ITEMS)) + 1) * 4));
    ORBX_CUDA(cudaMemsetAsync(dBitmap.p, 0, (size_t)words * 4, st));
    voxel_mark_kernel<<<nbPts, VX_THREADS, 0, st>>>((const orbx_cloud_point*)dPts.p, n, g, (uint32_t*)dIdx.p, (uint32_t*)dBitmap.p);
    int rc = device_scan<1>((const uint32_t*)dBitmap.p, words, (uint32_t*)dWordRank.p, (uint32_t*)dSums.p, nbWords, st);
    if (rc) return rc;
    uint32_t nocc = 0;
    ORBX_CUDA(cudaMemcpyAsync(&nocc, (uint32_t*)dSums.p + nbWords, 4, cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    *nout = (int32_t)nocc;
    if (!out) return ORBX_OK;                                            // count only
    if ((int)nocc > capacity) { orbx_set_error("capacity %d < %u voxels", capacity, nocc); return ORBX_E_CAPACITY; }
    // ---- members of every voxel, centroids --------------------------------------------------------------------------------
    ORBX_CUDA(cudaMalloc(&dRank.p, (size_t)n * 4));
    ORBX_CUDA(cudaMalloc(&dCounts.p, ((size_t)nocc + 1) * 4));
    ORBX_CUDA(cudaMalloc(&dSeg.p, ((size_t)nocc + 1) * 4));
    ORBX_CUDA(cudaMalloc(&dCursor.p, ((size_t)nocc + 1) * 4));
    ORBX_CUDA(cudaMalloc(&dMembers.p, (size_t)n * 4));
    ORBX_CUDA(cudaMalloc(&dOut.p, (size_t)nocc * sizeof(orbx_cloud_point)));
    ORBX_CUDA(cudaMemsetAsync(dCounts.p, 0, ((size_t)nocc + 1) * 4, st));
    ORBX_CUDA(cudaMemsetAsync(dCursor.p, 0, ((size_t)nocc + 1) * 4, st));
    voxel_count_kernel<<<nbPts, VX_THREADS, 0, st>>>((const uint32_t*)dIdx.p, n, (const uint32_t*)dBitmap.p, (const uint32_t*)dWordRank.p,
                                                     (uint32_t*)dRank.p, (uint32_t*)dCounts.p);
    const long long nseg = (long long)nocc + 1;
    const int nbSeg = (int)((nseg + VX_THREADS * VX_ITEMS - 1) / (VX_THREADS * VX_ITEMS));
    rc = device_scan<0>((const uint32_t*)dCounts.p, nseg, (uint32_t*)dSeg.p, (uint32_t*)dSums.p, nbSeg, st);
    if (rc) return rc;
    voxel_fill_kernel<<<nbPts, VX_THREADS, 0, st>>>((const uint32_t*)dRank.p, n, (const uint32_t*)dSeg.p, (uint32_t*)dCursor.p, (uint32_t*)dMembers.p);
    voxel_reduce_kernel<<<((int)nocc + VX_THREADS - 1) / VX_THREADS, VX_THREADS, 0, st>>>((const orbx_cloud_point*)dPts.p, (const uint32_t*)dSeg.p, (int)nocc,
                                                                                          (uint32_t*)dMembers.p, (orbx_cloud_point*)dOut.p);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(out, dOut.p, (size_t)nocc * sizeof(orbx_cloud_point), cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

} // extern "C"
