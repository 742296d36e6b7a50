// orbx_sor.cu — pcl::StatisticalOutlierRemoval<PointXYZRGBA> as PointCloudMapping applies it to the global map before saving it
// (reference src/PointCloudMap.cc:46-47 setMeanK(50) / setStddevMulThresh(1.0), :284-285 sor.filter), SURVEY.md §8(f) rank 4.
//
//   K28 sor_tile_bounds_kernel   bounding box of every tile of SOR_TILE consecutive points
//   K29 sor_mean_dist_kernel     EXACT k-nearest-neighbour search, thread per point: the mean_k + 1 smallest squared distances
//                                (the point itself included, as the kd-tree search returns it) in a sorted per-thread list;
//                                candidate tiles are visited outward from the thread block's own tile and a tile is skipped when
//                                its box is farther from the block's box than the block's current worst neighbour (clouds that
//                                come out of the voxel filter are spatially coherent in index order, so most tiles are skipped;
//                                a shuffled cloud degrades to the full n^2 scan, still exact)
//   K30 sor_partial_sums / sor_threshold_kernel   mean and standard deviation of the mean distances (double, fixed tree order)
//   K31 sor_flag_kernel + scan + sor_compact_kernel   keep points with distance <= mean + mul * stddev, input order preserved
//
// PCL is not in the reference tree (an external dependency), so the semantics restate PCL's published algorithm
// (filters/impl/statistical_outlier_removal.hpp): squared distances as FLANN's L2_Simple accumulates them in float
// ((dx*dx + dy*dy) + dz*dz, no FMA), dist_sum += sqrt(d2) in double over the neighbours in ascending order, mean distance
// = (float)(dist_sum / mean_k); non-finite points get distance 0 and are kept; fewer than mean_k + 1 points: everything is kept.

#include "orbx_internal.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <vector>

#define SOR_TILE 256
#define SOR_KMAX 128                 // mean_k + 1 <= SOR_KMAX

namespace {

__device__ __forceinline__ bool finite3(float x, float y, float z) { return isfinite(x) && isfinite(y) && isfinite(z); }

__global__ void __launch_bounds__(SOR_TILE)
sor_tile_bounds_kernel(const orbx_cloud_point* __restrict__ pts, long long n, float* __restrict__ box /* [tiles][6] */)
{
    __shared__ float red[6][SOR_TILE / 32];
    const long long i = (long long)blockIdx.x * SOR_TILE + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n) {
        const orbx_cloud_point p = pts[i];
        if (finite3(p.x, p.y, p.z)) { lo[0] = hi[0] = p.x; lo[1] = hi[1] = p.y; lo[2] = hi[2] = p.z; }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int d = 16; d > 0; d >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], d));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], d));
        }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) for (int k = 0; k < 3; ++k) { red[k][wid] = lo[k]; red[3 + k][wid] = hi[k]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = red[threadIdx.x][0];
        for (int w = 1; w < SOR_TILE / 32; ++w) v = threadIdx.x < 3 ? fminf(v, red[threadIdx.x][w]) : fmaxf(v, red[threadIdx.x][w]);
        box[(size_t)blockIdx.x * 6 + threadIdx.x] = v;
    }
}

// squared distance between two boxes (0 when they overlap); an empty box (lo > hi) is infinitely far
__device__ __forceinline__ float box_dist2(const float* a, const float* b)
{
    if (a[0] > a[3] || b[0] > b[3]) return FLT_MAX;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float g = fmaxf(0.f, fmaxf(a[k] - b[3 + k], b[k] - a[3 + k]));
        s += g * g;
    }
    return s;
}

__global__ void __launch_bounds__(SOR_TILE)
sor_mean_dist_kernel(const orbx_cloud_point* __restrict__ pts, long long n, const float* __restrict__ box, int ntiles, int K,
                     float* __restrict__ meanDist, unsigned char* __restrict__ valid, unsigned long long* __restrict__ tilesVisited)
{
    __shared__ float sx[SOR_TILE], sy[SOR_TILE], sz[SOR_TILE];
    __shared__ unsigned sWorst;
    __shared__ float sBoxQ[6];
    const int tid = threadIdx.x, t0 = blockIdx.x;
    const long long i = (long long)t0 * SOR_TILE + tid;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    bool live = false;                               // a finite query point
    if (i < n) { const orbx_cloud_point p = pts[i]; qx = p.x; qy = p.y; qz = p.z; live = finite3(qx, qy, qz); }
    if (tid < 6) sBoxQ[tid] = box[(size_t)t0 * 6 + tid];
    __syncthreads();
    float best[SOR_KMAX];                            // ascending; best[0 .. cnt)
    int cnt = 0;
    float worst = FLT_MAX;                           // best[K - 1] once the list is full
    unsigned long long visited = 0;
    // tiles outward from the block's own: t0, t0 + 1, t0 - 1, t0 + 2, ...
    float blockWorst = FLT_MAX;                      // the block's worst current neighbour distance; it only changes after a visit
    for (int step = 0; step < 2 * ntiles; ++step) {
        const int off = (step + 1) >> 1;
        const int tile = (step & 1) ? t0 + off : t0 - off;
        if (tile < 0 || tile >= ntiles) continue;
        // the margin covers the rounding of the box distance against the point distances
        if (box_dist2(sBoxQ, box + (size_t)tile * 6) * 0.9999f > blockWorst) continue;      // uniform over the block
        ++visited;
        const long long j = (long long)tile * SOR_TILE + tid;
        float cx = INFINITY, cy = 0.f, cz = 0.f;     // an absent or non-finite candidate is infinitely far
        if (j < n) { const orbx_cloud_point c = pts[j]; if (finite3(c.x, c.y, c.z)) { cx = c.x; cy = c.y; cz = c.z; } }
        __syncthreads();                             // the previous tile is consumed
        sx[tid] = cx; sy[tid] = cy; sz[tid] = cz;
        __syncthreads();
        if (live) {
            for (int c = 0; c < SOR_TILE; ++c) {
                const float dx = __fsub_rn(qx, sx[c]), dy = __fsub_rn(qy, sy[c]), dz = __fsub_rn(qz, sz[c]);
                const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                if (d2 < worst) {
                    int p = cnt < K ? cnt : K - 1;   // slot that is freed (the list grows or drops its last entry)
                    while (p > 0 && best[p - 1] > d2) { best[p] = best[p - 1]; --p; }
                    best[p] = d2;
                    if (cnt < K) ++cnt;
                    if (cnt == K) worst = best[K - 1];
                }
            }
        }
        // non-negative floats order like their bit patterns
        if (tid == 0) sWorst = 0u;
        __syncthreads();
        if (live) atomicMax(&sWorst, __float_as_uint(worst));
        __syncthreads();
        blockWorst = __uint_as_float(sWorst);
    }
    if (i < n) {
        float md = 0.f;
        unsigned char ok = 0;
        if (live && cnt == K) {
            double sum = 0.0;
            for (int k = 1; k < K; ++k) sum += sqrt((double)best[k]);        // k = 0 is the query point itself
            md = (float)(sum / (double)(K - 1));
            ok = 1;
        }
        meanDist[i] = md; valid[i] = ok;
    }
    if (tid == 0 && tilesVisited) atomicAdd(tilesVisited, visited);
}

// sum and sum of squares of the mean distances: per-block partials (fixed order), then one thread adds the partials in order
__global__ void __launch_bounds__(256)
sor_partial_sums_kernel(const float* __restrict__ meanDist, const unsigned char* __restrict__ valid, long long n, int per,
                        double* __restrict__ part /* [blocks][3] */)
{
    __shared__ double red[3][256];
    const long long b0 = ((long long)blockIdx.x * 256 + threadIdx.x) * per;
    double s = 0.0, q = 0.0, c = 0.0;
    for (long long i = b0; i < b0 + per && i < n; ++i) {
        const float d = meanDist[i];
        s += (double)d; q += (double)__fmul_rn(d, d);     // PCL: sq_sum += distances[i] * distances[i] (a float product)
        c += valid[i] ? 1.0 : 0.0;
    }
    red[0][threadIdx.x] = s; red[1][threadIdx.x] = q; red[2][threadIdx.x] = c;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x < 3) part[(size_t)blockIdx.x * 3 + threadIdx.x] = red[threadIdx.x][0];
}

__global__ void sor_threshold_kernel(const double* __restrict__ part, int nblocks, double mul, double* __restrict__ out /* thr, mean, stddev, valid */)
{
    double s = 0.0, q = 0.0, c = 0.0;
    for (int b = 0; b < nblocks; ++b) { s += part[3 * b]; q += part[3 * b + 1]; c += part[3 * b + 2]; }
    const double mean = s / c;
    const double variance = (q - s * s / c) / (c - 1.0);
    const double stddev = sqrt(variance);
    out[0] = mean + mul * stddev; out[1] = mean; out[2] = stddev; out[3] = c;
}

__global__ void __launch_bounds__(256)
sor_flag_kernel(const float* __restrict__ meanDist, long long n, const double* __restrict__ thr, uint32_t* __restrict__ keep)
{
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) keep[i] = ((double)meanDist[i] > thr[0]) ? 0u : 1u;         // a NaN threshold (no valid distance) keeps everything
}

// exclusive scan of 0/1 flags: block sums, one block over the sums, apply
__device__ __forceinline__ uint32_t sor_block_scan(uint32_t v, uint32_t* wsum, uint32_t& total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += u; }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    uint32_t off = 0, tot = 0;
    for (int w = 0; w < 8; ++w) { const uint32_t s = wsum[w]; if (w < wid) off += s; tot += s; }
    __syncthreads();
    total = tot;
    return off + incl - v;
}

__global__ void __launch_bounds__(256)
sor_scan_sums_kernel(const uint32_t* __restrict__ keep, long long n, uint32_t* __restrict__ sums)
{
    __shared__ uint32_t wsum[8];
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    uint32_t total;
    sor_block_scan(i < n ? keep[i] : 0u, wsum, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(256)
sor_scan_top_kernel(uint32_t* __restrict__ sums, int nb)
{
    __shared__ uint32_t wsum[8];
    uint32_t carry = 0;
    for (int b0 = 0; b0 < nb; b0 += 256) {
        const int b = b0 + threadIdx.x;
        uint32_t total;
        const uint32_t ex = sor_block_scan(b < nb ? sums[b] : 0u, wsum, total);
        if (b < nb) sums[b] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[nb] = carry;
}

__global__ void __launch_bounds__(256)
sor_compact_kernel(const orbx_cloud_point* __restrict__ pts, const uint32_t* __restrict__ keep, long long n, const uint32_t* __restrict__ sums,
                   orbx_cloud_point* __restrict__ out, long long capacity)
{
    __shared__ uint32_t wsum[8];
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    const uint32_t k = i < n ? keep[i] : 0u;
    uint32_t total;
    const uint32_t pos = sums[blockIdx.x] + sor_block_scan(k, wsum, total);
    if (k && out && (long long)pos < capacity) out[pos] = pts[i];
}

}  // namespace

extern "C" {

int orbm_statistical_outlier_removal(const orbx_cloud_point* points, int64_t n, int mean_k, double stddev_mul, int device,
                                     orbx_cloud_point* out, int64_t capacity, int64_t* nout, float* mean_dist, double* stats)
{
    ORBX_NVTX("orbm_statistical_outlier_removal");
    if (!nout || (n > 0 && !points)) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    *nout = 0;
    if (mean_k < 1 || mean_k + 1 > SOR_KMAX) { orbx_set_error("mean_k must be 1..%d", SOR_KMAX - 1); return ORBX_E_INVALID; }
    if (n <= 0) return ORBX_OK;
    if (n > 0x7fffffffLL - SOR_TILE) { orbx_set_error("more than 2^31 points"); return ORBX_E_CAPACITY; }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (ndev <= 0 || device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available; there is no CPU fallback", device); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    cudaStream_t st = orbx_call_stream();
    const int ntiles = (int)((n + SOR_TILE - 1) / SOR_TILE);
    const int K = mean_k + 1;
    const int per = 16, nbSum = (int)((n + 256LL * per - 1) / (256LL * per));
    struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } } dPts, dBox, dDist, dValid, dPart, dThr, dKeep, dSums, dOut, dVisited;
    ORBX_CUDA(cudaMalloc(&dPts.p, (size_t)n * sizeof(orbx_cloud_point)));
    ORBX_CUDA(cudaMalloc(&dBox.p, (size_t)ntiles * 6 * sizeof(float)));
    ORBX_CUDA(cudaMalloc(&dDist.p, (size_t)n * sizeof(float)));
    ORBX_CUDA(cudaMalloc(&dValid.p, (size_t)n));
    ORBX_CUDA(cudaMalloc(&dPart.p, (size_t)nbSum * 3 * sizeof(double)));
    ORBX_CUDA(cudaMalloc(&dThr.p, 4 * sizeof(double)));
    ORBX_CUDA(cudaMalloc(&dKeep.p, (size_t)n * sizeof(uint32_t)));
    ORBX_CUDA(cudaMalloc(&dSums.p, (size_t)(ntiles + 1) * sizeof(uint32_t)));
    ORBX_CUDA(cudaMalloc(&dVisited.p, 8));
    ORBX_CUDA(cudaMemsetAsync(dVisited.p, 0, 8, st));
    ORBX_CUDA(cudaMemcpyAsync(dPts.p, points, (size_t)n * sizeof(orbx_cloud_point), cudaMemcpyHostToDevice, st));
    const orbx_cloud_point* P = (const orbx_cloud_point*)dPts.p;
    sor_tile_bounds_kernel<<<ntiles, SOR_TILE, 0, st>>>(P, n, (float*)dBox.p);
    sor_mean_dist_kernel<<<ntiles, SOR_TILE, 0, st>>>(P, n, (const float*)dBox.p, ntiles, K, (float*)dDist.p, (unsigned char*)dValid.p,
                                                    (unsigned long long*)dVisited.p);
    sor_partial_sums_kernel<<<nbSum, 256, 0, st>>>((const float*)dDist.p, (const unsigned char*)dValid.p, n, per, (double*)dPart.p);
    sor_threshold_kernel<<<1, 1, 0, st>>>((const double*)dPart.p, nbSum, stddev_mul, (double*)dThr.p);
    sor_flag_kernel<<<ntiles, 256, 0, st>>>((const float*)dDist.p, n, (const double*)dThr.p, (uint32_t*)dKeep.p);
    sor_scan_sums_kernel<<<ntiles, 256, 0, st>>>((const uint32_t*)dKeep.p, n, (uint32_t*)dSums.p);
    sor_scan_top_kernel<<<1, 256, 0, st>>>((uint32_t*)dSums.p, ntiles);
    ORBX_CUDA(cudaGetLastError());
    uint32_t kept = 0;
    ORBX_CUDA(cudaMemcpyAsync(&kept, (const uint32_t*)dSums.p + ntiles, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    double hs[5] = {0, 0, 0, 0, 0};
    unsigned long long visited = 0;
    ORBX_CUDA(cudaMemcpyAsync(hs, dThr.p, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaMemcpyAsync(&visited, dVisited.p, 8, cudaMemcpyDeviceToHost, st));
    if (mean_dist) ORBX_CUDA(cudaMemcpyAsync(mean_dist, dDist.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    *nout = (int64_t)kept;
    if (stats) { for (int k = 0; k < 4; ++k) stats[k] = hs[k]; stats[4] = (double)visited / ((double)ntiles * (double)ntiles); }
    if (!out) return ORBX_OK;
    if ((int64_t)kept > capacity) { orbx_set_error("%u points kept, capacity %lld", kept, (long long)capacity); return ORBX_E_CAPACITY; }
    if (kept == 0) return ORBX_OK;
    ORBX_CUDA(cudaMalloc(&dOut.p, (size_t)kept * sizeof(orbx_cloud_point)));
    sor_compact_kernel<<<ntiles, 256, 0, st>>>(P, (const uint32_t*)dKeep.p, n, (const uint32_t*)dSums.p, (orbx_cloud_point*)dOut.p, (long long)kept);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(out, dOut.p, (size_t)kept * sizeof(orbx_cloud_point), cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

}  // extern "C"
