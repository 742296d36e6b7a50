// orbx_kfmatch.cu — the keyframe-rate matchers as sm_100a kernels behind the C ABI:
//
//   K19 area_best_kernel       the per-MapPoint search of ORBmatcher::Fuse (src/ORBmatcher.cc:825-973, 975-1100) and of
//                              either direction of SearchBySim3 (:1102-1326) over KeyFrame::GetFeaturesInArea
//                              (src/KeyFrame.cc:607-646)
//   K20 sim3_agree_kernel      SearchBySim3's agreement check (:1307-1322)
//   K21 triangulation_kernel   ORBmatcher::SearchForTriangulation (:657-823) with CheckDistEpipolarLine (:140-157)
//
// None of the three loops carries state from one MapPoint / feature to the next (Fuse's side effects are replayed by the
// host wrapper; SearchForTriangulation never sets vbMatched2), so every query is one warp's work and the grid is as
// wide as the query list.  K19: lanes own grid columns of the search window; "first least distance in scan order"
// (strict '<') = the smallest key dist << 16 | position in the cell-sorted index list.  K21: FeatureVectors as a
// block bitonic sort of (node << 32 | index) keys like K17; a candidate replaces the best when dist <= bestDist, so the
// winner is the least distance with the LAST position among equals.
// Per-keyframe sizes (the stateless calls): K19 has a grid-free sweep form (area_best_sweep_kernel: the lanes sweep the keyframe's
// keypoints with the window and cell-window tests, both SearchBySim3 legs in one launch), and K21 splits the single pair of a call
// over several CTAs of 1024 threads with a last-ticket ComputeThreeMaxima; both forms are bit-identical to the ones above.

#include "orbx_internal.h"
#include "orbx_match_common.cuh"

#include <algorithm>
#include <vector>

#define AB_WARPS 8

struct AreaBestArgs {
    FrameStore fs;
    GridParams gp;
    const float* uRight;        // mvuRight of the keyframe or nullptr (all -1)
    float invSigma2[ORBX_MAX_LEVELS];
    int variant, thDist;
};

__global__ void __launch_bounds__(AB_WARPS * 32)
area_best_kernel(AreaBestArgs A, const orbm_proj_query* __restrict__ queries, const u8* __restrict__ qdesc, int nq,
                 int* __restrict__ bestIdxOut, int* __restrict__ bestDistOut)
{
    const int lane = threadIdx.x & 31, q = blockIdx.x * AB_WARPS + (threadIdx.x >> 5);
    if (q >= nq) return;
    const orbx_keypoint* kps = A.fs.kps;
    const unsigned short* gs = A.fs.gridStart;
    const unsigned short* gi = A.fs.gridIdx;
    const orbm_proj_query Q = queries[q];
    unsigned best = 0xffffffffu;
    int cx0, cx1, cy0, cy1;
    if (area_cells(A.gp, Q.u, Q.v, Q.radius, cx0, cx1, cy0, cy1)) {
        uint32_t qd[8];
        load_desc(qd, qdesc + (size_t)q * 32);
        const int ncy = cy1 - cy0 + 1;
        for (int ix = cx0 + lane; ix <= cx1; ix += 32) {
            const int e = gs[ix * ORBX_GRID_ROWS + cy0 + ncy];
            for (int j = gs[ix * ORBX_GRID_ROWS + cy0]; j < e; ++j) {
                const int idx = gi[j];
                const orbx_keypoint kp = kps[idx];
                if (!(fabsf(__fsub_rn(kp.x, Q.u)) < Q.radius && fabsf(__fsub_rn(kp.y, Q.v)) < Q.radius)) continue;
                if (kp.octave < Q.level_hi - 1 || kp.octave > Q.level_hi) continue;
                if (A.variant == ORBM_FUSE_MAPPOINTS) {
                    const float kpr = A.uRight ? A.uRight[idx] : -1.f;
                    const float ex = __fsub_rn(Q.u, kp.x), ey = __fsub_rn(Q.v, kp.y);
                    float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                    double lim = 5.99;
                    if (kpr >= 0) { const float er = __fsub_rn(Q.u_right, kpr); e2 = __fadd_rn(e2, __fmul_rn(er, er)); lim = 7.8; }
                    if ((double)__fmul_rn(e2, A.invSigma2[kp.octave]) > lim) continue;
                }
                const int dist = hamming256(qd, reinterpret_cast<const uint32_t*>(A.fs.desc + (size_t)idx * 32));
                best = min(best, ((unsigned)dist << 16) | (unsigned)j);
            }
        }
    }
    for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(FULL, best, d));
    if (lane == 0) {
        const int dist = best == 0xffffffffu ? 256 : (int)(best >> 16);
        bestDistOut[q] = dist;
        bestIdxOut[q] = (best != 0xffffffffu && dist <= A.thDist) ? (int)gi[best & 0xffff] : -1;
    }
}

// K20, latency form (keyframes of at most AB_SWEEP_MAXN keypoints, the per-keyframe calls): no grid.  The lanes of a warp sweep
// the keyframe's keypoints (coalesced, no dependent loads) with the window test AND the cell-window test of
// KeyFrame::GetFeaturesInArea (a keypoint counts only if the cell it is filed under lies in the query's cell window); ties
// between equal distances go to the first candidate in the grid's enumeration order = the smallest (cell, index).  Up to two
// legs (SearchBySim3 searches KF1's points in KF2 and KF2's in KF1) share one launch: blockIdx.y picks the leg.
#define AB_SWEEP_MAXN 4096
struct AreaSweepLeg {
    AreaBestArgs A;
    const orbm_proj_query* queries; const u8* qdesc; int nq, n;
    int* bestIdx; int* bestDist;
};
struct AreaSweepLegs { AreaSweepLeg leg[2]; };

__global__ void __launch_bounds__(AB_WARPS * 32)
area_best_sweep_kernel(AreaSweepLegs L)
{
    const AreaSweepLeg& G = L.leg[blockIdx.y];
    const AreaBestArgs& A = G.A;
    const int lane = threadIdx.x & 31, q = blockIdx.x * AB_WARPS + (threadIdx.x >> 5);
    if (q >= G.nq) return;
    const orbx_keypoint* kps = A.fs.kps;
    const orbm_proj_query Q = G.queries[q];
    unsigned long long best = ~0ull;
    int cx0, cx1, cy0, cy1;
    if (area_cells(A.gp, Q.u, Q.v, Q.radius, cx0, cx1, cy0, cy1)) {
        uint32_t qd[8];
        load_desc(qd, G.qdesc + (size_t)q * 32);
        for (int i = lane; i < G.n; i += 32) {
            const float kx = kps[i].x, ky = kps[i].y;
            if (!(fabsf(__fsub_rn(kx, Q.u)) < Q.radius && fabsf(__fsub_rn(ky, Q.v)) < Q.radius)) continue;
            const int oct = kps[i].octave;
            if (oct < Q.level_hi - 1 || oct > Q.level_hi) continue;
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kx, A.gp.minX), A.gp.invW));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(ky, A.gp.minY), A.gp.invH));
            if (px < cx0 || px > cx1 || py < cy0 || py > cy1) continue;           // the cell window lies inside the grid
            if (A.variant == ORBM_FUSE_MAPPOINTS) {
                const float kpr = A.uRight ? A.uRight[i] : -1.f;
                const float ex = __fsub_rn(Q.u, kx), ey = __fsub_rn(Q.v, ky);
                float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                double lim = 5.99;
                if (kpr >= 0) { const float er = __fsub_rn(Q.u_right, kpr); e2 = __fadd_rn(e2, __fmul_rn(er, er)); lim = 7.8; }
                if ((double)__fmul_rn(e2, A.invSigma2[oct]) > lim) continue;
            }
            const int dist = hamming256(qd, reinterpret_cast<const uint32_t*>(A.fs.desc + (size_t)i * 32));
            const unsigned long long key = ((unsigned long long)dist << 32) | ((unsigned long long)(px * ORBX_GRID_ROWS + py) << 16) | (unsigned long long)i;
            best = key < best ? key : best;
        }
    }
    for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_xor_sync(FULL, best, d); best = o < best ? o : best; }
    if (lane == 0) {
        const int dist = best == ~0ull ? 256 : (int)(best >> 32);
        G.bestDist[q] = dist;
        G.bestIdx[q] = (best != ~0ull && dist <= A.thDist) ? (int)(best & 0xffff) : -1;
    }
}

// vnMatch1 / vnMatch2 scatter + agreement.  One CTA.
__global__ void sim3_agree_kernel(const int* __restrict__ best12, const int* __restrict__ index1, int nq12,
                                  const int* __restrict__ best21, const int* __restrict__ index2, int nq21, int n1, int n2,
                                  int* __restrict__ vnMatch1, int* __restrict__ vnMatch2, int* __restrict__ match12,
                                  int* __restrict__ nfound)
{
    __shared__ int cnt;
    if (threadIdx.x == 0) cnt = 0;
    for (int i = threadIdx.x; i < n1; i += blockDim.x) vnMatch1[i] = -1;
    for (int i = threadIdx.x; i < n2; i += blockDim.x) vnMatch2[i] = -1;
    __syncthreads();
    for (int q = threadIdx.x; q < nq12; q += blockDim.x) vnMatch1[index1[q]] = best12[q];
    for (int q = threadIdx.x; q < nq21; q += blockDim.x) vnMatch2[index2[q]] = best21[q];
    __syncthreads();
    int mine = 0;
    for (int i1 = threadIdx.x; i1 < n1; i1 += blockDim.x) {
        const int idx2 = vnMatch1[i1];
        const bool ok = idx2 >= 0 && vnMatch2[idx2] == i1;
        match12[i1] = ok ? idx2 : -1;
        mine += ok;
    }
    atomicAdd(&cnt, mine);
    __syncthreads();
    if (threadIdx.x == 0) *nfound = cnt;
}

#define TRI_THREADS 1024
struct TriPair {
    int offA, nA, offB, nB;
    float F[9], ex, ey;
    float scale[ORBX_MAX_LEVELS], sigma2[ORBX_MAX_LEVELS];
};

__device__ void tri_bitonic_sort64(unsigned long long* a, int cap)
{
    for (int k = 2; k <= cap; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < cap; i += TRI_THREADS) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long x = a[i], y = a[l];
                    if ((x > y) == ((i & k) == 0)) { a[i] = y; a[l] = x; }
                }
            }
            __syncthreads();
        }
}

// gridDim = (pairs, splits): the features of KF1 are independent of each other (the reference keeps no "already matched" state
// on the KF2 side of this search that a later feature reads), so a pair's KF1 features are split over `splits` CTAs -- one for
// a batch, many for the single pair of a per-keyframe call.  Every CTA sorts KF2's (node, index) keys for itself; the rotation
// histogram, the count and a ticket live in global memory (pairState: 64 ints per pair, zeroed by the host) and the CTA that
// draws the last ticket applies ComputeThreeMaxima to the whole pair.
#define TRI_STATE_INTS 64
__global__ void __launch_bounds__(TRI_THREADS)
triangulation_kernel(const TriPair* __restrict__ pairs, const orbx_keypoint* __restrict__ kpA, const u8* __restrict__ descA,
                     const uint32_t* __restrict__ nodeA, const u8* __restrict__ flagA, const float* __restrict__ urA,
                     const orbx_keypoint* __restrict__ kpB, const u8* __restrict__ descB, const uint32_t* __restrict__ nodeB,
                     const u8* __restrict__ flagB, const float* __restrict__ urB, int onlyStereo, int checkOri,
                     int capB, int32_t* __restrict__ matches12, int32_t* __restrict__ nmatchesOut, u8* __restrict__ binOfAll,
                     int* __restrict__ pairState)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    __shared__ int sInd[3], sLast, sHist[ORBX_HISTO_LENGTH];
    __shared__ TriPair P;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid < (int)(sizeof(TriPair) / 4)) reinterpret_cast<int*>(&P)[tid] = reinterpret_cast<const int*>(pairs + blockIdx.x)[tid];
    __syncthreads();
    unsigned long long* keysB = reinterpret_cast<unsigned long long*>(smraw);
    int* gHist = pairState + (size_t)blockIdx.x * TRI_STATE_INTS;         // [0..29] histogram, [32] count, [33] ticket
    u8* binOf = binOfAll + P.offA;
    kpA += P.offA; descA += (size_t)P.offA * 32; nodeA += P.offA; flagA += P.offA; urA += P.offA; matches12 += P.offA;
    kpB += P.offB; descB += (size_t)P.offB * 32; nodeB += P.offB; flagB += P.offB; urB += P.offB;
    int pb = 1; while (pb < P.nB) pb <<= 1;
    const int perSplit = (P.nA + (int)gridDim.y - 1) / (int)gridDim.y;
    const int aBeg = min(P.nA, (int)blockIdx.y * perSplit), aEnd = min(P.nA, aBeg + perSplit);
    // B: every FeatureVector entry (skips are per candidate)
    for (int i = tid; i < pb; i += TRI_THREADS)
        keysB[i] = (i < P.nB && (flagB[i] & 1)) ? (((unsigned long long)nodeB[i] << 32) | (unsigned)i) : ~0ull;
    for (int i = aBeg + tid; i < aEnd; i += TRI_THREADS) { matches12[i] = -1; binOf[i] = 0xff; }
    __syncthreads();
    if (pb <= 4 * TRI_THREADS) block_sort64<TRI_THREADS>(keysB, pb); else tri_bitonic_sort64(keysB, pb);
    int accepted = 0;
    for (int idx1 = aBeg + wid; idx1 < aEnd; idx1 += TRI_THREADS / 32) {
        // A: only the features the loop does not skip (:699-707)
        bool use = (flagA[idx1] & 1) && !(flagA[idx1] & 2);
        if (use && onlyStereo && !(urA[idx1] >= 0)) use = false;
        if (!use) continue;
        const unsigned node = nodeA[idx1];
        int lo = 0, hi = pb;
        const unsigned long long target = (unsigned long long)node << 32;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (keysB[mid] < target) lo = mid + 1; else hi = mid; }
        const orbx_keypoint kp1 = kpA[idx1];
        const bool stereo1 = urA[idx1] >= 0;
        // epipolar line in the second image l = x1' F12 = [a b c]
        const float a = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F[0]), __fmul_rn(kp1.y, P.F[3])), P.F[6]);
        const float b = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F[1]), __fmul_rn(kp1.y, P.F[4])), P.F[7]);
        const float c = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F[2]), __fmul_rn(kp1.y, P.F[5])), P.F[8]);
        const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
        uint32_t q[8];
        load_desc(q, descA + (size_t)idx1 * 32);
        unsigned best = 0xffffffffu;                               // dist << 16 | (0xffff - position): last of the least wins
        for (int jb0 = lo; jb0 < pb; jb0 += 32) {
            const int jb = jb0 + lane;
            const unsigned long long kB = jb < pb ? keysB[jb] : ~0ull;
            const bool inRun = kB != ~0ull && (unsigned)(kB >> 32) == node;
            if (inRun) {
                const int idx2 = (int)(kB & 0xffffffffu);
                const bool stereo2 = urB[idx2] >= 0;
                bool ok = !(flagB[idx2] & 2) && !(onlyStereo && !stereo2);
                int dist = 0;
                if (ok) { dist = hamming256(q, reinterpret_cast<const uint32_t*>(descB + (size_t)idx2 * 32)); ok = dist <= ORBX_TH_LOW; }
                if (ok) {
                    const orbx_keypoint kp2 = kpB[idx2];
                    if (!stereo1 && !stereo2) {
                        const float dx = __fsub_rn(P.ex, kp2.x), dy = __fsub_rn(P.ey, kp2.y);
                        if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, P.scale[kp2.octave])) ok = false;
                    }
                    if (ok) {
                        const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, kp2.x), __fmul_rn(b, kp2.y)), c);
                        ok = den != 0.f && (double)__fdiv_rn(__fmul_rn(num, num), den) < __dmul_rn(3.84, (double)P.sigma2[kp2.octave]);
                    }
                    if (ok) best = min(best, ((unsigned)dist << 16) | (unsigned)(0xffff - (jb - lo)));
                }
            }
            if (!__all_sync(FULL, inRun)) break;
        }
        for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(FULL, best, d));
        if (best != 0xffffffffu) {
            const int idx2 = (int)(keysB[lo + (0xffff - (int)(best & 0xffff))] & 0xffffffffu);
            if (lane == 0) {
                matches12[idx1] = idx2;
                if (checkOri) { const int bin = rot_bin(kp1.angle, kpB[idx2].angle); binOf[idx1] = (u8)bin; atomicAdd(&gHist[bin], 1); }
            }
            ++accepted;
        }
    }
    if (lane == 0 && accepted) atomicAdd(&gHist[32], accepted);
    __threadfence();
    __syncthreads();
    if (tid == 0) sLast = atomicAdd(&gHist[33], 1) == (int)gridDim.y - 1;
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    volatile int* vh = gHist;
    if (checkOri) {
        if (tid < ORBX_HISTO_LENGTH) sHist[tid] = vh[tid];                 // 30 loads in flight at once, not one after the other
        __syncthreads();
        if (tid == 0) { int i1, i2, i3; three_maxima(sHist, ORBX_HISTO_LENGTH, i1, i2, i3); sInd[0] = i1; sInd[1] = i2; sInd[2] = i3; }
        __syncthreads();
        int culled = 0;
        volatile u8* vb = binOf;
        for (int i = tid; i < P.nA; i += TRI_THREADS) {
            const int bnum = vb[i];
            if (bnum != 0xff && bnum != sInd[0] && bnum != sInd[1] && bnum != sInd[2]) { matches12[i] = -1; ++culled; }
        }
        if (culled) atomicSub(&gHist[32], culled);
        __syncthreads();
    }
    if (tid == 0) nmatchesOut[blockIdx.x] = vh[32];
}

// ---- host side -------------------------------------------------------------------------------------------------
static int kf_device(int device)
{
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    return ORBX_OK;
}

static int kf_check(const orbm_keyframe* kf)
{
    const orbm_frame* f = kf ? &kf->frame : nullptr;
    if (!f || f->n < 0 || (f->n > 0 && (!f->keys_un || !f->descriptors))) { orbx_set_error("bad orbm_keyframe"); return ORBX_E_INVALID; }
    if (f->n > 65000) { orbx_set_error("more than 65000 keypoints per keyframe unsupported"); return ORBX_E_CAPACITY; }
    if (f->nlevels < 0 || f->nlevels > ORBX_MAX_LEVELS) { orbx_set_error("nlevels out of range"); return ORBX_E_INVALID; }
    return ORBX_OK;
}

// one keyframe of a call's arena (CallArena): keypoints, descriptors, count and, for Fuse variant 0, mvuRight go up with the
// call's single upload; the grid is built on the device behind it
struct ArenaKeyFrame {
    int sCount = -1, sKps = -1, sDesc = -1, sUR = -1, sGS = -1, sGI = -1, n = 0;
    int declare(CallArena& ar, const orbm_keyframe* kf, int variant)
    {
        const orbm_frame* f = &kf->frame;
        if (!(f->max_x > f->min_x) || !(f->max_y > f->min_y)) { orbx_set_error("bad keyframe bounds"); return ORBX_E_INVALID; }
        if (variant == ORBM_FUSE_MAPPOINTS && (!kf->inv_level_sigma2 || f->nlevels <= 0)) { orbx_set_error("Fuse variant 0 needs inv_level_sigma2 / nlevels"); return ORBX_E_INVALID; }
        n = f->n;
        const size_t cap = (size_t)std::max(n, 8);
        sCount = ar.in(&n, 4);
        sKps = ar.in(n ? f->keys_un : nullptr, (size_t)n * sizeof(orbx_keypoint));
        sDesc = ar.in(n ? f->descriptors : nullptr, (size_t)n * 32);
        if (variant == ORBM_FUSE_MAPPOINTS && f->u_right && n > 0) sUR = ar.in(f->u_right, (size_t)n * 4);
        sGS = ar.work((ORBX_GRID_CELLS + 1) * sizeof(unsigned short));
        sGI = ar.work(cap * sizeof(unsigned short));
        return ORBX_OK;
    }
    bool sweep() const { return n <= AB_SWEEP_MAXN; }       // the latency form needs no grid
    int ready(CallArena& ar, const orbm_keyframe* kf, int variant, AreaBestArgs& args, bool needGrid)
    {
        const orbm_frame* f = &kf->frame;
        FrameStore fs;
        fs.kps = ar.dev<orbx_keypoint>(sKps); fs.desc = ar.dev<u8>(sDesc); fs.count = ar.dev<int>(sCount);
        fs.gridStart = ar.dev<unsigned short>(sGS); fs.gridIdx = ar.dev<unsigned short>(sGI); fs.cap = std::max(n, 8);
        args.gp = orbx_grid_params(f->min_x, f->max_x, f->min_y, f->max_y);
        if (needGrid) { const int rc = orbx_launch_grid_build(fs, 0, 1, args.gp, ar.st); if (rc) return rc; }
        args.fs = fs;
        args.uRight = sUR >= 0 ? ar.dev<float>(sUR) : nullptr;
        args.variant = variant;
        args.thDist = variant == ORBM_FUSE_SIM3_LEG ? ORBX_TH_HIGH : ORBX_TH_LOW;
        for (int l = 0; l < ORBX_MAX_LEVELS; ++l) args.invSigma2[l] = 0.f;
        if (variant == ORBM_FUSE_MAPPOINTS) for (int l = 0; l < f->nlevels; ++l) args.invSigma2[l] = kf->inv_level_sigma2[l];
        return ORBX_OK;
    }
};

struct ArenaQueries {
    int sQ = -1, sD = -1, sBI = -1, sBD = -1;
    // outIdx / outDist: host destinations (nullptr = the result stays on the device as work space)
    void declare(CallArena& ar, const orbm_proj_query* queries, const uint8_t* desc, int nq, int32_t* outIdx, int32_t* outDist)
    {
        sQ = ar.in(nq ? queries : nullptr, (size_t)nq * sizeof(orbm_proj_query));
        sD = ar.in(nq ? desc : nullptr, (size_t)nq * 32);
        sBI = outIdx ? ar.out(outIdx, (size_t)nq * 4) : ar.work((size_t)nq * 4);
        sBD = outDist ? ar.out(outDist, (size_t)nq * 4) : ar.work((size_t)nq * 4);
    }
    void launch(CallArena& ar, const AreaBestArgs& a, int nq)
    {
        if (nq > 0)
            area_best_kernel<<<(nq + AB_WARPS - 1) / AB_WARPS, AB_WARPS * 32, 0, ar.st>>>(a, ar.dev<orbm_proj_query>(sQ), ar.dev<u8>(sD), nq, ar.dev<int>(sBI), ar.dev<int>(sBD));
    }
    AreaSweepLeg leg(CallArena& ar, const AreaBestArgs& a, int nq, int n)
    {
        AreaSweepLeg g;
        g.A = a; g.queries = ar.dev<orbm_proj_query>(sQ); g.qdesc = ar.dev<u8>(sD); g.nq = nq; g.n = n;
        g.bestIdx = ar.dev<int>(sBI); g.bestDist = ar.dev<int>(sBD);
        return g;
    }
};

static void launch_area_sweep(const AreaSweepLeg& g0, const AreaSweepLeg* g1, cudaStream_t st)
{
    AreaSweepLegs L;
    L.leg[0] = g0; L.leg[1] = g1 ? *g1 : g0;
    const int nq = std::max(g0.nq, g1 ? g1->nq : 0);
    if (nq > 0) area_best_sweep_kernel<<<dim3((nq + AB_WARPS - 1) / AB_WARPS, g1 ? 2 : 1), AB_WARPS * 32, 0, st>>>(L);
}

extern "C" {

int orbm_fuse(int variant, const orbm_keyframe* kf, const orbm_proj_query* queries, const uint8_t* query_desc, int nq,
              int device, int32_t* best_idx, int32_t* best_dist)
{
    ORBX_NVTX("orbm_fuse");
    int rc = kf_check(kf); if (rc) return rc;
    if (variant < ORBM_FUSE_MAPPOINTS || variant > ORBM_FUSE_SIM3_LEG) { orbx_set_error("unknown Fuse variant %d", variant); return ORBX_E_INVALID; }
    if (nq < 0 || (nq > 0 && (!queries || !query_desc || !best_idx))) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (nq == 0) return ORBX_OK;
    rc = kf_device(device); if (rc) return rc;
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("fuse", acc, &calls); }
    ArenaKeyFrame dk;
    ArenaQueries dq;
    rc = dk.declare(ar, kf, variant); if (rc) return rc;
    dq.declare(ar, queries, query_desc, nq, best_idx, best_dist);
    rc = ar.commit(); if (rc) return rc;
    AreaBestArgs args;
    rc = dk.ready(ar, kf, variant, args, !dk.sweep()); if (rc) return rc;
    if (dk.sweep()) launch_area_sweep(dq.leg(ar, args, nq, dk.n), nullptr, ar.st);
    else dq.launch(ar, args, nq);
    return ar.finish();
}

int orbm_search_by_sim3(const orbm_keyframe* kf1, const orbm_keyframe* kf2, const orbm_proj_query* queries12,
                        const uint8_t* desc12, const int32_t* index1, int nq12, const orbm_proj_query* queries21,
                        const uint8_t* desc21, const int32_t* index2, int nq21, int device, int32_t* match12,
                        int32_t* nfound)
{
    ORBX_NVTX("orbm_search_by_sim3");
    int rc = kf_check(kf1); if (rc) return rc;
    rc = kf_check(kf2); if (rc) return rc;
    if (nq12 < 0 || nq21 < 0 || (nq12 > 0 && (!queries12 || !desc12 || !index1)) || (nq21 > 0 && (!queries21 || !desc21 || !index2)) ||
        (kf1->frame.n > 0 && !match12)) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    const int n1 = kf1->frame.n, n2 = kf2->frame.n;
    for (int q = 0; q < nq12; ++q) if (index1[q] < 0 || index1[q] >= n1) { orbx_set_error("index1[%d] out of range", q); return ORBX_E_INVALID; }
    for (int q = 0; q < nq21; ++q) if (index2[q] < 0 || index2[q] >= n2) { orbx_set_error("index2[%d] out of range", q); return ORBX_E_INVALID; }
    rc = kf_device(device); if (rc) return rc;
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("search_by_sim3", acc, &calls); }
    ArenaKeyFrame d1, d2;
    ArenaQueries q12, q21;
    rc = d1.declare(ar, kf1, ORBM_FUSE_SIM3_LEG); if (rc) return rc;
    rc = d2.declare(ar, kf2, ORBM_FUSE_SIM3_LEG); if (rc) return rc;
    q12.declare(ar, queries12, desc12, nq12, nullptr, nullptr);
    q21.declare(ar, queries21, desc21, nq21, nullptr, nullptr);
    const int sI1 = ar.in(nq12 ? index1 : nullptr, (size_t)nq12 * 4), sI2 = ar.in(nq21 ? index2 : nullptr, (size_t)nq21 * 4);
    int nf = 0;
    const int sM12 = ar.out(match12, (size_t)n1 * 4), sNf = ar.out(&nf, 4);
    const int sM1 = ar.work((size_t)n1 * 4), sM2 = ar.work((size_t)n2 * 4);
    rc = ar.commit(); if (rc) return rc;
    AreaBestArgs a1, a2;
    const bool sweep = d1.sweep() && d2.sweep();
    rc = d1.ready(ar, kf1, ORBM_FUSE_SIM3_LEG, a1, !sweep); if (rc) return rc;
    rc = d2.ready(ar, kf2, ORBM_FUSE_SIM3_LEG, a2, !sweep); if (rc) return rc;
    if (sweep) {
        const AreaSweepLeg g12 = q12.leg(ar, a2, nq12, d2.n), g21 = q21.leg(ar, a1, nq21, d1.n);   // points of KF1 searched in KF2 and back, one launch
        launch_area_sweep(g12, &g21, ar.st);
    } else {
        q12.launch(ar, a2, nq12);
        q21.launch(ar, a1, nq21);
    }
    sim3_agree_kernel<<<1, 256, 0, ar.st>>>(ar.dev<int>(q12.sBI), ar.dev<int>(sI1), nq12, ar.dev<int>(q21.sBI), ar.dev<int>(sI2), nq21, n1, n2,
                                  ar.dev<int>(sM1), ar.dev<int>(sM2), ar.dev<int>(sM12), ar.dev<int>(sNf));
    rc = ar.finish(); if (rc) return rc;
    if (nfound) *nfound = nf;
    return ORBX_OK;
}

int orbm_search_for_triangulation(const orbm_keyframe* kf1, const orbm_keyframe* kf2, int npairs, const float* f12,
                                  const float* epipole, int only_stereo, int check_orientation, int device,
                                  int32_t* matches12, int32_t* nmatches)
{
    ORBX_NVTX("orbm_search_for_triangulation");
    if (npairs <= 0) return ORBX_OK;
    if (!kf1 || !kf2 || !f12 || !epipole || !matches12 || !nmatches) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    std::vector<TriPair> pairs(npairs);
    long long totA = 0, totB = 0;
    int maxA = 1, maxB = 1;
    for (int p = 0; p < npairs; ++p) {
        int rc = kf_check(kf1 + p); if (rc) return rc;
        rc = kf_check(kf2 + p); if (rc) return rc;
        const orbm_frame &A = kf1[p].frame, &B = kf2[p].frame;
        if ((A.n > 0 && (!kf1[p].node_id || !kf1[p].flags)) || (B.n > 0 && (!kf2[p].node_id || !kf2[p].flags)) ||
            (B.n > 0 && (!B.scale_factors || !kf2[p].level_sigma2 || B.nlevels <= 0))) {
            orbx_set_error("pair %d: node_id / flags / scale_factors / level_sigma2 required", p); return ORBX_E_INVALID;
        }
        TriPair& P = pairs[p];
        P.offA = (int)totA; P.nA = A.n; P.offB = (int)totB; P.nB = B.n;
        for (int i = 0; i < 9; ++i) P.F[i] = f12[(size_t)p * 9 + i];
        P.ex = epipole[2 * p]; P.ey = epipole[2 * p + 1];
        for (int l = 0; l < ORBX_MAX_LEVELS; ++l) {
            P.scale[l] = l < B.nlevels ? B.scale_factors[l] : 0.f;
            P.sigma2[l] = l < B.nlevels ? kf2[p].level_sigma2[l] : 0.f;
        }
        totA += A.n; totB += B.n;
        maxA = std::max(maxA, A.n); maxB = std::max(maxB, B.n);
    }
    int capB = 1; while (capB < maxB) capB <<= 1;
    const size_t smem = (size_t)capB * 8 + 16;
    if (smem > 220 * 1024) { orbx_set_error("keyframes of %d / %d features exceed the shared-memory sort (%zu bytes)", maxA, maxB, smem); return ORBX_E_CAPACITY; }
    int rc = kf_device(device); if (rc) return rc;
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("search_for_triangulation", acc, &calls); }
    const int sPairs = ar.in(pairs.data(), npairs * sizeof(TriPair));
    const int sKA = ar.in(nullptr, (size_t)totA * sizeof(orbx_keypoint)), sDA = ar.in(nullptr, (size_t)totA * 32), sNA = ar.in(nullptr, (size_t)totA * 4),
              sFA = ar.in(nullptr, (size_t)totA), sUA = ar.in(nullptr, (size_t)totA * 4);
    const int sKB = ar.in(nullptr, (size_t)totB * sizeof(orbx_keypoint)), sDB = ar.in(nullptr, (size_t)totB * 32), sNB = ar.in(nullptr, (size_t)totB * 4),
              sFB = ar.in(nullptr, (size_t)totB), sUB = ar.in(nullptr, (size_t)totB * 4);
    const int sM = ar.out(matches12, (size_t)totA * 4), sNM = ar.out(nmatches, (size_t)npairs * 4);
    const int sBin = ar.work((size_t)totA), sState = ar.inout_fill(nullptr, (size_t)npairs * TRI_STATE_INTS * 4);   // zeroed in the staging buffer: rides on the upload
    rc = ar.layout(); if (rc) return rc;
    auto up = [&](const orbm_keyframe& kf, int off, int k, int d, int nd, int fl, int ur) {
        const orbm_frame& f = kf.frame;
        if (f.n <= 0) return;
        memcpy(ar.hostp<orbx_keypoint>(k) + off, f.keys_un, (size_t)f.n * sizeof(orbx_keypoint));
        memcpy(ar.hostp<u8>(d) + (size_t)off * 32, f.descriptors, (size_t)f.n * 32);
        memcpy(ar.hostp<uint32_t>(nd) + off, kf.node_id, (size_t)f.n * 4);
        memcpy(ar.hostp<u8>(fl) + off, kf.flags, (size_t)f.n);
        float* u = ar.hostp<float>(ur) + off;
        if (f.u_right) memcpy(u, f.u_right, (size_t)f.n * 4);
        else for (int i = 0; i < f.n; ++i) u[i] = -1.f;
    };
    for (int p = 0; p < npairs; ++p) {
        up(kf1[p], pairs[p].offA, sKA, sDA, sNA, sFA, sUA);
        up(kf2[p], pairs[p].offB, sKB, sDB, sNB, sFB, sUB);
    }
    memset(ar.hostp<u8>(sState), 0, ar.segs[sState].bytes);
    rc = ar.upload(); if (rc) return rc;
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(triangulation_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // a batch fills the GPU with one CTA per pair; the single pair of a per-keyframe call is spread over up to 16 SMs (32 warps each)
    const int splits = npairs >= 64 ? 1 : std::max(1, std::min(16, std::min(128 / npairs, (maxA + 127) / 128)));
    triangulation_kernel<<<dim3(npairs, splits), TRI_THREADS, smem, ar.st>>>(ar.dev<TriPair>(sPairs), ar.dev<orbx_keypoint>(sKA), ar.dev<u8>(sDA), ar.dev<uint32_t>(sNA),
                                                        ar.dev<u8>(sFA), ar.dev<float>(sUA), ar.dev<orbx_keypoint>(sKB), ar.dev<u8>(sDB),
                                                        ar.dev<uint32_t>(sNB), ar.dev<u8>(sFB), ar.dev<float>(sUB), only_stereo, check_orientation,
                                                        capB, ar.dev<int32_t>(sM), ar.dev<int32_t>(sNM), ar.dev<u8>(sBin), ar.dev<int>(sState));
    return ar.finish();
}

} // extern "C"
