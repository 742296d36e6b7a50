// orbx_bow.cu — the bag-of-words matcher path as sm_100a kernels behind the C ABI:
//
//   K16 bow_transform_kernel   DBoW2 TemplatedVocabulary::transform(feature, word, weight, nid, levelsup)
//                              Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1231-1271, FORB::distance FORB.cpp:81-101
//   K17 search_bow_kernel      ORBmatcher::SearchByBoW (KeyFrame, Frame) src/ORBmatcher.cc:159-288 and
//                              (KeyFrame, KeyFrame) src/ORBmatcher.cc:522-655
//   K18 distinctive_kernel     MapPoint::ComputeDistinctiveDescriptors src/MapPoint.cc:242-305 (rank 4)
//
// K16: one warp per feature, lane = child of the current node (first minimum wins, as the reference's scan).
// K17: one CTA per frame pair in a batch, several for the single pair of a per-frame call.  DBoW2's FeatureVector (std::map
// node -> ascending feature indices) becomes a register/shuffle bitonic sort of (node << 32 | index) keys; the reference's
// merge of the two maps visits exactly the common nodes, and a candidate feature belongs to one node only, so nodes are
// independent: one warp per node run replays the run's keyframe features in order (the "already matched" skip is sequential
// inside a node) with the run's candidates resident in its lanes, best / second best by two warp-wide integer minima.
// Rotation histogram + ComputeThreeMaxima cull by the CTA that finishes last.

#include "orbx_internal.h"
#include "orbx_match_common.cuh"

#include <algorithm>
#include <vector>

struct orbm_vocab_handle {
    int device, nnodes, depth;
    int32_t* childBegin; int32_t* childCount; int32_t* children;
    u8* desc; uint32_t* word; double* weight;
};

__global__ void __launch_bounds__(256)
bow_transform_kernel(const int32_t* __restrict__ childBegin, const int32_t* __restrict__ childCount,
                     const int32_t* __restrict__ children, const u8* __restrict__ nodeDesc,
                     const uint32_t* __restrict__ nodeWord, const double* __restrict__ nodeWeight, int depth,
                     const u8* __restrict__ feats, int n, int levelsup, uint32_t* __restrict__ wordOut,
                     double* __restrict__ weightOut, uint32_t* __restrict__ nodeOut)
{
    const int lane = threadIdx.x & 31, f = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (f >= n) return;
    uint32_t q[8];
    load_desc(q, feats + (size_t)f * 32);
    const int nidLevel = depth - levelsup;
    int finalId = 0, level = 0;
    uint32_t nid = 0;
    do {
        ++level;
        const int beg = childBegin[finalId], cnt = childCount[finalId];
        unsigned best = 0xffffffffu;                 // (distance << 16 | child position): first minimum wins
        for (int c = lane; c < cnt; c += 32) {
            const int id = children[beg + c];
            const unsigned key = ((unsigned)hamming256(q, reinterpret_cast<const uint32_t*>(nodeDesc + (size_t)id * 32)) << 16) | (unsigned)c;
            best = min(best, key);
        }
        for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(FULL, best, d));
        finalId = children[beg + (best & 0xffff)];
        if (level == nidLevel) nid = (uint32_t)finalId;
    } while (childCount[finalId] != 0);
    if (lane == 0) { wordOut[f] = nodeWord[finalId]; weightOut[f] = nodeWeight[finalId]; nodeOut[f] = nid; }
}

#define BOW_THREADS 1024
struct BowPair { int offA, nA, offB, nB; };

__device__ void block_bitonic_sort64(unsigned long long* a, int cap)
{
    for (int k = 2; k <= cap; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < cap; i += BOW_THREADS) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long x = a[i], y = a[l];
                    if ((x > y) == ((i & k) == 0)) { a[i] = y; a[l] = x; }
                }
            }
            __syncthreads();
        }
}

// gridDim = (pairs, splits).  Nodes are independent (a feature of B belongs to one node, so the "already matched" flags of one
// node are never read by another), which lets the node runs of a pair be dealt to the warps of `splits` CTAs: one CTA per
// pair in a batch, many for the single pair of a per-frame call.  Every CTA sorts both key lists for itself and lists A's node
// runs in ascending order by a block scan, so all CTAs hold the same list and deal run r to warp r mod (all warps) without talking.  Outputs, the
// per-feature bins and the pair state (histogram, count, ticket: BOW_STATE_INTS ints per pair) live in global memory and are
// preset to 0xff bytes by the host (match = -1, bin = none, counters = -1); the CTA that draws the last ticket applies
// ComputeThreeMaxima to the whole pair.
#define BOW_STATE_INTS 64
__global__ void __launch_bounds__(BOW_THREADS)
search_bow_kernel(int variant, const BowPair* __restrict__ pairs, const u8* __restrict__ descA, const float* __restrict__ angA,
                  const uint32_t* __restrict__ nodeA, const u8* __restrict__ flagA, const u8* __restrict__ descB,
                  const float* __restrict__ angB, const uint32_t* __restrict__ nodeB, const u8* __restrict__ flagB,
                  float nnratio, int checkOri, int capA, int capB, int32_t* __restrict__ matchA, int32_t* __restrict__ matchB,
                  int32_t* __restrict__ nmatchesOut, u8* __restrict__ binOfAll, int* __restrict__ pairState, int stage)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    __shared__ int sInd[3], sLast, sCull, sWarpRuns[BOW_THREADS / 32], sHist[ORBX_HISTO_LENGTH];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const BowPair P = pairs[blockIdx.x];
    unsigned long long* keysA = reinterpret_cast<unsigned long long*>(smraw);
    unsigned long long* keysB = keysA + capA;
    unsigned short* runs = reinterpret_cast<unsigned short*>(keysB + capB);      // start positions of A's node runs, ascending
    u8* matchedB = reinterpret_cast<u8*>(runs + capA);
    int* gState = pairState + (size_t)blockIdx.x * BOW_STATE_INTS;        // [0..29] histogram, [32] count, [33] ticket, all starting at -1
    u8* binOf = binOfAll + P.offA;
    descA += (size_t)P.offA * 32; angA += P.offA; nodeA += P.offA; flagA += P.offA; matchA += P.offA;
    descB += (size_t)P.offB * 32; angB += P.offB; nodeB += P.offB; flagB += P.offB; matchB += P.offB;
    int pa = 1; while (pa < P.nA) pa <<= 1;
    int pb = 1; while (pb < P.nB) pb <<= 1;
    for (int i = tid; i < pa; i += BOW_THREADS)
        keysA[i] = (i < P.nA && (flagA[i] & 1)) ? (((unsigned long long)nodeA[i] << 32) | (unsigned)i) : ~0ull;
    for (int i = tid; i < pb; i += BOW_THREADS)
        keysB[i] = (i < P.nB && (flagB[i] & 1)) ? (((unsigned long long)nodeB[i] << 32) | (unsigned)i) : ~0ull;
    for (int i = tid; i < P.nB; i += BOW_THREADS) matchedB[i] = 0;
    if (stage) {
        // per-frame sizes: descriptors, angles and flags of both frames are copied into shared memory (coalesced, in flight during
        // the sort), so that the replay of a node run never waits on a dependent global load
        u8* sp = reinterpret_cast<u8*>((reinterpret_cast<uintptr_t>(matchedB + capB) + 15) & ~(uintptr_t)15);
        uint4* sDA = reinterpret_cast<uint4*>(sp); sp += (size_t)P.nA * 32;
        uint4* sDB = reinterpret_cast<uint4*>(sp); sp += (size_t)P.nB * 32;
        float* sAA = reinterpret_cast<float*>(sp); sp += (size_t)P.nA * 4;
        float* sAB = reinterpret_cast<float*>(sp); sp += (size_t)P.nB * 4;
        u8* sFA = sp; sp += P.nA;
        u8* sFB = sp;
        for (int i = tid; i < P.nA * 2; i += BOW_THREADS) sDA[i] = reinterpret_cast<const uint4*>(descA)[i];
        for (int i = tid; i < P.nB * 2; i += BOW_THREADS) sDB[i] = reinterpret_cast<const uint4*>(descB)[i];
        for (int i = tid; i < P.nA; i += BOW_THREADS) { sAA[i] = angA[i]; sFA[i] = flagA[i]; }
        for (int i = tid; i < P.nB; i += BOW_THREADS) { sAB[i] = angB[i]; sFB[i] = flagB[i]; }
        descA = reinterpret_cast<const u8*>(sDA); descB = reinterpret_cast<const u8*>(sDB);
        angA = sAA; angB = sAB; flagA = sFA; flagB = sFB;
    }
    __syncthreads();
    if (pa == pb && pa <= 2 * BOW_THREADS) block_sort64_pair<BOW_THREADS>(keysA, keysB, pa);      // the usual case: both frames of a size
    else {
        if (pa <= 4 * BOW_THREADS) block_sort64<BOW_THREADS>(keysA, pa); else block_bitonic_sort64(keysA, pa);
        if (pb <= 4 * BOW_THREADS) block_sort64<BOW_THREADS>(keysB, pb); else block_bitonic_sort64(keysB, pb);
    }
    // A's node runs in ascending order (every CTA builds the same list, so the deal below needs no communication): thread t owns
    // a contiguous range of positions, counts its run starts, a block scan places them
    int nRuns;
    {
        const int per = (pa + BOW_THREADS - 1) / BOW_THREADS;
        const int p0 = tid * per, p1 = min(p0 + per, P.nA);
        auto is_start = [&](int i) { const unsigned long long k = keysA[i]; return k != ~0ull && (i == 0 || (keysA[i - 1] >> 32) != (k >> 32)); };
        int cnt = 0;
        for (int i = p0; i < p1; ++i) cnt += is_start(i);
        int incl = cnt;
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) sWarpRuns[wid] = incl;
        __syncthreads();
        int base = 0, total = 0;
        for (int w = 0; w < BOW_THREADS / 32; ++w) { const int v = sWarpRuns[w]; if (w < wid) base += v; total += v; }
        int o = base + incl - cnt;
        for (int i = p0; i < p1; ++i) if (is_start(i)) runs[o++] = (unsigned short)i;
        nRuns = total;
        __syncthreads();
    }
    int accepted = 0;
    const int warpsAll = (int)gridDim.y * (BOW_THREADS / 32);
    for (int r = (int)blockIdx.y * (BOW_THREADS / 32) + wid; r < nRuns; r += warpsAll) {
        const int a0 = runs[r];
        const unsigned long long k0 = keysA[a0];
        const unsigned node = (unsigned)(k0 >> 32);
        // B's run of the same node: [b0, b1)
        int lo = 0, hi = pb;
        const unsigned long long target = (unsigned long long)node << 32;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (keysB[mid] < target) lo = mid + 1; else hi = mid; }
        const int b0 = lo;
        int b1 = b0;
        while (b1 < pb && (unsigned)(keysB[b1] >> 32) == node && keysB[b1] != ~0ull) ++b1;      // runs are short (tens of features)
        if (b1 == b0) continue;
        const bool resident = b1 - b0 <= 32;                       // the usual case: lane j keeps B's j-th feature of the node in registers
        uint32_t db[8];
        int myB = -1;
        bool freeB = false;
        float myAngB = 0.f;
        if (resident && b0 + lane < b1) {
            myB = (int)(keysB[b0 + lane] & 0xffffffffu);
            freeB = !(variant == 1 && !(flagB[myB] & 2));
            load_desc(db, descB + (size_t)myB * 32);
            if (checkOri) myAngB = angB[myB];
        }
        // end of A's run (the run's entries are a prefix of every 32-entry window)
        int aEnd = a0;
        for (;;) {
            const int i = aEnd + lane;
            const unsigned m = __ballot_sync(FULL, i < P.nA && keysA[i] != ~0ull && (unsigned)(keysA[i] >> 32) == node);
            aEnd += __popc(m);
            if (m != FULL) break;
        }
        // A's features of the node, 32 at a time: every lane fetches one feature (flags, angle, descriptor) so that the loads
        // of a whole chunk fly together; the features are then replayed in index order from registers (shuffles)
        for (int c0 = a0; c0 < aEnd; c0 += 32) {
            const int cnt = min(32, aEnd - c0);
            int myA = 0, myFlag = 0;
            float myAngA = 0.f;
            uint32_t da[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (lane < cnt) {
                myA = (int)(keysA[c0 + lane] & 0xffffffffu);
                myFlag = flagA[myA];
                if (checkOri) myAngA = angA[myA];
                load_desc(da, descA + (size_t)myA * 32);
            }
            for (int t = 0; t < cnt; ++t) {
                if (!(__shfl_sync(FULL, myFlag, t) & 2)) continue;        // !pMP || pMP->isBad()
                const int idxA = __shfl_sync(FULL, myA, t);
                uint32_t q[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) q[i] = __shfl_sync(FULL, da[i], t);
                int best = 0x7fffffff, second = 0x7fffffff;
                unsigned key = 0xffffffffu;
                if (resident) {
                    // one candidate per lane: two warp-wide integer minima (REDUX) instead of a shuffle tree
                    const unsigned mine = freeB ? ((unsigned)hamming256_regs(q, db) << 8) | (unsigned)lane : 0xffffffffu;
                    const unsigned w1 = __reduce_min_sync(FULL, mine);
                    const unsigned w2 = __reduce_min_sync(FULL, mine == w1 ? 0xffffffffu : mine);
                    if (w1 != 0xffffffffu) { best = (int)(w1 >> 8); key = (unsigned)b0 + (w1 & 0xff); }
                    if (w2 != 0xffffffffu) second = (int)(w2 >> 8);
                } else {
                    for (int jb = b0 + lane; jb < b1; jb += 32) {
                        const int idxB = (int)(keysB[jb] & 0xffffffffu);
                        if (matchedB[idxB]) continue;
                        if (variant == 1 && !(flagB[idxB] & 2)) continue;
                        const int dist = hamming256(q, reinterpret_cast<const uint32_t*>(descB + (size_t)idxB * 32));
                        if (dist < best) { second = best; best = dist; key = (unsigned)jb; }
                        else if (dist < second) second = dist;
                    }
                }
                if (!resident) warp_top2(best, key, second);
                best = min(best, 256); second = min(second, 256);     // bestDist1 = bestDist2 = 256 initially; a distance of 256 never wins
                const bool found = best < 256;
                const bool close = variant == 0 ? best <= ORBX_TH_LOW : best < ORBX_TH_LOW;
                if (found && close && (float)best < __fmul_rn(nnratio, (float)second)) {
                    const int idxB = (int)(keysB[key] & 0xffffffffu);
                    float aB = 0.f;
                    if (checkOri) aB = resident ? __shfl_sync(FULL, myAngB, (int)key - b0) : angB[idxB];
                    const float aA = __shfl_sync(FULL, myAngA, t);
                    if (resident) { if ((int)key == b0 + lane) freeB = false; }
                    if (lane == 0) {
                        if (!resident) matchedB[idxB] = 1;
                        matchA[idxA] = idxB; matchB[idxB] = idxA;
                        if (checkOri) { const int bin = rot_bin(aA, aB); binOf[idxA] = (u8)bin; atomicAdd(&gState[bin], 1); }
                    }
                    ++accepted;
                    __syncwarp();
                }
            }
        }
    }
    if (lane == 0 && accepted) atomicAdd(&gState[32], accepted);
    __threadfence();
    __syncthreads();
    if (tid == 0) sLast = atomicAdd(&gState[33], 1) == (int)gridDim.y - 2;        // tickets run -1, 0, 1, ...
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    // everything the finish reads is requested at once (histogram, count, this thread's first bin and match): one trip to L2
    volatile int* vs = gState;
    volatile u8* vb = binOf;
    volatile int32_t* vmA = matchA;
    if (tid == 0) sCull = 0;
    const int count = tid == 0 ? vs[32] + 1 : 0;
    int b0 = 0xff, m0 = -1;
    if (checkOri) {
        if (tid < ORBX_HISTO_LENGTH) sHist[tid] = vs[tid] + 1;
        if (tid < P.nA) { b0 = vb[tid]; m0 = vmA[tid]; }
        __syncthreads();
        if (tid == 0) { int i1, i2, i3; three_maxima(sHist, ORBX_HISTO_LENGTH, i1, i2, i3); sInd[0] = i1; sInd[1] = i2; sInd[2] = i3; }
        __syncthreads();
        int culled = 0;
        for (int i = tid; i < P.nA; i += BOW_THREADS) {
            const int bb = i == tid ? b0 : (int)vb[i];
            if (bb != 0xff && bb != sInd[0] && bb != sInd[1] && bb != sInd[2]) { matchB[i == tid ? m0 : (int)vmA[i]] = -1; matchA[i] = -1; ++culled; }
        }
        if (culled) atomicAdd(&sCull, culled);
    }
    __syncthreads();
    if (tid == 0) nmatchesOut[blockIdx.x] = count - sCull;
}

// K18: one CTA per map point.  Thread i owns row i of the N x N distance matrix and finds its median
// (the element at sorted position int(0.5*(N-1))) by bisection over the value range 0..256 — distances are
// recomputed per step from the descriptors staged in shared memory, nothing of size N x N is stored.
#define DD_THREADS 128
#define DD_SMEM_DESC 2048           // descriptors staged in shared memory (64 KB); larger groups read global memory
__global__ void __launch_bounds__(DD_THREADS)
distinctive_kernel(const u8* __restrict__ desc, const int32_t* __restrict__ offsets, int32_t* __restrict__ bestIdx,
                   int32_t* __restrict__ bestMedian)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    __shared__ unsigned sBest;
    const int p = blockIdx.x, tid = threadIdx.x;
    const int o = offsets[p], N = offsets[p + 1] - o;
    if (N <= 0) { if (tid == 0) { bestIdx[p] = -1; if (bestMedian) bestMedian[p] = -1; } return; }
    const u8* d = desc + (size_t)o * 32;
    const bool staged = N <= DD_SMEM_DESC;
    if (staged) {
        for (int i = tid; i < N * 2; i += DD_THREADS) reinterpret_cast<uint4*>(smraw)[i] = reinterpret_cast<const uint4*>(d)[i];
        d = smraw;
    }
    if (tid == 0) sBest = 0xffffffffu;
    __syncthreads();
    const int m = (int)(0.5 * (double)(N - 1));                 // vDists[0.5*(N-1)]
    unsigned mine = 0xffffffffu;
    for (int i = tid; i < N; i += DD_THREADS) {
        uint32_t q[8];
        load_desc(q, d + (size_t)i * 32);
        int lo = 0, hi = 256;                                   // smallest t with #(dist <= t) >= m + 1
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < N; ++j) cnt += hamming256(q, reinterpret_cast<const uint32_t*>(d + (size_t)j * 32)) <= mid;
            if (cnt >= m + 1) hi = mid; else lo = mid + 1;
        }
        mine = min(mine, ((unsigned)lo << 16) | (unsigned)i);   // N <= 65535 (host check); first minimum wins
    }
    for (int s = 16; s > 0; s >>= 1) mine = min(mine, __shfl_xor_sync(FULL, mine, s));
    if ((tid & 31) == 0) atomicMin(&sBest, mine);
    __syncthreads();
    if (tid == 0) { bestIdx[p] = (int)(sBest & 0xffff); if (bestMedian) bestMedian[p] = (int)(sBest >> 16); }
}

struct BBuf {
    void* p = nullptr;
    ~BBuf() { orbx_scratch_free(p); }
    int alloc(size_t bytes) { return orbx_scratch_alloc(&p, bytes); }
    template <typename T> T* as() { return (T*)p; }
};

extern "C" {

int orbm_vocabulary_upload(const orbm_vocabulary* voc, int device, orbm_vocab_handle** out)
{
    if (!voc || !out || voc->nnodes < 2 || voc->depth < 1) { orbx_set_error("bad vocabulary"); return ORBX_E_INVALID; }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    long long nchildren = 0;
    for (int i = 0; i < voc->nnodes; ++i) {
        if (voc->child_count[i] < 0 || voc->child_begin[i] < 0) { orbx_set_error("bad child table at node %d", i); return ORBX_E_INVALID; }
        if (voc->child_count[i] > 0xffff) { orbx_set_error("more than 65535 children per node unsupported"); return ORBX_E_INVALID; }
        nchildren = std::max(nchildren, (long long)voc->child_begin[i] + voc->child_count[i]);
    }
    if (voc->child_count[0] == 0) { orbx_set_error("the root has no children"); return ORBX_E_INVALID; }
    orbm_vocab_handle* v = new orbm_vocab_handle();
    v->device = device; v->nnodes = voc->nnodes; v->depth = voc->depth;
    const size_t N = (size_t)voc->nnodes;
    ORBX_CUDA(cudaMalloc((void**)&v->childBegin, N * 4)); ORBX_CUDA(cudaMalloc((void**)&v->childCount, N * 4));
    ORBX_CUDA(cudaMalloc((void**)&v->children, std::max<size_t>(nchildren, 1) * 4));
    ORBX_CUDA(cudaMalloc((void**)&v->desc, N * 32)); ORBX_CUDA(cudaMalloc((void**)&v->word, N * 4)); ORBX_CUDA(cudaMalloc((void**)&v->weight, N * 8));
    ORBX_CUDA(cudaMemcpy(v->childBegin, voc->child_begin, N * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->childCount, voc->child_count, N * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->children, voc->children, (size_t)nchildren * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->desc, voc->descriptors, N * 32, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->word, voc->word_id, N * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->weight, voc->weight, N * 8, cudaMemcpyHostToDevice));
    *out = v;
    return ORBX_OK;
}

int orbm_vocabulary_free(orbm_vocab_handle* v)
{
    if (!v) return ORBX_OK;
    cudaSetDevice(v->device);
    cudaFree(v->childBegin); cudaFree(v->childCount); cudaFree(v->children); cudaFree(v->desc); cudaFree(v->word); cudaFree(v->weight);
    delete v;
    return ORBX_OK;
}

int orbm_bow_transform(orbm_vocab_handle* v, const uint8_t* features, int n, int levelsup, uint32_t* word_id,
                       double* weight, uint32_t* node_id)
{
    if (!v || !features || !word_id || !weight || !node_id) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (n <= 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(v->device));
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("bow_transform", acc, &calls); }
    const int sF = ar.in(features, (size_t)n * 32);
    const int sWt = ar.out(weight, (size_t)n * 8), sW = ar.out(word_id, (size_t)n * 4), sN = ar.out(node_id, (size_t)n * 4);
    int rc = ar.commit(); if (rc) return rc;
    bow_transform_kernel<<<(n + 7) / 8, 256, 0, ar.st>>>(v->childBegin, v->childCount, v->children, v->desc, v->word, v->weight, v->depth,
                                               ar.dev<u8>(sF), n, levelsup, ar.dev<uint32_t>(sW), ar.dev<double>(sWt), ar.dev<uint32_t>(sN));
    return ar.finish();
}

int orbm_search_by_bow(int variant, const orbm_bow_frame* a, const orbm_bow_frame* b, int npairs, float nnratio,
                       int check_orientation, int device, int32_t* match_a, int32_t* match_b, int32_t* nmatches)
{
    ORBX_NVTX("orbm_search_by_bow");
    if (!a || !b || !match_a || !match_b || !nmatches) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (variant != 0 && variant != 1) { orbx_set_error("variant must be 0 (KeyFrame, Frame) or 1 (KeyFrame, KeyFrame)"); return ORBX_E_INVALID; }
    if (npairs <= 0) return ORBX_OK;
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    std::vector<BowPair> pairs(npairs);
    long long totA = 0, totB = 0;
    int maxA = 1, maxB = 1;
    for (int p = 0; p < npairs; ++p) {
        if (a[p].n < 0 || b[p].n < 0 || a[p].n > 0xffff || b[p].n > 0xffff) { orbx_set_error("pair %d: 0..65535 features per frame", p); return ORBX_E_INVALID; }
        pairs[p] = BowPair{(int)totA, a[p].n, (int)totB, b[p].n};
        totA += a[p].n; totB += b[p].n;
        maxA = std::max(maxA, a[p].n); maxB = std::max(maxB, b[p].n);
    }
    int capA = 1; while (capA < maxA) capA <<= 1;
    int capB = 1; while (capB < maxB) capB <<= 1;
    size_t smem = (size_t)(capA + capB) * 8 + (size_t)capA * 2 + ((capB + 15) & ~15) + 16;
    const size_t stageBytes = (size_t)(maxA + maxB) * 37 + 64;
    const int stage = npairs <= 16 && smem + stageBytes <= 200 * 1024;      // per-frame calls; a batch keeps its occupancy instead
    if (stage) smem += stageBytes;
    if (smem > 220 * 1024) { orbx_set_error("frames of %d / %d features exceed the shared-memory sort (%zu bytes)", maxA, maxB, smem); return ORBX_E_CAPACITY; }
    for (int p = 0; p < npairs; ++p) {
        if (pairs[p].nA && (!a[p].descriptors || !a[p].angles || !a[p].node_id || !a[p].flags)) { orbx_set_error("pair %d: null array in a", p); return ORBX_E_INVALID; }
        if (pairs[p].nB && (!b[p].descriptors || !b[p].angles || !b[p].node_id || !b[p].flags)) { orbx_set_error("pair %d: null array in b", p); return ORBX_E_INVALID; }
    }
    // one staged upload, one download (CallArena): the pairs' arrays are concatenated in the staging buffer
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("search_by_bow", acc, &calls); }
    const int sPairs = ar.in(pairs.data(), npairs * sizeof(BowPair));
    const int sDescA = ar.in(nullptr, (size_t)totA * 32), sAngA = ar.in(nullptr, (size_t)totA * 4), sNodeA = ar.in(nullptr, (size_t)totA * 4), sFlagA = ar.in(nullptr, (size_t)totA);
    const int sDescB = ar.in(nullptr, (size_t)totB * 32), sAngB = ar.in(nullptr, (size_t)totB * 4), sNodeB = ar.in(nullptr, (size_t)totB * 4), sFlagB = ar.in(nullptr, (size_t)totB);
    // outputs, bins and pair state start as 0xff bytes (see the kernel's header): they are in/out segments filled in the staging
    // buffer, so the preset rides on the call's one upload instead of a memset node
    const int sMA = ar.inout_fill(match_a, (size_t)totA * 4), sMB = ar.inout_fill(match_b, (size_t)totB * 4), sNM = ar.inout_fill(nmatches, (size_t)npairs * 4);
    const int sBin = ar.inout_fill(nullptr, (size_t)totA), sState = ar.inout_fill(nullptr, (size_t)npairs * BOW_STATE_INTS * 4);
    int rc = ar.layout(); if (rc) return rc;
    for (int p = 0; p < npairs; ++p) {
        const BowPair& P = pairs[p];
        if (P.nA) {
            memcpy(ar.hostp<u8>(sDescA) + (size_t)P.offA * 32, a[p].descriptors, (size_t)P.nA * 32);
            memcpy(ar.hostp<float>(sAngA) + P.offA, a[p].angles, (size_t)P.nA * 4);
            memcpy(ar.hostp<uint32_t>(sNodeA) + P.offA, a[p].node_id, (size_t)P.nA * 4);
            memcpy(ar.hostp<u8>(sFlagA) + P.offA, a[p].flags, (size_t)P.nA);
        }
        if (P.nB) {
            memcpy(ar.hostp<u8>(sDescB) + (size_t)P.offB * 32, b[p].descriptors, (size_t)P.nB * 32);
            memcpy(ar.hostp<float>(sAngB) + P.offB, b[p].angles, (size_t)P.nB * 4);
            memcpy(ar.hostp<uint32_t>(sNodeB) + P.offB, b[p].node_id, (size_t)P.nB * 4);
            memcpy(ar.hostp<u8>(sFlagB) + P.offB, b[p].flags, (size_t)P.nB);
        }
    }
    for (int sg : {sMA, sMB, sNM, sBin, sState}) memset(ar.hostp<u8>(sg), 0xff, ar.segs[sg].bytes);
    rc = ar.upload(); if (rc) return rc;
    static bool attrSet = false;             // once: the largest size the host checks above allow
    if (!attrSet) { ORBX_CUDA(cudaFuncSetAttribute(search_bow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024)); attrSet = true; }
    // a batch fills the GPU with one CTA per pair; the single pair of a per-frame call is spread over up to 8 SMs (32 warps each)
    const int splits = npairs >= 64 ? 1 : std::max(1, std::min(8, std::min(128 / npairs, (maxA + 255) / 256)));
    search_bow_kernel<<<dim3(npairs, splits), BOW_THREADS, smem, ar.st>>>(variant, ar.dev<BowPair>(sPairs), ar.dev<u8>(sDescA), ar.dev<float>(sAngA), ar.dev<uint32_t>(sNodeA),
                                                     ar.dev<u8>(sFlagA), ar.dev<u8>(sDescB), ar.dev<float>(sAngB), ar.dev<uint32_t>(sNodeB), ar.dev<u8>(sFlagB),
                                                     nnratio, check_orientation, capA, capB, ar.dev<int32_t>(sMA), ar.dev<int32_t>(sMB), ar.dev<int32_t>(sNM),
                                                     ar.dev<u8>(sBin), ar.dev<int>(sState), stage);
    return ar.finish();
}

int orbm_distinctive_descriptors(const uint8_t* descriptors, const int32_t* offsets, int npoints, int device,
                                 int32_t* best_idx, int32_t* best_median)
{
    if (!offsets || !best_idx) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (npoints <= 0) return ORBX_OK;
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    int maxN = 0;
    for (int p = 0; p < npoints; ++p) {
        const int n = offsets[p + 1] - offsets[p];
        if (n < 0 || n > 0xffff) { orbx_set_error("map point %d: 0..65535 observations", p); return ORBX_E_INVALID; }
        maxN = std::max(maxN, n);
    }
    const size_t total = (size_t)(offsets[npoints] - offsets[0]);
    if (total && !descriptors) { orbx_set_error("null descriptors"); return ORBX_E_INVALID; }
    std::vector<int32_t> rel(npoints + 1);
    for (int p = 0; p <= npoints; ++p) rel[p] = offsets[p] - offsets[0];
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("distinctive_descriptors", acc, &calls); }
    const int sDesc = ar.in(total ? descriptors + (size_t)offsets[0] * 32 : nullptr, total * 32), sOff = ar.in(rel.data(), (size_t)(npoints + 1) * 4);
    const int sIdx = ar.out(best_idx, (size_t)npoints * 4), sMed = ar.out(best_median, (size_t)npoints * 4);
    int rc = ar.commit(); if (rc) return rc;
    const int smem = std::min(maxN, DD_SMEM_DESC) * 32 + 16;
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(distinctive_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    distinctive_kernel<<<npoints, DD_THREADS, smem, ar.st>>>(ar.dev<u8>(sDesc), ar.dev<int32_t>(sOff), ar.dev<int32_t>(sIdx), ar.dev<int32_t>(sMed));
    return ar.finish();
}

} // extern "C"
