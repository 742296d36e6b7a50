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
// The stateless per-frame calls (one frame of <= 4096 keypoints) take the LATENCY forms K7a/K7b and K9a/K9b further down instead:
// grid-free keypoint sweeps that keep a sorted 32-key candidate list per query on all SMs, then ONE CTA that iterates the
// reference's sequential rule for all queries at once until nothing changes (the fixed point is the sequential result).

#include "orbx_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "orbx_match_common.cuh"

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

// =============================================================================================
// K7: SearchForInitialization, one CTA (4 warps) per frame pair (slotA -> slotB).
// F2's keypoint positions, descriptors and grid are staged in shared memory once.  Queries are
// processed in blocks of SI_QBLOCK:
//   phase A (parallel, warp per query): enumerate the window's candidates, evaluate
//     DescriptorDistance, keep (dist << 16 | grid position) for dist <= Dkeep.  The grid position
//     (index into Frame::mGrid flattened cell-major) IS the reference's scan order, and a dropped
//     candidate (dist > Dkeep >= TH_LOW / nnratio) can neither win nor fail the ratio test.
//   phase B (warp 0, sequential in i1 like the reference): apply the vMatchedDistance skip
//     (src/ORBmatcher.cc:444), take best / second best, steal-back bookkeeping (:463-471).
// =============================================================================================
#define SI_WARPS 8
#define SI_QBLOCK 32
#define SI_MAXC 32
#define SI_DESC_SMEM 256            // F2 descriptors staged in shared memory (the octave-0 ones come first)
#define SI_PEND 64                  // per-warp buffer of window hits waiting for their Hamming distance

// The searches only ever return octave-0 keypoints of F2 (level window [0,0], src/ORBmatcher.cc:425), about a
// fifth of the frame.  They are compacted once per pair into the list L, in the grid's enumeration order (cell-major,
// index-minor = the order Frame::GetFeaturesInArea walks), each with its grid cell.  A query then tests all of L
// with full lanes — cell inside the query's cell window AND |dx| < r AND |dy| < r, exactly the two filters of
// GetFeaturesInArea — instead of walking the grid with one lane per column (6 of 32 lanes busy, 2.5 in the Hamming
// code); the hits are compacted by ballot and their Hamming distances computed 32 at a time.  The list position
// is the tie-break key (it is monotonic in the enumeration order).
struct SiSmem {
    float* lx; float* ly;            // L: positions
    unsigned short* lidx;            // L: keypoint index in F2
    unsigned short* lcell;           // L: cell x | cell y << 8
    unsigned short* md; unsigned short* m21;
    unsigned short* qlist;           // octave-0 keypoints of F1 in index order
    unsigned short* mbi;             // per F1 keypoint: the F2 keypoint it took when its turn came (0xffff = none)
    u8* binOf;                       // per F1 keypoint
    uint32_t* cand;                  // [2][SI_QBLOCK][SI_MAXC]
    int* ccount;                     // [2][SI_QBLOCK] (> SI_MAXC = overflow)
    int* hist;
    unsigned short* pend;            // [SI_WARPS][SI_PEND]
    u8* desc;                        // first SI_DESC_SMEM descriptors of F2
};

__device__ __forceinline__ SiSmem si_carve(unsigned char* base, int cap)
{
    SiSmem s;
    unsigned char* p = base;
    const size_t h = ((size_t)cap + 1) / 2 * 4;      // bytes of a u16[cap] array, 4-byte granular
    s.desc = p; p += (size_t)min(cap, SI_DESC_SMEM) * 32;
    s.lx = reinterpret_cast<float*>(p); p += (size_t)cap * 4;
    s.ly = reinterpret_cast<float*>(p); p += (size_t)cap * 4;
    s.cand = reinterpret_cast<uint32_t*>(p); p += 2 * SI_QBLOCK * SI_MAXC * 4;
    s.ccount = reinterpret_cast<int*>(p); p += 2 * SI_QBLOCK * 4;
    s.hist = reinterpret_cast<int*>(p); p += 32 * 4;
    s.pend = reinterpret_cast<unsigned short*>(p); p += SI_WARPS * SI_PEND * 2;
    s.lidx = reinterpret_cast<unsigned short*>(p); p += h;
    s.lcell = reinterpret_cast<unsigned short*>(p); p += h;
    s.md = reinterpret_cast<unsigned short*>(p); p += h;
    s.m21 = reinterpret_cast<unsigned short*>(p); p += h;
    s.qlist = reinterpret_cast<unsigned short*>(p); p += h;
    s.mbi = reinterpret_cast<unsigned short*>(p); p += h;
    s.binOf = p;
    return s;
}
static size_t si_smem_bytes(int cap)
{
    return (size_t)std::min(cap, SI_DESC_SMEM) * 32 + (size_t)cap * 8 + 2 * SI_QBLOCK * SI_MAXC * 4 + 2 * SI_QBLOCK * 4 + 128 +
           SI_WARPS * SI_PEND * 2 + 6 * (((size_t)cap + 1) / 2 * 4) + ((size_t)cap + 3) / 4 * 4 + 16;
}

struct SiCtx {
    SiSmem S;
    const orbx_keypoint* k1; const orbx_keypoint* k2;
    const u8* d1; const u8* d2g;
    const float* prev;
    GridParams gp;
    float r;
    int dkeep, ndescSmem, nL;
};

__device__ __forceinline__ const uint32_t* si_desc2(const SiCtx& c, int idx)
{
    return reinterpret_cast<const uint32_t*>(idx < c.ndescSmem ? c.S.desc + (size_t)idx * 32 : c.d2g + (size_t)idx * 32);
}

// is list entry p returned by GetFeaturesInArea(x, y, r, 0, 0)?  (cell window first, then the distance test)
__device__ __forceinline__ bool si_hit(const SiSmem& S, int p, float x, float y, float r, int cx0, int cx1, int cy0, int cy1)
{
    const int cell = S.lcell[p], cx = cell & 0xff, cy = cell >> 8;
    return cx >= cx0 && cx <= cx1 && cy >= cy0 && cy <= cy1 &&
           fabsf(__fsub_rn(S.lx[p], x)) < r && fabsf(__fsub_rn(S.ly[p], y)) < r;
}

// phase A for the queries qlist[q0 .. q0+nq) into candidate buffer `buf`, warps wfirst..SI_WARPS-1
__device__ __forceinline__ void si_phase_a(const SiCtx& c, int q0, int nq, int buf, int wfirst, int wid, int lane,
                                           unsigned long long& ncmp)
{
    const SiSmem& S = c.S;
    unsigned short* pend = S.pend + wid * SI_PEND;
    for (int qi = wid - wfirst; qi < nq; qi += SI_WARPS - wfirst) {
        const int i1 = S.qlist[q0 + qi];
        const float x = c.prev ? c.prev[2 * i1] : c.k1[i1].x, y = c.prev ? c.prev[2 * i1 + 1] : c.k1[i1].y;
        int cx0, cx1, cy0, cy1;
        int cnt = 0;
        if (area_cells(c.gp, x, y, c.r, cx0, cx1, cy0, cy1)) {
            uint32_t q[8];
            load_desc(q, c.d1 + (size_t)i1 * 32);
            uint32_t* out = S.cand + (buf * SI_QBLOCK + qi) * SI_MAXC;
            // Hamming distances of the first n pending hits, one per lane
            auto flush = [&](int n) {
                uint32_t ent = 0xffffffffu;
                if (lane < n) {
                    const int p = pend[lane];
                    const int dist = hamming256(q, si_desc2(c, S.lidx[p]));
                    ++ncmp;
                    if (dist <= c.dkeep) ent = ((uint32_t)dist << 16) | (uint32_t)p;
                }
                const unsigned m = __ballot_sync(FULL, ent != 0xffffffffu);
                if (ent != 0xffffffffu) {
                    const int o = cnt + __popc(m & ((1u << lane) - 1));
                    if (o < SI_MAXC) out[o] = ent;
                }
                cnt += __popc(m);
            };
            int npend = 0;
            for (int p0 = 0; p0 < c.nL; p0 += 32) {
                const int p = p0 + lane;
                const bool hit = p < c.nL && si_hit(S, p, x, y, c.r, cx0, cx1, cy0, cy1);
                const unsigned m = __ballot_sync(FULL, hit);
                if (hit) pend[npend + __popc(m & ((1u << lane) - 1))] = (unsigned short)p;
                npend += __popc(m);
                __syncwarp();
                if (npend >= 32) {
                    flush(32);
                    __syncwarp();
                    const int rest = npend - 32;
                    const unsigned short mv = lane < rest ? pend[32 + lane] : (unsigned short)0;
                    __syncwarp();
                    if (lane < rest) pend[lane] = mv;
                    npend = rest;
                    __syncwarp();
                }
            }
            if (npend > 0) flush(npend);
            __syncwarp();
        }
        if (lane == 0) S.ccount[buf * SI_QBLOCK + qi] = cnt;
    }
}

__global__ void __launch_bounds__(SI_WARPS * 32, 5)
search_init_kernel(FrameStore fs, int firstSlot, int slotStride, GridParams gp, float* prevMatchedBase,
                   int windowSize, float nnratio, int checkOri, int* __restrict__ matches12Base,
                   int* __restrict__ nmatchesOut, unsigned long long* __restrict__ ncompares, int dkeep)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    __shared__ int sNq, sCull, sNL, sWq[SI_WARPS], sWl[SI_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int pair = blockIdx.x;
    const int sA = firstSlot + pair * slotStride, sB = sA + 1;
    const int n1 = min(fs.count[sA], fs.cap), n2 = min(fs.count[sB], fs.cap);
    SiCtx c;
    c.S = si_carve(smraw, fs.cap);
    const SiSmem& S = c.S;
    c.k1 = fs.kps + (size_t)sA * fs.cap; c.k2 = fs.kps + (size_t)sB * fs.cap;
    c.d1 = fs.desc + (size_t)sA * fs.cap * 32; c.d2g = fs.desc + (size_t)sB * fs.cap * 32;
    c.prev = prevMatchedBase ? prevMatchedBase + (size_t)pair * fs.cap * 2 : nullptr;
    c.gp = gp; c.r = (float)windowSize; c.dkeep = dkeep; c.ndescSmem = min(n2, SI_DESC_SMEM);
    const unsigned short* gsg = fs.gridStart + (size_t)sB * (ORBX_GRID_CELLS + 1);
    const unsigned short* gig = fs.gridIdx + (size_t)sB * fs.cap;
    int* m12 = matches12Base + (size_t)pair * fs.cap;
    // ---- stage F2 -------------------------------------------------------------------------------
    for (int i = tid; i < n2; i += SI_WARPS * 32) { S.md[i] = 0xffff; S.m21[i] = 0xffff; }
    for (int i = tid; i < c.ndescSmem * 2; i += SI_WARPS * 32)
        reinterpret_cast<uint4*>(S.desc)[i] = reinterpret_cast<const uint4*>(c.d2g)[i];
    for (int i = tid; i < n1; i += SI_WARPS * 32) { S.binOf[i] = 0xff; S.mbi[i] = 0xffff; m12[i] = -1; }
    if (tid < ORBX_HISTO_LENGTH) S.hist[tid] = 0;
    if (tid == 0) sCull = 0;
    // Two ordered compactions, spread over all warps (each warp owns a contiguous range; counts, then a prefix over
    // the warps, then the writes):
    //   qlist: octave-0 keypoints of F1 (src/ORBmatcher.cc:421-423), ascending index
    //   L:     octave-0 keypoints of F2 that are in the grid, in enumeration order, with their cell
    const int nIn = min((int)gsg[ORBX_GRID_CELLS], n2);
    const int qr = ((n1 + SI_WARPS - 1) / SI_WARPS + 31) & ~31, lr = ((nIn + SI_WARPS - 1) / SI_WARPS + 31) & ~31;
    const int qa = min(wid * qr, n1), qb = min(qa + qr, n1), la = min(wid * lr, nIn), lb = min(la + lr, nIn);
    {
        int cq = 0, cl = 0;
        for (int i0 = qa; i0 < qb; i0 += 32) { const int i = i0 + lane; cq += __popc(__ballot_sync(FULL, i < qb && c.k1[i].octave <= 0)); }
        for (int j0 = la; j0 < lb; j0 += 32) { const int j = j0 + lane; cl += __popc(__ballot_sync(FULL, j < lb && c.k2[gig[j]].octave <= 0)); }
        if (lane == 0) { sWq[wid] = cq; sWl[wid] = cl; }
    }
    __syncthreads();
    {
        int nq = 0, nl = 0;
        for (int w = 0; w < wid; ++w) { nq += sWq[w]; nl += sWl[w]; }
        if (wid == SI_WARPS - 1 && lane == 0) { sNq = nq + sWq[wid]; sNL = nl + sWl[wid]; }
        for (int i0 = qa; i0 < qb; i0 += 32) {
            const int i = i0 + lane;
            const bool isq = i < qb && c.k1[i].octave <= 0;
            const unsigned m = __ballot_sync(FULL, isq);
            if (isq) S.qlist[nq + __popc(m & ((1u << lane) - 1))] = (unsigned short)i;
            nq += __popc(m);
        }
        for (int j0 = la; j0 < lb; j0 += 32) {
            const int j = j0 + lane;
            bool keep = false;
            int idx = 0;
            orbx_keypoint kp;
            if (j < lb) { idx = gig[j]; kp = c.k2[idx]; keep = kp.octave <= 0; }
            const unsigned m = __ballot_sync(FULL, keep);
            if (keep) {
                const int o = nl + __popc(m & ((1u << lane) - 1));
                // Frame::PosInGrid (src/Frame.cc:382-392): the cell the grid build filed this keypoint under
                const int px = (int)roundf(__fmul_rn(__fsub_rn(kp.x, gp.minX), gp.invW));
                const int py = (int)roundf(__fmul_rn(__fsub_rn(kp.y, gp.minY), gp.invH));
                S.lx[o] = kp.x; S.ly[o] = kp.y;
                S.lidx[o] = (unsigned short)idx;
                S.lcell[o] = (unsigned short)(px | (py << 8));
            }
            nl += __popc(m);
        }
    }
    __syncthreads();
    const int nqTotal = sNq;
    c.nL = sNL;
    int nmatches = 0;                      // meaningful in warp 0
    unsigned long long ncmp = 0;
    si_phase_a(c, 0, min(SI_QBLOCK, nqTotal), 0, 0, wid, lane, ncmp);
    for (int q0 = 0, buf = 0; q0 < nqTotal; q0 += SI_QBLOCK, buf ^= 1) {
        __syncthreads();
        const int nq = min(SI_QBLOCK, nqTotal - q0);
        if (wid != 0) {
            // warps 1.. : candidates of the NEXT block while warp 0 replays this one
            if (q0 + SI_QBLOCK < nqTotal) si_phase_a(c, q0 + SI_QBLOCK, min(SI_QBLOCK, nqTotal - q0 - SI_QBLOCK), buf ^ 1, 1, wid, lane, ncmp);
        } else {
            // ---- phase B: replay in i1 order ---------------------------------------------------------
            for (int qi = 0; qi < nq; ++qi) {
                const int i1 = S.qlist[q0 + qi];
                const int cnt = S.ccount[buf * SI_QBLOCK + qi];
                int best = 0x7fffffff, second = 0x7fffffff;
                unsigned key = 0xffffffffu;
                if (cnt <= SI_MAXC) {
                    if (lane < cnt) {
                        const uint32_t ent = S.cand[(buf * SI_QBLOCK + qi) * SI_MAXC + lane];
                        const int dist = (int)(ent >> 16), idx = S.lidx[ent & 0xffff];
                        if ((int)S.md[idx] > dist) { best = dist; key = ent & 0xffff; }
                    }
                } else {
                    // overflow: enumerate again, filtering on the fly (rare)
                    const float x = c.prev ? c.prev[2 * i1] : c.k1[i1].x, y = c.prev ? c.prev[2 * i1 + 1] : c.k1[i1].y;
                    int cx0, cx1, cy0, cy1;
                    area_cells(gp, x, y, c.r, cx0, cx1, cy0, cy1);
                    uint32_t q[8];
                    load_desc(q, c.d1 + (size_t)i1 * 32);
                    for (int p = lane; p < c.nL; p += 32) {
                        if (!si_hit(S, p, x, y, c.r, cx0, cx1, cy0, cy1)) continue;
                        const int idx = S.lidx[p];
                        const int dist = hamming256(q, si_desc2(c, idx));
                        if ((int)S.md[idx] <= dist) continue;
                        if (dist < best || (dist == best && (unsigned)p < key)) { second = min(second, best); best = dist; key = (unsigned)p; }
                        else second = min(second, dist);
                    }
                }
                warp_top2(best, key, second);
                if (best <= ORBX_TH_LOW && (float)best < __fmul_rn((float)second, nnratio)) {
                    const int bi = S.lidx[key];
                    const int old = S.m21[bi];
                    __syncwarp();
                    if (lane == 0) {
                        if (old != 0xffff) m12[old] = -1;
                        m12[i1] = bi;
                        S.m21[bi] = (unsigned short)i1;
                        S.md[bi] = (unsigned short)best;
                        S.mbi[i1] = (unsigned short)bi;          // its rotation bin is worked out after the replay (two dependent global loads)
                    }
                    if (old != 0xffff) --nmatches;
                    ++nmatches;
                    __syncwarp();
                }
            }
        }
    }
    __threadfence_block();
    __syncthreads();
    if (checkOri) {
        // rotHist[bin].push_back(i1) happened for every accepted query, also for the ones a later query displaced (src/ORBmatcher.cc:470-482)
        for (int i = tid; i < n1; i += SI_WARPS * 32) {
            const int bi = S.mbi[i];
            if (bi != 0xffff) { const int bin = rot_bin(c.k1[i].angle, c.k2[bi].angle); S.binOf[i] = (u8)bin; atomicAdd(&S.hist[bin], 1); }
        }
        __syncthreads();
        int i1a, i2a, i3a;
        three_maxima(S.hist, ORBX_HISTO_LENGTH, i1a, i2a, i3a);
        int culled = 0;
        for (int i = tid; i < n1; i += SI_WARPS * 32) {
            const int b = S.binOf[i];
            if (b != 0xff && b != i1a && b != i2a && b != i3a && m12[i] >= 0) { m12[i] = -1; ++culled; }
        }
        for (int d = 16; d > 0; d >>= 1) culled += __shfl_xor_sync(FULL, culled, d);
        if (lane == 0 && culled) atomicAdd(&sCull, culled);
    }
    __syncthreads();
    if (prevMatchedBase) {
        float* pw = prevMatchedBase + (size_t)pair * fs.cap * 2;
        for (int i = tid; i < n1; i += SI_WARPS * 32) {
            const int m = m12[i];
            if (m >= 0) { pw[2 * i] = c.k2[m].x; pw[2 * i + 1] = c.k2[m].y; }
        }
    }
    for (int d = 16; d > 0; d >>= 1) ncmp += __shfl_xor_sync(FULL, ncmp, d);
    if (lane == 0 && ncompares && ncmp) atomicAdd(ncompares, ncmp);
    if (tid == 0) nmatchesOut[pair] = nmatches - sCull;
}

// =============================================================================================
// K9: the four ORBmatcher::SearchByProjection variants on flattened queries.  The reference loops
// over map points in order and a match changes what later queries may take (occupied flags), so
// one warp replays the query order; the candidates of one query (grid cells of
// Frame::GetFeaturesInArea) are spread over the lanes.  "First minimum / second minimum in scan
// order" (strict '<' updates, src/ORBmatcher.cc:100-114) = the two smallest 64-bit keys
// dist | cell | idx | octave, because the scan order is (cell, idx) ascending.
// =============================================================================================
__device__ __forceinline__ void top2_insert(unsigned long long& k1, unsigned long long& k2, unsigned long long k)
{
    if (k < k1) { k2 = k1; k1 = k; }
    else if (k < k2) k2 = k;
}

// The PROJ_K smallest candidate keys of one query, ascending (warp-cooperative: lanes over the grid cells of
// Frame::GetFeaturesInArea, each lane keeps its own PROJ_K smallest, then PROJ_K rounds of warp-minimum extraction;
// every lane returns the same keys).  LIVE_OCC: read the occupancy flags as they are now (replay) rather than as
// read-only data (candidate phase).
#define PROJ_K 6
struct ProjCtx {
    const orbx_keypoint* kps; const u8* desc; const unsigned short* gs; const unsigned short* gi;
    const float* uRight; const u8* occupied; GridParams gp; bool gate;
};
template <bool LIVE_OCC>
__device__ __forceinline__ void proj_topk(const ProjCtx& c, const orbm_proj_query& Q, const u8* qdescRow, int lane,
                                          unsigned long long (&out)[PROJ_K])
{
    const unsigned long long NONE = ~0ull;
#pragma unroll
    for (int k = 0; k < PROJ_K; ++k) out[k] = NONE;
    int cx0, cx1, cy0, cy1;
    if (!area_cells(c.gp, Q.u, Q.v, Q.radius, cx0, cx1, cy0, cy1)) return;
    unsigned long long loc[PROJ_K];
#pragma unroll
    for (int k = 0; k < PROJ_K; ++k) loc[k] = NONE;
    uint32_t qd[8];
    load_desc(qd, qdescRow);
    const bool checkLevels = (Q.min_level > 0) || (Q.max_level >= 0);
    const int ncy = cy1 - cy0 + 1, nc = (cx1 - cx0 + 1) * ncy;
    for (int cc = lane; cc < nc; cc += 32) {
        const int ix = cx0 + cc / ncy, iy = cy0 + cc % ncy;
        const int cell = ix * ORBX_GRID_ROWS + iy;
        const int s = c.gs[cell], e = c.gs[cell + 1];
        for (int j = s; j < e; ++j) {
            const int idx = c.gi[j];
            const orbx_keypoint kp = c.kps[idx];
            if (checkLevels) {
                if (kp.octave < Q.min_level) continue;
                if (Q.max_level >= 0 && kp.octave > Q.max_level) continue;
            }
            if (!(fabsf(__fsub_rn(kp.x, Q.u)) < Q.radius && fabsf(__fsub_rn(kp.y, Q.v)) < Q.radius)) continue;
            if (LIVE_OCC ? *reinterpret_cast<const volatile u8*>(c.occupied + idx) : c.occupied[idx]) continue;
            if (c.gate) {
                const float ur = c.uRight[idx];
                if (ur > 0 && Q.max_right_err >= 0 && fabsf(__fsub_rn(Q.u_right, ur)) > Q.max_right_err) continue;
            }
            if (Q.level_lo >= 0 && (kp.octave < Q.level_lo || kp.octave > Q.level_hi)) continue;
            const int dist = hamming256(qd, reinterpret_cast<const uint32_t*>(c.desc + (size_t)idx * 32));
            unsigned long long key = ((unsigned long long)dist << 40) | ((unsigned long long)cell << 24) |
                                     ((unsigned long long)idx << 8) | (unsigned long long)(kp.octave & 0xff);
#pragma unroll
            for (int k = 0; k < PROJ_K; ++k)             // sorted insertion (keys are distinct: idx is part of the key)
                if (key < loc[k]) { const unsigned long long t = loc[k]; loc[k] = key; key = t; }
        }
    }
#pragma unroll
    for (int k = 0; k < PROJ_K; ++k) {
        unsigned long long m = loc[0];
        for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_xor_sync(FULL, m, d); m = o < m ? o : m; }
        out[k] = m;
        if (m != NONE && loc[0] == m) {                  // the owner pops its head
#pragma unroll
            for (int j = 0; j + 1 < PROJ_K; ++j) loc[j] = loc[j + 1];
            loc[PROJ_K - 1] = NONE;
        }
    }
}

// Phase 1 (all SMs, warp per query): the PROJ_K smallest keys of every query under the occupancy the call STARTS with.
#define PROJ_WARPS 4
#define PROJ_CLAIMS 1024
__global__ void __launch_bounds__(PROJ_WARPS * 32)
search_proj_cand_kernel(FrameStore fs, int slot, GridParams gp, const float* __restrict__ uRight,
                        const orbm_proj_query* __restrict__ queries, const u8* __restrict__ qdesc, int nq,
                        const u8* __restrict__ occupied, int variant, unsigned long long* __restrict__ keys,
                        float* __restrict__ keyAngles)
{
    const int lane = threadIdx.x & 31, q = blockIdx.x * PROJ_WARPS + (threadIdx.x >> 5);
    if (q >= nq) return;
    ProjCtx c;
    c.kps = fs.kps + (size_t)slot * fs.cap; c.desc = fs.desc + (size_t)slot * fs.cap * 32;
    c.gs = fs.gridStart + (size_t)slot * (ORBX_GRID_CELLS + 1); c.gi = fs.gridIdx + (size_t)slot * fs.cap;
    c.uRight = uRight; c.occupied = occupied; c.gp = gp;
    c.gate = (variant == ORBM_PROJ_LOCALMAP || variant == ORBM_PROJ_LASTFRAME) && uRight != nullptr;
    const orbm_proj_query Q = queries[q];
    unsigned long long K[PROJ_K];
    proj_topk<false>(c, Q, qdesc + (size_t)q * 32, lane, K);
    if (lane < PROJ_K) {
        unsigned long long v = K[0];
#pragma unroll
        for (int k = 1; k < PROJ_K; ++k) if (lane == k) v = K[k];
        keys[(size_t)PROJ_K * q + lane] = v;
        // the candidate's orientation rides along, so the replay needs no dependent keypoint load
        keyAngles[(size_t)PROJ_K * q + lane] = v != ~0ull ? c.kps[(int)((v >> 8) & 0xffff)].angle : 0.f;
    }
}

// Phase 2 (one warp): replay of the query order.  Occupancy only grows during a call, so the candidates of query q
// under the occupancy it actually meets are its phase-1 candidates minus those taken by earlier queries of this
// call: its best and second best are the first two still-free keys of its phase-1 list.  That is exact whenever the
// list is complete (fewer than PROJ_K candidates) or still has the free keys the variant needs (two for the ratio
// test of the local-map variant, one otherwise), because everything beyond the list is larger than all of it; only
// the remaining queries are searched again (proj_topk under the current occupancy).  32 queries are loaded at a
// time, one per lane, and handed round by shuffle; the occupancy flags live in shared memory during the replay.
__global__ void __launch_bounds__(32)
search_proj_kernel(FrameStore fs, int slot, GridParams gp, const float* __restrict__ uRight,
                   const orbm_proj_query* __restrict__ queries, const u8* __restrict__ qdesc, int nq,
                   const u8* __restrict__ queryBlocks, u8* __restrict__ occupied, int* __restrict__ assigned,
                   int* __restrict__ bestIdxOut, int* __restrict__ bestDistOut, u8* __restrict__ binOf,
                   int variant, int thDist, float nnratio, int checkOri, int* __restrict__ nmatchesOut,
                   const unsigned long long* __restrict__ keys, const float* __restrict__ keyAngles)
{
    __shared__ int hist[ORBX_HISTO_LENGTH];
    __shared__ unsigned claim[PROJ_CLAIMS];
    extern __shared__ __align__(16) u8 occS[];
    const int lane = threadIdx.x;
    const int n = min(fs.count[slot], fs.cap);
    for (int i = lane; i < PROJ_CLAIMS; i += 32) claim[i] = 0xffffffffu;
    ProjCtx c;
    c.kps = fs.kps + (size_t)slot * fs.cap; c.desc = fs.desc + (size_t)slot * fs.cap * 32;
    c.gs = fs.gridStart + (size_t)slot * (ORBX_GRID_CELLS + 1); c.gi = fs.gridIdx + (size_t)slot * fs.cap;
    c.uRight = uRight; c.occupied = occS; c.gp = gp;
    c.gate = (variant == ORBM_PROJ_LOCALMAP || variant == ORBM_PROJ_LASTFRAME) && uRight != nullptr;
    const orbx_keypoint* kps = c.kps;
    if (lane < ORBX_HISTO_LENGTH) hist[lane] = 0;
    for (int i = lane; i < n; i += 32) { assigned[i] = -1; occS[i] = occupied[i]; }
    __syncwarp();
    const bool rot = checkOri && (variant == ORBM_PROJ_LASTFRAME || variant == ORBM_PROJ_KEYFRAME);
    const bool alwaysBlocks = variant == ORBM_PROJ_KEYFRAME || variant == ORBM_PROJ_SIM3;
    const int needFree = variant == ORBM_PROJ_LOCALMAP ? 2 : 1;
    const unsigned long long NONE = ~0ull;
    volatile u8* occ = occS;
    int nmatches = 0;
    // per-lane query state, loaded one batch ahead so the loads of batch b+1 fly while batch b is replayed
    unsigned long long Kn[PROJ_K];
    float An[PROJ_K];
    float angleN = 0.f;
    bool blocksN = true;
    auto load_batch = [&](int q0) {
        const int myq = q0 + lane;
#pragma unroll
        for (int k = 0; k < PROJ_K; ++k) { Kn[k] = NONE; An[k] = 0.f; }
        angleN = 0.f; blocksN = true;
        if (myq < nq) {
#pragma unroll
            for (int k = 0; k < PROJ_K; ++k) Kn[k] = keys[(size_t)PROJ_K * myq + k];
            if (rot) {
                angleN = queries[myq].angle;
#pragma unroll
                for (int k = 0; k < PROJ_K; ++k) An[k] = keyAngles[(size_t)PROJ_K * myq + k];
            }
            blocksN = alwaysBlocks ? true : (queryBlocks ? queryBlocks[myq] != 0 : true);
        }
    };
    load_batch(0);
    for (int q0 = 0; q0 < nq; q0 += 32) {
        const int myq = q0 + lane;
        unsigned long long K[PROJ_K];
        float A[PROJ_K];
#pragma unroll
        for (int k = 0; k < PROJ_K; ++k) { K[k] = Kn[k]; A[k] = An[k]; }
        const float myAngle = angleN;
        const bool myBlocks = blocksN;
        if (q0 + 32 < nq) load_batch(q0 + 32);
        int outIdx = -1, outDist = 256, outBin = 0xff;
        const int cnt = min(32, nq - q0);
        // The 32 queries of the batch are decided lane-parallel against the current occupancy, then committed in
        // order as far as that is safe: a lane may commit once every earlier lane has, or when no earlier
        // uncommitted lane takes a keypoint of its candidate list.  The first lane that cannot (or whose list ran
        // out of free keys: it is searched again by the whole warp) ends the round; the rest is decided again.
        unsigned done = cnt < 32 ? ~((1u << cnt) - 1u) : 0u;
        while (done != FULL) {
            unsigned freeMask = 0;
            bool complete = false;
#pragma unroll
            for (int k = 0; k < PROJ_K; ++k) {
                const bool valid = K[k] != NONE;
                const unsigned o = occ[valid ? (int)((K[k] >> 8) & 0xffff) : 0];
                freeMask |= (valid && !o) ? (1u << k) : 0u;
                complete |= !valid;
            }
            const int nfree = __popc(freeMask);
            const int p1 = nfree > 0 ? __ffs(freeMask) - 1 : 0;
            const unsigned rest = freeMask & (freeMask - 1);
            const int p2 = nfree > 1 ? __ffs(rest) - 1 : -1;
            unsigned long long k1 = NONE, k2 = NONE;
            float kpAngle = A[0];
#pragma unroll
            for (int k = 0; k < PROJ_K; ++k) {
                if (nfree > 0 && p1 == k) { k1 = K[k]; kpAngle = A[k]; }
                if (p2 == k) k2 = K[k];
            }
            const bool pending = !((done >> lane) & 1u);
            const bool dirty = pending && !complete && nfree < needFree;
            const unsigned pendMask = ~done;
            const int first = __ffs(pendMask) - 1;                       // pendMask != 0 here
            float qAngle = myAngle;
            bool blocks = myBlocks;
            int q = myq;
            const bool slow = (__ballot_sync(FULL, dirty) >> first) & 1u;
            if (slow) {
                // the first uncommitted query has no free key left in a truncated list: search it again (whole warp)
                q = q0 + first;
                const orbm_proj_query Q = queries[q];
                unsigned long long R[PROJ_K];
                proj_topk<true>(c, Q, qdesc + (size_t)q * 32, lane, R);
                k1 = R[0]; k2 = R[1];
                qAngle = Q.angle;
                blocks = __shfl_sync(FULL, (int)myBlocks, first) != 0;
                if (rot && k1 != NONE) kpAngle = kps[(int)((k1 >> 8) & 0xffff)].angle;
            }
            int bIdx = -1, bDist = 256;
            if (k1 != NONE) {
                const int bestDist = (int)(k1 >> 40), bestLevel = (int)(k1 & 0xff), bestIdx = (int)((k1 >> 8) & 0xffff);
                // reference initial values: bestDist2 = 256, bestLevel2 = -1 (src/ORBmatcher.cc:73-77)
                const int bestDist2 = (k2 != NONE && (int)(k2 >> 40) < 256) ? (int)(k2 >> 40) : 256;
                const int bestLevel2 = (k2 != NONE && (int)(k2 >> 40) < 256) ? (int)(k2 & 0xff) : -1;
                bool ok = bestDist < 256 && bestDist <= thDist;
                if (ok && variant == ORBM_PROJ_LOCALMAP && bestLevel == bestLevel2 &&
                    (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) ok = false;
                if (ok) { bIdx = bestIdx; bDist = bestDist; }
            }
            const int bin = (rot && bIdx >= 0) ? rot_bin(qAngle, kpAngle) : 0xff;
            unsigned commit;
            if (slow) commit = 1u << first;                              // every lane holds query `first`'s decision
            else {
                // an earlier uncommitted lane that takes a keypoint my decision depends on (or is itself undecided) stops me.  Takes are
                // posted in a small hashed claim table (lowest lane wins); a hash collision only stops a lane that
                // could have gone on, which costs a round, never exactness.
                const int take = (pending && !dirty) ? bIdx : -1;
                if (take >= 0) atomicMin(&claim[take & (PROJ_CLAIMS - 1)], (unsigned)lane);
                __syncwarp();
                const unsigned dirtyMask = __ballot_sync(FULL, dirty);
                bool stop = dirty || (dirtyMask & ((1u << lane) - 1u)) != 0;
                volatile unsigned* cl = claim;
                // only the keys my decision reads matter: the best one, and the second best where the ratio test uses it
                stop |= k1 != NONE && cl[(int)((k1 >> 8) & 0xffff) & (PROJ_CLAIMS - 1)] < (unsigned)lane;
                stop |= needFree == 2 && k2 != NONE && cl[(int)((k2 >> 8) & 0xffff) & (PROJ_CLAIMS - 1)] < (unsigned)lane;
                __syncwarp();
                if (take >= 0) cl[take & (PROJ_CLAIMS - 1)] = 0xffffffffu;
                const unsigned bad = __ballot_sync(FULL, pending && stop);
                const int f = bad ? __ffs(bad) - 1 : 32;
                commit = pendMask & (f >= 32 ? FULL : ((1u << f) - 1u));
            }
            const bool mine = slow ? (lane == first) : (((commit >> lane) & 1u) != 0);
            if (mine) {
                outIdx = bIdx; outDist = bDist; outBin = bin;
                if (bIdx >= 0) {
                    assigned[bIdx] = q;
                    occ[bIdx] = blocks ? 1 : 0;
                    if (rot) atomicAdd(&hist[bin], 1);
                }
            }
            nmatches += __popc(__ballot_sync(FULL, mine && bIdx >= 0));
            done |= commit;
            __syncwarp();
        }
        if (myq < nq) { bestIdxOut[myq] = outIdx; bestDistOut[myq] = outDist; if (rot) binOf[myq] = (u8)outBin; }
    }
    __threadfence_block();
    __syncwarp();
    for (int i = lane; i < n; i += 32) occupied[i] = occS[i];
    if (rot) {
        int i1, i2, i3;
        three_maxima(hist, ORBX_HISTO_LENGTH, i1, i2, i3);
        for (int q0 = 0; q0 < nq; q0 += 32) {
            const int q = q0 + lane;
            bool cull = false;
            if (q < nq) {
                const int b = binOf[q];
                cull = b != 0xff && b != i1 && b != i2 && b != i3;
                if (cull) assigned[bestIdxOut[q]] = -1;
            }
            nmatches -= __popc(__ballot_sync(FULL, cull));
        }
    }
    if (lane == 0) *nmatchesOut = nmatches;
}

// =============================================================================================
// K9, latency form (one frame of at most PROJ_SCAN_MAXN keypoints: the per-frame tracking call).
//  K9a  proj_list_kernel     warp per query, all SMs.  No grid: the lanes sweep the frame's keypoints (coalesced, no
//       dependent loads) with the window test AND the cell-range test of Frame::GetFeaturesInArea (a keypoint counts only
//       if the cell AssignFeaturesToGrid files it under lies in the query's cell window, src/Frame.cc:332-377), compact the
//       hits, and keep the PROJ_L = 32 smallest keys dist|cell|idx|octave sorted across the lanes (bitonic sort + merge).
//  K9b  proj_fixpoint_kernel  one CTA.  The reference's loop is "query q takes the first key of its list that no EARLIER
//       query has blocked".  Iterate that rule for all queries at once from the takes of the previous iteration (owner[k] =
//       lowest blocking query that takes keypoint k): query q is final once every earlier query is, so the prefix of final
//       queries grows by at least one per iteration and the fixed point is the sequential result; real frames converge in a
//       handful of iterations.  A query whose 32-key list is truncated and used up is searched again under the occupancy
//       it sees (same sweep, by a warp).
// =============================================================================================
#define PROJ_L 32
#define PROJ_SCAN_MAXN 4096
#define PROJ_FIX_THREADS 1024
#define PROJ_SCAN_BUF 64

struct ProjScan {
    const orbx_keypoint* kps; const u8* desc; const float* uRight; GridParams gp; int n; bool gate;
};

__device__ __forceinline__ unsigned long long min64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
__device__ __forceinline__ unsigned long long max64(unsigned long long a, unsigned long long b) { return a < b ? b : a; }

// ascending bitonic sort of one key per lane
__device__ __forceinline__ unsigned long long warp_sort32(unsigned long long v, int lane)
{
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const unsigned long long o = __shfl_xor_sync(FULL, v, j);
            const bool takeMin = ((lane & j) == 0) == ((lane & k) == 0);
            v = takeMin ? min64(o, v) : max64(o, v);
        }
    }
    return v;
}

// the 32 smallest of two ascending 32-key sequences, ascending
__device__ __forceinline__ unsigned long long warp_merge32(unsigned long long a, unsigned long long b, int lane)
{
    unsigned long long v = min64(a, __shfl_sync(FULL, b, 31 - lane));       // bitonic
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const unsigned long long o = __shfl_xor_sync(FULL, v, j);
        v = (lane & j) == 0 ? min64(o, v) : max64(o, v);
    }
    return v;
}

// Lane l returns the l-th smallest candidate key of the query (~0 where there are fewer).  isFree(idx): the occupancy
// the query sees.  buf: PROJ_SCAN_BUF entries of shared memory owned by this warp.
struct KeepAll { __device__ __forceinline__ bool operator()(int, int) const { return true; } };
// keep(idx, dist): a last filter that needs the distance (SearchForInitialization's vMatchedDistance test)
template <typename Free, typename Keep = KeepAll>
__device__ __forceinline__ unsigned long long proj_scan_sorted(const ProjScan& c, const orbm_proj_query& Q, const u8* qdescRow,
                                                                int lane, unsigned long long* buf, Free isFree, Keep keep = Keep())
{
    const unsigned long long NONE = ~0ull;
    unsigned long long top = NONE;
    int cx0, cx1, cy0, cy1;
    if (!area_cells(c.gp, Q.u, Q.v, Q.radius, cx0, cx1, cy0, cy1)) return top;
    uint32_t qd[8];
    load_desc(qd, qdescRow);
    const bool checkLevels = (Q.min_level > 0) || (Q.max_level >= 0);
    bool haveTop = false;
    int nbuf = 0;
    auto flush = [&](int cnt) {
        unsigned long long key = NONE;
        if (lane < cnt) {
            const unsigned long long part = buf[lane];
            const int idx = (int)((part >> 8) & 0xffff);
            const int dist = hamming256(qd, reinterpret_cast<const uint32_t*>(c.desc + (size_t)idx * 32));
            if (keep(idx, dist)) key = ((unsigned long long)dist << 40) | part;
        }
        key = warp_sort32(key, lane);
        top = haveTop ? warp_merge32(top, key, lane) : key;
        haveTop = true;
    };
    for (int i0 = 0; i0 < c.n; i0 += 32) {
        const int i = i0 + lane;
        bool cand = false;
        unsigned long long part = 0;
        if (i < c.n) {
            const float kx = c.kps[i].x, ky = c.kps[i].y;
            if (fabsf(__fsub_rn(kx, Q.u)) < Q.radius && fabsf(__fsub_rn(ky, Q.v)) < Q.radius) {
                const int oct = c.kps[i].octave;
                const int px = (int)roundf(__fmul_rn(__fsub_rn(kx, c.gp.minX), c.gp.invW));
                const int py = (int)roundf(__fmul_rn(__fsub_rn(ky, c.gp.minY), c.gp.invH));
                cand = px >= cx0 && px <= cx1 && py >= cy0 && py <= cy1;       // the cell window lies inside the grid
                if (checkLevels && (oct < Q.min_level || (Q.max_level >= 0 && oct > Q.max_level))) cand = false;
                if (Q.level_lo >= 0 && (oct < Q.level_lo || oct > Q.level_hi)) cand = false;
                if (cand && c.gate) {
                    const float ur = c.uRight[i];
                    if (ur > 0 && Q.max_right_err >= 0 && fabsf(__fsub_rn(Q.u_right, ur)) > Q.max_right_err) cand = false;
                }
                if (cand && !isFree(i)) cand = false;
                part = ((unsigned long long)(px * ORBX_GRID_ROWS + py) << 24) | ((unsigned long long)i << 8) | (unsigned long long)(oct & 0xff);
            }
        }
        const unsigned m = __ballot_sync(FULL, cand);
        if (m) {
            if (cand) buf[nbuf + __popc(m & ((1u << lane) - 1u))] = part;
            nbuf += __popc(m);
            __syncwarp();
            if (nbuf >= 32) {
                flush(32);
                const unsigned long long carry = (32 + lane < nbuf) ? buf[32 + lane] : 0ull;
                __syncwarp();
                buf[lane] = carry;
                nbuf -= 32;
                __syncwarp();
            }
        }
    }
    if (nbuf > 0) flush(nbuf);
    return top;
}

#define PROJ_LIST_WARPS 4
__global__ void __launch_bounds__(PROJ_LIST_WARPS * 32)
proj_list_kernel(ProjScan c, const orbm_proj_query* __restrict__ queries, const u8* __restrict__ qdesc, int nq,
                 const u8* __restrict__ occupied, unsigned long long* __restrict__ keys)
{
    __shared__ unsigned long long bufS[PROJ_LIST_WARPS][PROJ_SCAN_BUF];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, q = blockIdx.x * PROJ_LIST_WARPS + w;
    if (q >= nq) return;
    const orbm_proj_query Q = queries[q];
    const unsigned long long k = proj_scan_sorted(c, Q, qdesc + (size_t)q * 32, lane, bufS[w],
                                                  [&](int i) { return occupied[i] == 0; });
    keys[(size_t)PROJ_L * q + lane] = k;
}

// best / second best free key -> the reference's acceptance (src/ORBmatcher.cc:100-127, :1428-1433, :1561-1566, :384-389)
__device__ __forceinline__ int proj_decide(unsigned long long k1, unsigned long long k2, int variant, int thDist, float nnratio, int& bDist)
{
    const unsigned long long NONE = ~0ull;
    bDist = 256;
    if (k1 == NONE) return -1;
    const int bestDist = (int)(k1 >> 40), bestLevel = (int)(k1 & 0xff), bestIdx = (int)((k1 >> 8) & 0xffff);
    const int bestDist2 = (k2 != NONE && (int)(k2 >> 40) < 256) ? (int)(k2 >> 40) : 256;      // initial values 256 / -1 (:73-77)
    const int bestLevel2 = (k2 != NONE && (int)(k2 >> 40) < 256) ? (int)(k2 & 0xff) : -1;
    bool ok = bestDist < 256 && bestDist <= thDist;
    if (ok && variant == ORBM_PROJ_LOCALMAP && bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) ok = false;
    if (!ok) return -1;
    bDist = bestDist;
    return bestIdx;
}

__global__ void __launch_bounds__(PROJ_FIX_THREADS)
proj_fixpoint_kernel(ProjScan c, const orbm_proj_query* __restrict__ queries, const u8* __restrict__ qdesc, int nq,
                     const u8* __restrict__ queryBlocks, u8* __restrict__ occupied, int* __restrict__ assigned,
                     int* __restrict__ bestIdxOut, int* __restrict__ bestDistOut, int variant, int thDist, float nnratio,
                     int checkOri, int* __restrict__ nmatchesOut, const unsigned long long* __restrict__ keys,
                     int* __restrict__ dirtyList, int* __restrict__ iterationsOut)
{
    extern __shared__ __align__(16) u8 fixS[];
    __shared__ int hist[ORBX_HISTO_LENGTH];
    __shared__ int sChanged, sDirty, sTakes, sMax[3];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, n = c.n;
    int* ownerA = reinterpret_cast<int*>(fixS);
    int* ownerB = ownerA + n;
    unsigned long long* scanBuf = reinterpret_cast<unsigned long long*>(fixS + (((size_t)2 * n * sizeof(int) + 15) & ~(size_t)15));
    const unsigned long long NONE = ~0ull;
    const int INF = 0x7fffffff;
    const bool rot = checkOri && (variant == ORBM_PROJ_LASTFRAME || variant == ORBM_PROJ_KEYFRAME);
    const bool alwaysBlocks = variant == ORBM_PROJ_KEYFRAME || variant == ORBM_PROJ_SIM3;
    const int needFree = variant == ORBM_PROJ_LOCALMAP ? 2 : 1;
    for (int i = tid; i < n; i += PROJ_FIX_THREADS) { ownerA[i] = INF; assigned[i] = -1; }
    for (int q = tid; q < nq; q += PROJ_FIX_THREADS) { bestIdxOut[q] = -2; bestDistOut[q] = 256; }   // -2: no decision yet
    if (tid < ORBX_HISTO_LENGTH) hist[tid] = 0;
    if (tid == 0) sTakes = 0;
    __syncthreads();
    int* prev = ownerA;
    int* next = ownerB;
    int iterations = 0;
    for (;;) {
        ++iterations;
        for (int i = tid; i < n; i += PROJ_FIX_THREADS) next[i] = INF;
        if (tid == 0) { sChanged = 0; sDirty = 0; }
        __syncthreads();
        // decisions under the takes of the previous iteration
        for (int q = tid; q < nq; q += PROJ_FIX_THREADS) {
            const unsigned long long* K = keys + (size_t)PROJ_L * q;
            unsigned long long k1 = NONE, k2 = NONE;
            int nfree = 0;
            bool complete = false;
            for (int j = 0; j < PROJ_L; ++j) {
                const unsigned long long k = K[j];
                if (k == NONE) { complete = true; break; }
                if (prev[(int)((k >> 8) & 0xffff)] >= q) {
                    if (nfree == 0) k1 = k; else k2 = k;
                    if (++nfree == needFree) break;
                }
            }
            if (!complete && nfree < needFree) dirtyList[atomicAdd(&sDirty, 1)] = q;
            else {
                int dist;
                const int take = proj_decide(k1, k2, variant, thDist, nnratio, dist);
                if (take != bestIdxOut[q]) sChanged = 1;
                bestIdxOut[q] = take; bestDistOut[q] = dist;
            }
        }
        __syncthreads();
        // truncated lists without enough free keys: the sweep again, under the occupancy the query sees now
        const int nd = sDirty;
        for (int d = wid; d < nd; d += PROJ_FIX_THREADS / 32) {
            const int q = dirtyList[d];
            const orbm_proj_query Q = queries[q];
            const unsigned long long k = proj_scan_sorted(c, Q, qdesc + (size_t)q * 32, lane, scanBuf + wid * PROJ_SCAN_BUF,
                                                          [&](int i) { return occupied[i] == 0 && prev[i] >= q; });
            const unsigned long long k1 = __shfl_sync(FULL, k, 0), k2 = __shfl_sync(FULL, k, 1);
            if (lane == 0) {
                int dist;
                const int take = proj_decide(k1, k2, variant, thDist, nnratio, dist);
                if (take != bestIdxOut[q]) sChanged = 1;
                bestIdxOut[q] = take; bestDistOut[q] = dist;
            }
        }
        __syncthreads();
        if (!sChanged) break;                      // every take repeats: prev already holds exactly these takes
        for (int q = tid; q < nq; q += PROJ_FIX_THREADS) {
            const int take = bestIdxOut[q];
            const bool blocks = alwaysBlocks ? true : (queryBlocks ? queryBlocks[q] != 0 : true);
            if (take >= 0 && blocks) atomicMin(&next[take], q);
        }
        __syncthreads();
        int* t = prev; prev = next; next = t;
    }
    // results: the query that writes a keypoint last owns it; blocked keypoints become occupied
    int mine = 0;
    for (int q = tid; q < nq; q += PROJ_FIX_THREADS) {
        const int take = bestIdxOut[q];
        if (take >= 0) {
            ++mine;
            atomicMax(&assigned[take], q);
            if (rot) atomicAdd(&hist[rot_bin(queries[q].angle, c.kps[take].angle)], 1);
        }
    }
    for (int i = tid; i < n; i += PROJ_FIX_THREADS) if (prev[i] != INF) occupied[i] = 1;
    if (mine) atomicAdd(&sTakes, mine);
    __syncthreads();
    if (rot) {
        if (tid == 0) three_maxima(hist, ORBX_HISTO_LENGTH, sMax[0], sMax[1], sMax[2]);
        __syncthreads();
        int culled = 0;
        for (int q = tid; q < nq; q += PROJ_FIX_THREADS) {
            const int take = bestIdxOut[q];
            if (take >= 0) {
                const int b = rot_bin(queries[q].angle, c.kps[take].angle);
                if (b != sMax[0] && b != sMax[1] && b != sMax[2]) { assigned[take] = -1; ++culled; }
            }
        }
        if (culled) atomicSub(&sTakes, culled);
        __syncthreads();
    }
    if (tid == 0) { *nmatchesOut = sTakes; if (iterationsOut) *iterationsOut = iterations; }
}

// =============================================================================================
// K7, latency form (one pair of at most PROJ_SCAN_MAXN keypoints per frame: the stateless call of the initializer).
//  K7a  si_list_kernel      warp per octave-0 keypoint of F1, all SMs: the sweep of K9a over F2 with the level window [0, 0]
//       (src/ORBmatcher.cc:425) and only the candidates within dkeep = TH_LOW / nnratio + 2 (farther ones can neither win nor
//       make the ratio test fail): the 32 smallest keys dist|cell|idx, sorted.
//  K7b  si_fixpoint_kernel  one CTA.  The reference skips a candidate i2 when an EARLIER query took it with a distance <= the
//       candidate's (vMatchedDistance, :447).  As in K9b the rule is iterated for all queries at once from the takes of the
//       previous iteration (takers of a keypoint: a linked list head[i2] -> next[q]); the prefix of final queries grows by at
//       least one per iteration.  Then the reference's bookkeeping in closed form: the LAST taker of a keypoint owns it
//       (earlier ones were reset to -1, :462-466), nmatches = keypoints taken, every taker votes in the rotation histogram
//       (displaced ones too, :470-482), the cull only resets owners.
// =============================================================================================
__global__ void __launch_bounds__(PROJ_LIST_WARPS * 32)
si_list_kernel(ProjScan c, const orbx_keypoint* __restrict__ k1, const u8* __restrict__ d1, const float* __restrict__ prev, int n1,
               float window, int dkeep, unsigned long long* __restrict__ keys)
{
    __shared__ unsigned long long bufS[PROJ_LIST_WARPS][PROJ_SCAN_BUF];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, i1 = blockIdx.x * PROJ_LIST_WARPS + w;
    if (i1 >= n1 || k1[i1].octave > 0) return;
    orbm_proj_query Q;
    Q.u = prev[2 * i1]; Q.v = prev[2 * i1 + 1]; Q.radius = window;
    Q.min_level = 0; Q.max_level = 0; Q.u_right = 0.f; Q.max_right_err = -1.f; Q.angle = 0.f; Q.level_lo = -1; Q.level_hi = -1;
    const unsigned long long k = proj_scan_sorted(c, Q, d1 + (size_t)i1 * 32, lane, bufS[w], [](int) { return true; },
                                                  [dkeep](int, int dist) { return dist <= dkeep; });
    keys[(size_t)PROJ_L * i1 + lane] = k;
}

__global__ void __launch_bounds__(PROJ_FIX_THREADS)
si_fixpoint_kernel(ProjScan c, const orbx_keypoint* __restrict__ k1, const u8* __restrict__ d1, float* __restrict__ prev, int n1,
                   float window, int dkeep, float nnratio, int checkOri, const unsigned long long* __restrict__ keys,
                   int* __restrict__ m12, int* __restrict__ nmatchesOut, int* __restrict__ dirtyList)
{
    extern __shared__ __align__(16) u8 fixS[];
    __shared__ int hist[ORBX_HISTO_LENGTH];
    __shared__ int sChanged, sDirty, sCount, sMax[3];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, n2 = c.n;
    // take[q]: keypoint of F2 query q takes (-1 none, -2 no decision yet); tdist[q]: its distance; head / next: takers of a keypoint
    // (tprev: the distances the lists were built with, so that an iteration reads one consistent state)
    int* take = reinterpret_cast<int*>(fixS);
    int* tdist = take + n1;
    int* tprev = tdist + n1;
    int* next = tprev + n1;
    int* head = next + n1;
    unsigned long long* scanBuf = reinterpret_cast<unsigned long long*>(fixS + ((((size_t)4 * n1 + n2) * sizeof(int) + 15) & ~(size_t)15));
    const unsigned long long NONE = ~0ull;
    for (int i = tid; i < n1; i += PROJ_FIX_THREADS) { take[i] = -2; tdist[i] = 0; next[i] = -1; }
    for (int i = tid; i < n2; i += PROJ_FIX_THREADS) head[i] = -1;
    if (tid < ORBX_HISTO_LENGTH) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    // is candidate (i2, dist) still on offer when query q's turn comes?
    auto visible = [&](int q, int i2, int dist) {
        for (int t = head[i2]; t >= 0; t = next[t]) if (t < q && tprev[t] <= dist) return false;
        return true;
    };
    auto decide = [&](int q, unsigned long long ka, unsigned long long kb) {
        int t = -1, dist = 0;
        if (ka != NONE) {
            const int best = (int)(ka >> 40);
            // bestDist2 starts at INT_MAX (:435): without a second candidate the ratio test passes
            const bool ratioOk = kb == NONE || (float)best < __fmul_rn((float)(int)(kb >> 40), nnratio);
            if (best <= ORBX_TH_LOW && ratioOk) { t = (int)((ka >> 8) & 0xffff); dist = best; }
        }
        if (t != take[q] || dist != tdist[q]) sChanged = 1;
        take[q] = t; tdist[q] = dist;
    };
    for (;;) {
        if (tid == 0) { sChanged = 0; sDirty = 0; }
        __syncthreads();
        for (int q = tid; q < n1; q += PROJ_FIX_THREADS) {
            if (k1[q].octave > 0) { take[q] = -1; continue; }
            const unsigned long long* K = keys + (size_t)PROJ_L * q;
            unsigned long long ka = NONE, kb = NONE;
            int nvis = 0;
            bool complete = false;
            for (int j = 0; j < PROJ_L; ++j) {
                const unsigned long long k = K[j];
                if (k == NONE) { complete = true; break; }
                if (visible(q, (int)((k >> 8) & 0xffff), (int)(k >> 40))) {
                    if (nvis == 0) ka = k; else kb = k;
                    if (++nvis == 2) break;
                }
            }
            if (!complete && nvis < 2) dirtyList[atomicAdd(&sDirty, 1)] = q;       // a full list used up: searched again below
            else decide(q, ka, kb);
        }
        __syncthreads();
        const int nd = sDirty;
        for (int d = wid; d < nd; d += PROJ_FIX_THREADS / 32) {
            const int q = dirtyList[d];
            orbm_proj_query Q;
            Q.u = prev[2 * q]; Q.v = prev[2 * q + 1]; Q.radius = window;
            Q.min_level = 0; Q.max_level = 0; Q.u_right = 0.f; Q.max_right_err = -1.f; Q.angle = 0.f; Q.level_lo = -1; Q.level_hi = -1;
            const unsigned long long k = proj_scan_sorted(c, Q, d1 + (size_t)q * 32, lane, scanBuf + wid * PROJ_SCAN_BUF, [](int) { return true; },
                                                          [&](int i2, int dist) { return dist <= dkeep && visible(q, i2, dist); });
            const unsigned long long ka = __shfl_sync(FULL, k, 0), kb = __shfl_sync(FULL, k, 1);
            if (lane == 0) decide(q, ka, kb);
        }
        __syncthreads();
        if (!sChanged) break;
        // takers of the next iteration
        for (int i = tid; i < n2; i += PROJ_FIX_THREADS) head[i] = -1;
        __syncthreads();
        for (int q = tid; q < n1; q += PROJ_FIX_THREADS) if (take[q] >= 0) { tprev[q] = tdist[q]; next[q] = atomicExch(&head[take[q]], q); }
        __syncthreads();
    }
    // the lists hold the final takes (the last iteration changed nothing).  Owner of a keypoint = its last taker.
    for (int q = tid; q < n1; q += PROJ_FIX_THREADS) {
        const int t = take[q];
        int m = -1;
        if (t >= 0) {
            int last = q;
            for (int o = head[t]; o >= 0; o = next[o]) last = max(last, o);
            if (last == q) { m = t; atomicAdd(&sCount, 1); }
            if (checkOri) atomicAdd(&hist[rot_bin(k1[q].angle, c.kps[t].angle)], 1);
        }
        m12[q] = m;
    }
    __syncthreads();
    if (checkOri) {
        if (tid == 0) three_maxima(hist, ORBX_HISTO_LENGTH, sMax[0], sMax[1], sMax[2]);
        __syncthreads();
        for (int q = tid; q < n1; q += PROJ_FIX_THREADS) {
            const int t = m12[q];
            if (t >= 0) {
                const int b = rot_bin(k1[q].angle, c.kps[t].angle);
                if (b != sMax[0] && b != sMax[1] && b != sMax[2]) { m12[q] = -1; atomicSub(&sCount, 1); }
            }
        }
        __syncthreads();
    }
    // vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt (:510-513)
    for (int q = tid; q < n1; q += PROJ_FIX_THREADS) {
        const int t = m12[q];
        if (t >= 0) { prev[2 * q] = c.kps[t].x; prev[2 * q + 1] = c.kps[t].y; }
    }
    if (tid == 0) *nmatchesOut = sCount;
}

// =============================================================================================
// K8: Frame::ComputeStereoMatches (src/Frame.cc:466-638).  One CTA per stereo pair.
//  phase 1  every right keypoint -> (minr, maxr, octave, x): the rows it is filed under (:480-491)
//  phase 2  warp per left keypoint: lanes scan the right keypoints (ascending index = the order
//           of vRowIndices[row]), Hamming argmin with strict '<' from TH_HIGH (:502-547), then the
//           11x11 SAD over 11 shifts on the raw pyramids + parabola (:550-621); SAD is integer
//           valued so int arithmetic equals the reference's float L1 norm.
//  phase 3  median of the accepted SADs by bisection on the value, cull >= 1.5*1.4*median (:624-637)
// =============================================================================================
struct PyrView {
    const u8* base[2][ORBX_MAX_LEVELS];        // [eye][level], pair p at base + p*pairStride
    long long pairStride[2][ORBX_MAX_LEVELS];
    int pitch[2][ORBX_MAX_LEVELS];
    int w[ORBX_MAX_LEVELS], h[ORBX_MAX_LEVELS];
    float scale[ORBX_MAX_LEVELS], invScale[ORBX_MAX_LEVELS];
    int nlevels;
};

__device__ __forceinline__ int reflect101i(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * (n - 1) - i;
    return i;
}

#define STEREO_THREADS 256
#define STEREO_SPLITS 8           // CTAs per pair in a batch (left keypoints are split; the median cull is a second kernel); one pair: up to 64
__global__ void __launch_bounds__(STEREO_THREADS)
stereo_match_kernel(FrameStore fs, int firstSlot, PyrView pv, float bf, float fx, int rightCap, int listCap,
                    float* __restrict__ uRightOut, float* __restrict__ depthOut, int* __restrict__ sadScratch,
                    unsigned long long* __restrict__ ncompares)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    __shared__ int sPart[STEREO_THREADS / 32], sTotal;
    const int nRows = pv.h[0];
    // vRowIndices (src/Frame.cc:478-492) as a CSR table in shared memory: rowEnd[r] = end of row r's list
    int* rowEnd = reinterpret_cast<int*>(smraw);
    float* rX = reinterpret_cast<float*>(rowEnd + ((nRows + 4) & ~3));
    unsigned short* rowList = reinterpret_cast<unsigned short*>(rX + rightCap);
    u8* rOct = reinterpret_cast<u8*>(rowList + listCap);
    const int pair = blockIdx.x, split = blockIdx.y;
    const int sL = firstSlot + 2 * pair, sR = sL + 1;
    const int N = min(fs.count[sL], fs.cap), Nr = min(min(fs.count[sR], fs.cap), rightCap);
    const int perSplit = (N + (int)gridDim.y - 1) / (int)gridDim.y;
    const int iBeg = split * perSplit, iEnd = min(N, iBeg + perSplit);
    const orbx_keypoint* kL = fs.kps + (size_t)sL * fs.cap;
    const orbx_keypoint* kR = fs.kps + (size_t)sR * fs.cap;
    const u8* dL = fs.desc + (size_t)sL * fs.cap * 32;
    const u8* dR = fs.desc + (size_t)sR * fs.cap * 32;
    float* uOut = uRightOut + (size_t)pair * fs.cap;
    float* zOut = depthOut + (size_t)pair * fs.cap;
    int* sad = sadScratch + (size_t)pair * fs.cap;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int r = tid; r < nRows; r += STEREO_THREADS) rowEnd[r] = 0;
    if (split == 0)
        for (int i = N + tid; i < fs.cap; i += STEREO_THREADS) { uOut[i] = -1.f; zOut[i] = -1.f; }   // unused tail
    __syncthreads();
    // rows minr..maxr of every right keypoint (clipped to the image), r = 2 * scale[octave]
    auto row_range = [&](const orbx_keypoint& kp, int& minr, int& maxr) {
        const float r = __fmul_rn(2.0f, pv.scale[kp.octave]);
        maxr = min((int)ceilf(__fadd_rn(kp.y, r)), nRows - 1);
        minr = max((int)floorf(__fsub_rn(kp.y, r)), 0);
    };
    for (int i = tid; i < Nr; i += STEREO_THREADS) {
        const orbx_keypoint kp = kR[i];
        int minr, maxr;
        row_range(kp, minr, maxr);
        for (int r = minr; r <= maxr; ++r) atomicAdd(&rowEnd[r], 1);
        rX[i] = kp.x;
        rOct[i] = (u8)kp.octave;
    }
    __syncthreads();
    {   // exclusive scan of the row counts: thread t owns a contiguous run of rows
        const int per = (nRows + STEREO_THREADS - 1) / STEREO_THREADS;
        const int r0 = min(tid * per, nRows), r1 = min(r0 + per, nRows);
        int sum = 0;
        for (int r = r0; r < r1; ++r) sum += rowEnd[r];
        int incl = sum;
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) sPart[wid] = incl;
        __syncthreads();
        int off = 0, tot = 0;
        for (int w = 0; w < STEREO_THREADS / 32; ++w) { const int v = sPart[w]; if (w < wid) off += v; tot += v; }
        if (tid == 0) sTotal = tot;
        int run = off + incl - sum;
        for (int r = r0; r < r1; ++r) { const int c = rowEnd[r]; rowEnd[r] = run; run += c; }
    }
    __syncthreads();
    const bool useRows = sTotal <= listCap;          // otherwise (pathological scales) every right keypoint is scanned
    if (useRows)
        for (int i = tid; i < Nr; i += STEREO_THREADS) {
            int minr, maxr;
            row_range(kR[i], minr, maxr);
            for (int r = minr; r <= maxr; ++r) rowList[atomicAdd(&rowEnd[r], 1)] = (unsigned short)i;   // cursor ends at the row's end
        }
    __syncthreads();
    const float mb = __fdiv_rn(bf, fx);
    const float minD = -3.f, maxD = __fdiv_rn(bf, mb);
    unsigned long long ncmp = 0;
    for (int iL = iBeg + wid; iL < iEnd; iL += STEREO_THREADS / 32) {
        const orbx_keypoint kp = kL[iL];
        float uR = -1.f, z = -1.f;
        int bestSad = -1;
        const int row = (int)kp.y;
        const float minU = __fsub_rn(kp.x, maxD), maxU = __fsub_rn(kp.x, minD);
        if (row >= 0 && row < nRows && !(maxU < 0)) {
            uint32_t q[8];
            load_desc(q, dL + (size_t)iL * 32);
            unsigned best = ((unsigned)ORBX_TH_HIGH << 16);
            const int jBeg = useRows ? (row > 0 ? rowEnd[row - 1] : 0) : 0, jEnd = useRows ? rowEnd[row] : Nr;
            for (int j = jBeg + lane; j < jEnd; j += 32) {
                int iR = j;
                if (useRows) iR = rowList[j];
                else { int minr, maxr; row_range(kR[iR], minr, maxr); if (row < minr || row > maxr) continue; }
                const int o = rOct[iR];
                if (o < kp.octave - 1 || o > kp.octave + 1) continue;
                const float x = rX[iR];
                if (!(x >= minU && x <= maxU)) continue;
                const unsigned dist = (unsigned)hamming256(q, reinterpret_cast<const uint32_t*>(dR + (size_t)iR * 32));
                ++ncmp;
                const unsigned key = (dist << 16) | (unsigned)iR;
                if (key < best) best = key;        // dist < TH_HIGH, ties -> lowest iR
            }
            for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(FULL, best, d));
            if ((best >> 16) < (unsigned)ORBX_TH_HIGH) {
                const int bestIdxR = (int)(best & 0xffff);
                const int lvl = kp.octave;
                const float uR0 = rX[bestIdxR];
                const float sf = pv.invScale[lvl];
                const float scaleduL = roundf(__fmul_rn(kp.x, sf)), scaledvL = roundf(__fmul_rn(kp.y, sf));
                const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
                const int w = 5, L = 5;
                const float iniu = __fsub_rn(__fadd_rn(scaleduR0, (float)L), (float)w);
                const float endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, (float)L), (float)w), 1.f);
                if (!(iniu < 0 || endu >= (float)pv.w[lvl])) {
                    const int cuL = (int)scaleduL, cvL = (int)scaledvL, cuR = (int)scaleduR0;
                    const int lw = pv.w[lvl], lh = pv.h[lvl];
                    const u8* pl = pv.base[0][lvl] + (long long)pair * pv.pairStride[0][lvl];
                    const u8* pr = pv.base[1][lvl] + (long long)pair * pv.pairStride[1][lvl];
                    const int pitL = pv.pitch[0][lvl], pitR = pv.pitch[1][lvl];
                    int acc[11];
#pragma unroll
                    for (int s = 0; s < 11; ++s) acc[s] = 0;
                    if (cuL >= w && cuL + w < lw && cvL >= w && cvL + w < lh && cuR >= w + L && cuR + w + L < lw) {
                        // both windows inside the level (always, for keypoints the extractor produced): plain
                        // addressing, |(a - cL) - (b - cR_s)| = |(a + (cR_s - cL)) - b| as one add + one SAD
                        const u8* lc = pl + (long long)cvL * pitL + cuL;
                        const u8* rc = pr + (long long)cvL * pitR + cuR - L;
                        const int cL = lc[0];
                        int kS[11];
#pragma unroll
                        for (int s = 0; s < 11; ++s) kS[s] = (int)rc[s] - cL;
                        for (int p = lane; p < 121; p += 32) {
                            const int dy = p / 11 - w, dx = p - (dy + w) * 11 - w;
                            const int a = lc[dy * pitL + dx];
                            const u8* rq = rc + dy * pitR + dx;
#pragma unroll
                            for (int s = 0; s < 11; ++s) acc[s] = (int)__sad(a + kS[s], (int)rq[s], (unsigned)acc[s]);
                        }
                    } else {
                        const int cL = pl[(long long)reflect101i(cvL, lh) * pitL + reflect101i(cuL, lw)];
                        int cR[11];
#pragma unroll
                        for (int s = 0; s < 11; ++s) cR[s] = pr[(long long)reflect101i(cvL, lh) * pitR + reflect101i(cuR + s - L, lw)];
                        for (int p = lane; p < 121; p += 32) {
                            const int dy = p / 11 - w, dx = p % 11 - w;
                            const int yy = reflect101i(cvL + dy, lh);
                            const int a = (int)pl[(long long)yy * pitL + reflect101i(cuL + dx, lw)] - cL;
                            const u8* rrow = pr + (long long)yy * pitR;
#pragma unroll
                            for (int s = 0; s < 11; ++s) {
                                const int b = (int)rrow[reflect101i(cuR + s - L + dx, lw)] - cR[s];
                                acc[s] += abs(a - b);
                            }
                        }
                    }
#pragma unroll
                    for (int s = 0; s < 11; ++s)
                        for (int d = 16; d > 0; d >>= 1) acc[s] += __shfl_xor_sync(FULL, acc[s], d);
                    int bs = 0x7fffffff, binc = 0;
#pragma unroll
                    for (int s = 0; s < 11; ++s) if (acc[s] < bs) { bs = acc[s]; binc = s - L; }
                    if (!(binc == -L || binc == L)) {
                        float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                        for (int s = 0; s < 11; ++s) {
                            if (s == binc + L - 1) d1 = (float)acc[s];
                            if (s == binc + L) d2 = (float)acc[s];
                            if (s == binc + L + 1) d3 = (float)acc[s];
                        }
                        const float deltaR = __fdiv_rn(__fsub_rn(d1, d3),
                                                       __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                        if (!(deltaR < -1 || deltaR > 1)) {
                            float bestuR = __fmul_rn(pv.scale[lvl], __fadd_rn(__fadd_rn(scaleduR0, (float)binc), deltaR));
                            float disparity = __fsub_rn(kp.x, bestuR);
                            if (disparity >= 0 && disparity < maxD) {
                                if (disparity <= 0) { disparity = (float)0.01; bestuR = (float)((double)kp.x - 0.01); }
                                z = __fdiv_rn(bf, disparity);
                                uR = bestuR;
                                bestSad = bs;
                            }
                        }
                    }
                }
            }
        }
        if (lane == 0) { uOut[iL] = uR; zOut[iL] = z; sad[iL] = bestSad; }
    }
    for (int d = 16; d > 0; d >>= 1) ncmp += __shfl_xor_sync(FULL, ncmp, d);
    if (lane == 0 && ncompares && ncmp) atomicAdd(ncompares, ncmp);
}

// phase 3 of Frame::ComputeStereoMatches (src/Frame.cc:624-637): median of the accepted SADs of one
// pair (k-th smallest with k = size/2, found by bisection on the value), cull SAD >= 1.5*1.4*median.
__global__ void __launch_bounds__(STEREO_THREADS)
stereo_cull_kernel(FrameStore fs, int firstSlot, float* __restrict__ uRightOut, float* __restrict__ depthOut,
                   const int* __restrict__ sadScratch)
{
    __shared__ int sCount[2];
    const int pair = blockIdx.x;
    const int N = min(fs.count[firstSlot + 2 * pair], fs.cap);
    float* uOut = uRightOut + (size_t)pair * fs.cap;
    float* zOut = depthOut + (size_t)pair * fs.cap;
    const int* sad = sadScratch + (size_t)pair * fs.cap;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < 2) sCount[tid] = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < N; i += STEREO_THREADS) cnt += sad[i] >= 0;
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(FULL, cnt, d);
    if (lane == 0 && cnt) atomicAdd(&sCount[0], cnt);
    __syncthreads();
    const int nValid = sCount[0];
    if (nValid == 0) return;
    const int k = nValid / 2;
    int lo = 0, hi = 121 * 510;          // smallest v with count(sad <= v) > k
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        __syncthreads();
        if (tid == 0) sCount[1] = 0;
        __syncthreads();
        int c = 0;
        for (int i = tid; i < N; i += STEREO_THREADS) { const int s = sad[i]; c += (s >= 0 && s <= mid); }
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(FULL, c, d);
        if (lane == 0 && c) atomicAdd(&sCount[1], c);
        __syncthreads();
        if (sCount[1] > k) hi = mid; else lo = mid + 1;
    }
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)lo);
    for (int i = tid; i < N; i += STEREO_THREADS) {
        const int s = sad[i];
        if (s >= 0 && !((float)s < thDist)) { uOut[i] = -1.f; zOut[i] = -1.f; }
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

// K6, latency form (a handful of frames, the stateless matcher calls): one CTA per frame instead of one warp.  Counts and
// fill positions by shared-memory atomics, a block scan over the 3072 cells, then every cell's short list is put back
// into insertion order (ascending keypoint index, what Frame::AssignFeaturesToGrid's push_back produces) by its thread.
#define GB_THREADS 256
__global__ void __launch_bounds__(GB_THREADS)
grid_build_cta_kernel(FrameStore fs, int firstSlot, GridParams gp)
{
    __shared__ int cnt[ORBX_GRID_CELLS];            // counts, then fill cursors
    __shared__ unsigned short startS[ORBX_GRID_CELLS + 1];
    __shared__ int warpSum[GB_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int slot = firstSlot + blockIdx.x;
    const int n = min(fs.count[slot], fs.cap);
    const orbx_keypoint* kps = fs.kps + (size_t)slot * fs.cap;
    unsigned short* gs = fs.gridStart + (size_t)slot * (ORBX_GRID_CELLS + 1);
    unsigned short* gi = fs.gridIdx + (size_t)slot * fs.cap;
    for (int c = tid; c < ORBX_GRID_CELLS; c += GB_THREADS) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += GB_THREADS) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
        if (px >= 0 && px < ORBX_GRID_COLS && py >= 0 && py < ORBX_GRID_ROWS) atomicAdd(&cnt[px * ORBX_GRID_ROWS + py], 1);
    }
    __syncthreads();
    // exclusive scan: 12 consecutive cells per thread
    const int per = ORBX_GRID_CELLS / GB_THREADS;
    int sum = 0;
    for (int c = 0; c < per; ++c) sum += cnt[tid * per + c];
    int incl = sum;
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
    if (lane == 31) warpSum[wid] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < wid; ++w) base += warpSum[w];
    int run = base + incl - sum;
    for (int c = 0; c < per; ++c) {
        const int k = cnt[tid * per + c];
        startS[tid * per + c] = (unsigned short)run;
        gs[tid * per + c] = (unsigned short)run;
        cnt[tid * per + c] = 0;                     // becomes the fill cursor
        run += k;
    }
    if (tid == GB_THREADS - 1) { startS[ORBX_GRID_CELLS] = (unsigned short)run; gs[ORBX_GRID_CELLS] = (unsigned short)run; }
    __syncthreads();
    for (int i = tid; i < n; i += GB_THREADS) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
        if (px >= 0 && px < ORBX_GRID_COLS && py >= 0 && py < ORBX_GRID_ROWS) {
            const int cell = px * ORBX_GRID_ROWS + py;
            gi[startS[cell] + atomicAdd(&cnt[cell], 1)] = (unsigned short)i;
        }
    }
    __syncthreads();
    for (int c = tid * per; c < tid * per + per; ++c) {
        const int s = startS[c], e = startS[c + 1];
        for (int a = s + 1; a < e; ++a) {           // insertion sort of a list of a few entries
            const unsigned short v = gi[a];
            int b = a - 1;
            while (b >= s && gi[b] > v) { gi[b + 1] = gi[b]; --b; }
            gi[b + 1] = v;
        }
    }
}

// =============================================================================================
// host side
// =============================================================================================
int orbx_ensure_match_buffers(orbx_handle* h)
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
    fs.kps = h->storeUndistorted ? h->dKpsUn : h->dKps; fs.desc = h->dDesc; fs.count = h->dCount;
    fs.gridStart = h->dGridStart; fs.gridIdx = h->dGridIdx; fs.cap = h->kpCap;
    return fs;
}

GridParams orbx_grid_params(float minX, float maxX, float minY, float maxY)
{
    GridParams gp;
    gp.minX = minX; gp.minY = minY;
    gp.invW = static_cast<float>(ORBX_GRID_COLS) / (maxX - minX);   // src/Frame.cc:101-102
    gp.invH = static_cast<float>(ORBX_GRID_ROWS) / (maxY - minY);
    return gp;
}

int orbx_launch_grid_build(const FrameStore& fs, int firstSlot, int nframes, const GridParams& gp, cudaStream_t st)
{
    if (nframes <= 4) {                      // a few frames leave the GPU idle: a CTA per frame is 5x faster than a warp
        grid_build_cta_kernel<<<nframes, GB_THREADS, 0, st>>>(fs, firstSlot, gp);
        return ORBX_OK;
    }
    const int smem = (ORBX_GRID_CELLS + fs.cap) * (int)sizeof(unsigned short);
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    grid_build_kernel<<<nframes, 32, smem, st>>>(fs, firstSlot, gp);
    return ORBX_OK;
}

static int launch_search_init(const FrameStore& fs, int firstSlot, int slotStride, int npairs, const GridParams& gp,
                              float* prev, int window, float nnratio, int checkOri, int* m12, int* nm,
                              unsigned long long* ncmp, cudaStream_t st)
{
    const size_t smem = si_smem_bytes(fs.cap);
    if (smem > 220 * 1024) { orbx_set_error("frame too large for SearchForInitialization (%d keypoints)", fs.cap); return ORBX_E_CAPACITY; }
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(search_init_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // candidates farther than TH_LOW / nnratio can neither be the match nor make the ratio test fail
    int dkeep = 256;
    if (nnratio > 0.f) { const double d = (double)ORBX_TH_LOW / (double)nnratio + 2.0; dkeep = d < 256.0 ? (int)d : 256; }
    search_init_kernel<<<npairs, SI_WARPS * 32, smem, st>>>(fs, firstSlot, slotStride, gp, prev, window, nnratio, checkOri, m12, nm, ncmp, dkeep);
    return ORBX_OK;
}

int orbx_build_grids(orbx_handle* h, int firstSlot, int nframes, int width, int height)
{
    int rc = orbx_ensure_match_buffers(h);
    if (rc) return rc;
    const GridParams gp = h->storeUndistorted ? orbx_grid_params(h->bounds[0], h->bounds[1], h->bounds[2], h->bounds[3])
                                              : orbx_grid_params(0.f, (float)width, 0.f, (float)height);
    return orbx_launch_grid_build(store_of(h), firstSlot, nframes, gp, h->stream);
}

static int launch_stereo(const FrameStore& fs, int firstSlot, int npairs, const PyrView& pv, float bf, float fx,
                         int rightCap, float* uRight, float* depth, int* sadScratch, unsigned long long* ncmp,
                         cudaStream_t st)
{
    // row table: (rows + 4) ints, then per right keypoint x (4 B), octave (1 B) and room for 10 row entries (2 B each;
    // the standard 8-level pyramid needs 7.8 on average) — beyond that the kernel falls back to scanning every keypoint
    int listCap = rightCap * 10;
    int smem = ((pv.h[0] + 4) & ~3) * 4 + rightCap * 5 + listCap * 2 + 16;
    if (smem > 200 * 1024) { listCap = 0; smem = ((pv.h[0] + 4) & ~3) * 4 + rightCap * 5 + 16; }
    if (smem > 200 * 1024) { orbx_set_error("more than %d right keypoints per frame unsupported by the stereo matcher", 180 * 1024 / 5); return ORBX_E_CAPACITY; }
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(stereo_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    // a batch fills the GPU with 8 CTAs per pair; a single pair (the per-frame call) is spread over more SMs
    const int splits = npairs >= 16 ? STEREO_SPLITS : std::min(64, std::max(STEREO_SPLITS, 128 / npairs));
    stereo_match_kernel<<<dim3(npairs, splits), STEREO_THREADS, smem, st>>>(fs, firstSlot, pv, bf, fx, rightCap, listCap, uRight, depth, sadScratch, ncmp);
    stereo_cull_kernel<<<npairs, STEREO_THREADS, 0, st>>>(fs, firstSlot, uRight, depth, sadScratch);
    return ORBX_OK;
}

int orbx_match_chunk(orbx_handle* h, int gridFirst, int gridCount, int pairFirst, int npairs, int window, float nnratio, int checkOri)
{
    ORBX_NVTX("orbx_match_chunk");
    const FrameStore fs = store_of(h);
    const GridParams gp = h->storeUndistorted ? orbx_grid_params(h->bounds[0], h->bounds[1], h->bounds[2], h->bounds[3])
                                              : orbx_grid_params(0.f, (float)h->curW, 0.f, (float)h->curH);
    orbx_prof_mark(h, -1);
    int rc = orbx_launch_grid_build(fs, gridFirst, gridCount, gp, h->stream);
    if (rc) return rc;
    ORBX_COUNT(h, ORBX_STAGE_GRID, 1);
    orbx_prof_mark(h, ORBX_STAGE_GRID);
    if (npairs > 0) {
        rc = launch_search_init(fs, pairFirst, 1, npairs, gp, nullptr, window, nnratio, checkOri,
                                h->dMatches12 + (size_t)pairFirst * h->kpCap, h->dNMatches + pairFirst, nullptr, h->stream);
        if (rc) return rc;
        ORBX_COUNT(h, ORBX_STAGE_SEARCH_INIT, 1);
        orbx_prof_mark(h, ORBX_STAGE_SEARCH_INIT);
    }
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
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
    int rc = orbx_ensure_match_buffers(h);
    if (rc) return rc;
    const FrameStore fs = store_of(h);
    const GridParams gp = h->storeUndistorted ? orbx_grid_params(h->bounds[0], h->bounds[1], h->bounds[2], h->bounds[3])
                                              : orbx_grid_params(0.f, (float)h->curW, 0.f, (float)h->curH);
    orbx_prof_mark(h, -1);
    rc = orbx_launch_grid_build(fs, first_slot, npairs + 1, gp, h->stream);
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
    CallArena ar;
    const int sA = ar.in(a, (size_t)count * 32), sB = ar.in(b, (size_t)count * 32), sD = ar.out(dist, (size_t)count * 4);
    if ((rc = ar.commit())) return rc;
    hamming_pairs_kernel<<<(count + 255) / 256, 256, 0, ar.st>>>(ar.dev<u8>(sA), ar.dev<u8>(sB), count, ar.dev<int>(sD));
    if ((rc = ar.finish())) return rc;
    return ORBX_OK;
}

int orbm_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_matched,
                                   int32_t* matches12, int window_size, float nnratio,
                                   int check_orientation, int32_t* nmatches)
{
    ORBX_NVTX("orbm_search_for_initialization");
    int rc = check_frame(f1); if (rc) return rc;
    rc = check_frame(f2); if (rc) return rc;
    if (!prev_matched || !matches12) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    rc = require_device(); if (rc) return rc;
    // one arena, one upload, one download (see orbm_search_by_projection)
    const size_t cap = (size_t)std::max(std::max(f1->n, f2->n), 8);
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oCount = take(8), oKps = take(2 * cap * sizeof(orbx_keypoint)), oDesc = take(2 * cap * 32);
    const size_t oPrev = take(cap * 2 * sizeof(float));          // in/out: last of the upload, first of the download
    const size_t upBytes = off;
    const size_t oNm = take(4), oM12 = take(cap * 4);
    const size_t downBytes = off - oPrev;
    const size_t oGS = take(2 * (ORBX_GRID_CELLS + 1) * sizeof(unsigned short)), oGI = take(2 * cap * sizeof(unsigned short));
    const bool sweep = f1->n <= PROJ_SCAN_MAXN && f2->n <= PROJ_SCAN_MAXN;
    const size_t oKeys = take(sweep ? cap * PROJ_L * 8 : 8), oDirty = take(sweep ? cap * 4 : 4);
    DevBuf arena;
    if ((rc = arena.alloc(off))) return rc;
    u8* host = nullptr;
    if ((rc = orbx_scratch_host((void**)&host, oPrev + downBytes))) return rc;
    u8* dev = arena.as<u8>();
    const int counts[2] = {f1->n, f2->n};
    memcpy(host + oCount, counts, 8);
    if (f1->n) {
        memcpy(host + oKps, f1->keys_un, (size_t)f1->n * sizeof(orbx_keypoint));
        memcpy(host + oDesc, f1->descriptors, (size_t)f1->n * 32);
        memcpy(host + oPrev, prev_matched, (size_t)f1->n * 2 * sizeof(float));
    }
    if (f2->n) {
        memcpy(host + oKps + cap * sizeof(orbx_keypoint), f2->keys_un, (size_t)f2->n * sizeof(orbx_keypoint));
        memcpy(host + oDesc + cap * 32, f2->descriptors, (size_t)f2->n * 32);
    }
    PhaseTrace pt(orbx_lat_trace_on());
    cudaStream_t st = orbx_call_stream();
    pt.mark("start", st);
    ORBX_CUDA(cudaMemcpyAsync(dev, host, upBytes, cudaMemcpyHostToDevice, st));
    pt.mark("h2d", st);
    FrameStore fs;
    fs.kps = reinterpret_cast<orbx_keypoint*>(dev + oKps); fs.desc = dev + oDesc; fs.count = reinterpret_cast<int*>(dev + oCount);
    fs.gridStart = reinterpret_cast<unsigned short*>(dev + oGS); fs.gridIdx = reinterpret_cast<unsigned short*>(dev + oGI);
    fs.cap = (int)cap;
    const GridParams gp = orbx_grid_params(f2->min_x, f2->max_x, f2->min_y, f2->max_y);
    if (sweep) {
        // latency form (K7a / K7b): no grid, sorted candidate lists on all SMs, fixed-point replay in one CTA
        ProjScan ps;
        ps.kps = fs.kps + cap; ps.desc = fs.desc + cap * 32; ps.uRight = nullptr; ps.gp = gp; ps.n = f2->n; ps.gate = false;
        int dkeep = 256;
        if (nnratio > 0.f) { const double d = (double)ORBX_TH_LOW / (double)nnratio + 2.0; dkeep = d < 256.0 ? (int)d : 256; }
        unsigned long long* dKeys = reinterpret_cast<unsigned long long*>(dev + oKeys);
        float* dPrev = reinterpret_cast<float*>(dev + oPrev);
        if (f1->n > 0)
            si_list_kernel<<<(f1->n + PROJ_LIST_WARPS - 1) / PROJ_LIST_WARPS, PROJ_LIST_WARPS * 32, 0, st>>>(ps, fs.kps, fs.desc, dPrev, f1->n, (float)window_size, dkeep, dKeys);
        pt.mark("lists", st);
        const int fixSmem = (int)(((((size_t)4 * f1->n + f2->n) * sizeof(int) + 15) & ~(size_t)15) + (PROJ_FIX_THREADS / 32) * PROJ_SCAN_BUF * 8);
        static bool attrSet = false;
        if (!attrSet) { ORBX_CUDA(cudaFuncSetAttribute(si_fixpoint_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * PROJ_SCAN_MAXN * 4 + 16 + (PROJ_FIX_THREADS / 32) * PROJ_SCAN_BUF * 8)); attrSet = true; }
        si_fixpoint_kernel<<<1, PROJ_FIX_THREADS, fixSmem, st>>>(ps, fs.kps, fs.desc, dPrev, f1->n, (float)window_size, dkeep, nnratio, check_orientation, dKeys,
                                                             reinterpret_cast<int*>(dev + oM12), reinterpret_cast<int*>(dev + oNm), reinterpret_cast<int*>(dev + oDirty));
        pt.mark("fixpoint", st);
    } else {
        rc = orbx_launch_grid_build(fs, 1, 1, gp, st); if (rc) return rc;
        pt.mark("grid", st);
        rc = launch_search_init(fs, 0, 1, 1, gp, reinterpret_cast<float*>(dev + oPrev), window_size, nnratio, check_orientation,
                                reinterpret_cast<int*>(dev + oM12), reinterpret_cast<int*>(dev + oNm), nullptr, st);
        if (rc) return rc;
        pt.mark("search", st);
    }
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(host + oPrev, dev + oPrev, downBytes, cudaMemcpyDeviceToHost, st));
    pt.mark("d2h", st);
    ORBX_CUDA(cudaStreamSynchronize(st));
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; pt.report("search_for_initialization", acc, calls); }
    int nm = 0;
    memcpy(&nm, host + oNm, 4);
    if (f1->n) {
        memcpy(matches12, host + oM12, (size_t)f1->n * sizeof(int));
        memcpy(prev_matched, host + oPrev, (size_t)f1->n * 2 * sizeof(float));
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

int orbm_search_by_projection(int variant, const orbm_frame* frame, const orbm_proj_query* queries,
                              const uint8_t* query_desc, int nq, const uint8_t* query_blocks,
                              uint8_t* occupied, int32_t* assigned, int32_t* best_idx,
                              int32_t* best_dist, int th_dist, float nnratio,
                              int check_orientation, int32_t* nmatches)
{
    ORBX_NVTX("orbm_search_by_projection");
    int rc = check_frame(frame); if (rc) return rc;
    if (variant < ORBM_PROJ_LOCALMAP || variant > ORBM_PROJ_SIM3) { orbx_set_error("unknown SearchByProjection variant %d", variant); return ORBX_E_INVALID; }
    if (nq < 0 || (nq > 0 && (!queries || !query_desc || !best_idx || !best_dist)) || (frame->n > 0 && (!occupied || !assigned))) {
        orbx_set_error("null argument"); return ORBX_E_INVALID;
    }
    rc = require_device(); if (rc) return rc;
    // One device arena, one pinned staging buffer: [inputs][in/out + outputs][work].  Everything the call uploads goes
    // up in ONE copy and everything it returns comes back in ONE copy (a dozen small synchronous copies cost ~8 us each).
    const size_t n = (size_t)std::max(frame->n, 1), Q = (size_t)std::max(nq, 1);
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oCount = take(4), oKps = take(n * sizeof(orbx_keypoint)), oDesc = take(n * 32), oQ = take(Q * sizeof(orbm_proj_query)),
                 oQD = take(Q * 32), oQB = take(Q), oUR = take(n * 4);
    const size_t oOcc = take(n);                       // in/out: last of the upload, first of the download
    const size_t upBytes = off;
    const size_t oNm = take(4), oAsg = take(n * 4), oBI = take(Q * 4), oBD = take(Q * 4);
    const size_t downBytes = off - oOcc;
    // per-frame tracking sizes take the latency form (K9a/K9b: no grid, 32-key lists, fixed-point replay)
    const bool sweep = frame->n <= PROJ_SCAN_MAXN;
    const size_t oGS = take((ORBX_GRID_CELLS + 1) * sizeof(unsigned short)), oGI = take(n * sizeof(unsigned short)),
                 oKeys = take(Q * (sweep ? PROJ_L : PROJ_K) * 8), oKA = take(Q * PROJ_K * 4), oBin = take(Q * (sweep ? 4 : 1));
    DevBuf arena;
    if ((rc = arena.alloc(off))) return rc;
    u8* host = nullptr;
    if ((rc = orbx_scratch_host((void**)&host, oOcc + downBytes))) return rc;
    u8* dev = arena.as<u8>();
    const int nkp = frame->n;
    memcpy(host + oCount, &nkp, 4);
    if (frame->n > 0) {
        memcpy(host + oKps, frame->keys_un, (size_t)frame->n * sizeof(orbx_keypoint));
        memcpy(host + oDesc, frame->descriptors, (size_t)frame->n * 32);
        memcpy(host + oOcc, occupied, (size_t)frame->n);
        if (frame->u_right) memcpy(host + oUR, frame->u_right, (size_t)frame->n * 4);
    }
    if (nq > 0) {
        memcpy(host + oQ, queries, (size_t)nq * sizeof(orbm_proj_query));
        memcpy(host + oQD, query_desc, (size_t)nq * 32);
        if (query_blocks) memcpy(host + oQB, query_blocks, (size_t)nq);
    }
    PhaseTrace pt(orbx_lat_trace_on());
    cudaStream_t st = orbx_call_stream();
    pt.mark("start", st);
    ORBX_CUDA(cudaMemcpyAsync(dev, host, upBytes, cudaMemcpyHostToDevice, st));
    pt.mark("h2d", st);
    FrameStore fs;
    fs.kps = reinterpret_cast<orbx_keypoint*>(dev + oKps); fs.desc = dev + oDesc; fs.count = reinterpret_cast<int*>(dev + oCount);
    fs.gridStart = reinterpret_cast<unsigned short*>(dev + oGS); fs.gridIdx = reinterpret_cast<unsigned short*>(dev + oGI);
    fs.cap = (int)n;
    const GridParams gp = orbx_grid_params(frame->min_x, frame->max_x, frame->min_y, frame->max_y);
    const float* dUR = frame->u_right ? reinterpret_cast<const float*>(dev + oUR) : nullptr;
    const orbm_proj_query* dQ = reinterpret_cast<const orbm_proj_query*>(dev + oQ);
    unsigned long long* dKeys = reinterpret_cast<unsigned long long*>(dev + oKeys);
    float* dKA = reinterpret_cast<float*>(dev + oKA);
    if (sweep) {
        ProjScan ps;
        ps.kps = fs.kps; ps.desc = fs.desc; ps.uRight = dUR; ps.gp = gp; ps.n = frame->n;
        ps.gate = (variant == ORBM_PROJ_LOCALMAP || variant == ORBM_PROJ_LASTFRAME) && dUR != nullptr;
        if (nq > 0)
            proj_list_kernel<<<(nq + PROJ_LIST_WARPS - 1) / PROJ_LIST_WARPS, PROJ_LIST_WARPS * 32, 0, st>>>(ps, dQ, dev + oQD, nq, dev + oOcc, dKeys);
        pt.mark("lists", st);
        const int fixSmem = (int)((((size_t)2 * frame->n * sizeof(int) + 15) & ~(size_t)15) + (PROJ_FIX_THREADS / 32) * PROJ_SCAN_BUF * 8);
        static bool attrSet = false;
        if (!attrSet) { ORBX_CUDA(cudaFuncSetAttribute(proj_fixpoint_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * PROJ_SCAN_MAXN * 4 + 16 + (PROJ_FIX_THREADS / 32) * PROJ_SCAN_BUF * 8)); attrSet = true; }
        proj_fixpoint_kernel<<<1, PROJ_FIX_THREADS, fixSmem, st>>>(ps, dQ, dev + oQD, nq, query_blocks ? dev + oQB : nullptr, dev + oOcc,
                                                               reinterpret_cast<int*>(dev + oAsg), reinterpret_cast<int*>(dev + oBI),
                                                               reinterpret_cast<int*>(dev + oBD), variant, th_dist, nnratio, check_orientation,
                                                               reinterpret_cast<int*>(dev + oNm), dKeys, reinterpret_cast<int*>(dev + oBin), nullptr);
        pt.mark("fixpoint", st);
    } else {
        rc = orbx_launch_grid_build(fs, 0, 1, gp, st); if (rc) return rc;
        pt.mark("grid", st);
        if (nq > 0)
            search_proj_cand_kernel<<<(nq + PROJ_WARPS - 1) / PROJ_WARPS, PROJ_WARPS * 32, 0, st>>>(
                fs, 0, gp, dUR, dQ, dev + oQD, nq, dev + oOcc, variant, dKeys, dKA);
        pt.mark("cand", st);
        const int occSmem = (std::max(frame->n, 1) + 15) & ~15;          // n <= 65000 (check_frame)
        if (occSmem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(search_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, occSmem));
        search_proj_kernel<<<1, 32, occSmem, st>>>(fs, 0, gp, dUR, dQ, dev + oQD, nq, query_blocks ? dev + oQB : nullptr, dev + oOcc,
                                               reinterpret_cast<int*>(dev + oAsg), reinterpret_cast<int*>(dev + oBI),
                                               reinterpret_cast<int*>(dev + oBD), dev + oBin, variant, th_dist, nnratio,
                                               check_orientation, reinterpret_cast<int*>(dev + oNm), dKeys, dKA);
        pt.mark("replay", st);
    }
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(host + oOcc, dev + oOcc, downBytes, cudaMemcpyDeviceToHost, st));
    pt.mark("d2h", st);
    ORBX_CUDA(cudaStreamSynchronize(st));
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; pt.report("search_by_projection", acc, calls); }
    int nm = 0;
    memcpy(&nm, host + oNm, 4);
    if (frame->n > 0) {
        memcpy(occupied, host + oOcc, (size_t)frame->n);
        memcpy(assigned, host + oAsg, (size_t)frame->n * 4);
    }
    if (nq > 0) {
        memcpy(best_idx, host + oBI, (size_t)nq * 4);
        memcpy(best_dist, host + oBD, (size_t)nq * 4);
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}
int orbm_compute_stereo_matches(const orbm_stereo_input* in, float* u_right, float* depth,
                                uint64_t* ncompares)
{
    ORBX_NVTX("orbm_compute_stereo_matches");
    if (!in || in->n_left < 0 || in->n_right < 0 || in->nlevels < 1 || in->nlevels > ORBX_MAX_LEVELS ||
        !in->scale_factors || !in->inv_scale_factors || !in->pyr_left || !in->pyr_right || !in->level_width ||
        !in->level_height || !in->pyr_stride_left || !in->pyr_stride_right ||
        (in->n_left > 0 && (!in->keys_left || !in->desc_left || !u_right || !depth)) ||
        (in->n_right > 0 && (!in->keys_right || !in->desc_right))) {
        orbx_set_error("bad orbm_stereo_input"); return ORBX_E_INVALID;
    }
    if (in->n_left > 65000 || in->n_right > 22000) { orbx_set_error("too many keypoints for the stereo matcher"); return ORBX_E_CAPACITY; }
    int rc = require_device(); if (rc) return rc;
    for (int i = 0; i < in->n_left + in->n_right; ++i) {
        const orbx_keypoint& k = i < in->n_left ? in->keys_left[i] : in->keys_right[i - in->n_left];
        if (k.octave < 0 || k.octave >= in->nlevels) { orbx_set_error("keypoint octave outside [0,nlevels)"); return ORBX_E_INVALID; }
    }
    // one arena: both eyes' keypoints, descriptors and pyramids go up in ONE copy (the levels are packed row by row into the
    // page-locked staging buffer), uRight / depth / the compare count come back in one
    for (int l = 0; l < in->nlevels; ++l) {
        if (in->level_width[l] <= 0 || in->level_height[l] <= 0) { orbx_set_error("empty pyramid level %d", l); return ORBX_E_INVALID; }
        for (int e = 0; e < 2; ++e) {
            const uint8_t* src = e ? in->pyr_right[l] : in->pyr_left[l];
            const int64_t ss = e ? in->pyr_stride_right[l] : in->pyr_stride_left[l];
            if (!src || ss < in->level_width[l]) { orbx_set_error("bad pyramid level %d", l); return ORBX_E_INVALID; }
        }
    }
    const int cap = std::max(std::max(in->n_left, in->n_right), 8);
    const int counts[2] = {in->n_left, in->n_right};
    CallArena ar;
    { static double acc[PhaseTrace::MAXP]; static int calls = 0; ar.trace("compute_stereo_matches", acc, &calls); }
    const int sCount = ar.in(counts, 8), sKps = ar.in(nullptr, (size_t)2 * cap * sizeof(orbx_keypoint)), sDesc = ar.in(nullptr, (size_t)2 * cap * 32);
    int sLv[2][ORBX_MAX_LEVELS];
    for (int l = 0; l < in->nlevels; ++l)
        for (int e = 0; e < 2; ++e) sLv[e][l] = ar.in(nullptr, (size_t)((in->level_width[l] + 15) & ~15) * in->level_height[l]);
    unsigned long long c = 0;
    // the kernel fills uRight / depth up to the store's capacity (max of both eyes), the caller gets the left keypoints' part
    const int sU = ar.out(u_right, (size_t)in->n_left * 4, (size_t)cap * 4), sZ = ar.out(depth, (size_t)in->n_left * 4, (size_t)cap * 4), sC = ar.out(&c, 8);
    const int sS = ar.work((size_t)cap * 4);
    rc = ar.layout(); if (rc) return rc;
    if (in->n_left) {
        memcpy(ar.hostp<orbx_keypoint>(sKps), in->keys_left, (size_t)in->n_left * sizeof(orbx_keypoint));
        memcpy(ar.hostp<u8>(sDesc), in->desc_left, (size_t)in->n_left * 32);
    }
    if (in->n_right) {
        memcpy(ar.hostp<orbx_keypoint>(sKps) + cap, in->keys_right, (size_t)in->n_right * sizeof(orbx_keypoint));
        memcpy(ar.hostp<u8>(sDesc) + (size_t)cap * 32, in->desc_right, (size_t)in->n_right * 32);
    }
    PyrView pv;
    memset(&pv, 0, sizeof(pv));
    pv.nlevels = in->nlevels;
    for (int l = 0; l < in->nlevels; ++l) {
        pv.w[l] = in->level_width[l]; pv.h[l] = in->level_height[l];
        pv.scale[l] = in->scale_factors[l]; pv.invScale[l] = in->inv_scale_factors[l];
        const int pitch = (pv.w[l] + 15) & ~15;
        for (int e = 0; e < 2; ++e) {
            const uint8_t* src = e ? in->pyr_right[l] : in->pyr_left[l];
            const int64_t ss = e ? in->pyr_stride_right[l] : in->pyr_stride_left[l];
            u8* dst = ar.hostp<u8>(sLv[e][l]);
            if (ss == pitch) memcpy(dst, src, (size_t)pitch * pv.h[l] - (pitch - pv.w[l]));
            else for (int y = 0; y < pv.h[l]; ++y) memcpy(dst + (size_t)y * pitch, src + (size_t)y * ss, (size_t)pv.w[l]);
            pv.base[e][l] = ar.dev<u8>(sLv[e][l]); pv.pitch[e][l] = pitch; pv.pairStride[e][l] = 0;
        }
    }
    rc = ar.upload(); if (rc) return rc;
    ORBX_CUDA(cudaMemsetAsync(ar.dev<u8>(sC), 0, 8, ar.st));
    FrameStore fs;
    fs.kps = ar.dev<orbx_keypoint>(sKps); fs.desc = ar.dev<u8>(sDesc); fs.count = ar.dev<int>(sCount);
    fs.gridStart = nullptr; fs.gridIdx = nullptr; fs.cap = cap;
    rc = launch_stereo(fs, 0, 1, pv, in->bf, in->fx, cap, ar.dev<float>(sU), ar.dev<float>(sZ), ar.dev<int>(sS),
                       ar.dev<unsigned long long>(sC), ar.st);
    if (rc) return rc;
    rc = ar.finish(); if (rc) return rc;
    if (ncompares) *ncompares = c;
    return ORBX_OK;
}

int orbx_extract_stereo_batch(orbx_handle* h, const uint8_t* left, const uint8_t* right,
                              size_t frame_stride, size_t row_stride, int width, int height,
                              int npairs, int images_on_device, int first_slot, float bf, float fx,
                              float* u_right, float* depth, int capacity, uint64_t* ncompares)
{
    ORBX_NVTX("orbx_extract_stereo_batch");
    if (!h || !left || !right) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || npairs <= 0) return ORBX_OK;
    if (width < 64 || height < 64) { orbx_set_error("image smaller than 64x64 is outside the extractor's domain"); return ORBX_E_INVALID; }
    if (row_stride < (size_t)width) { orbx_set_error("row_stride < width"); return ORBX_E_INVALID; }
    if (h->chunkFrames < 2) { orbx_set_error("stereo needs a handle with chunk_frames >= 2"); return ORBX_E_CAPACITY; }
    if (first_slot < 0 || first_slot + 2 * npairs > h->resultSlots) { orbx_set_error("slots [%d,%d) exceed result store (%d)", first_slot, first_slot + 2 * npairs, h->resultSlots); return ORBX_E_CAPACITY; }
    if ((u_right || depth) && capacity < h->kpCap) { orbx_set_error("output capacity %d < %d", capacity, h->kpCap); return ORBX_E_CAPACITY; }
    if (!(fx > 0) || !(bf > 0)) { orbx_set_error("bf and fx must be positive"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    int rc = orbx_setup_shape(h, width, height); if (rc) return rc;
    h->storeUndistorted = false;
    rc = orbx_ensure_match_buffers(h); if (rc) return rc;
    const Geom& g = h->geom;
    const bool direct = images_on_device && (row_stride % 4 == 0) && (frame_stride % 4 == 0) &&
                        (((uintptr_t)left) % 4 == 0) && (((uintptr_t)right) % 4 == 0);
    const int pairsPerChunk = h->chunkFrames / 2;
    if (ncompares) ORBX_CUDA(cudaMemsetAsync(h->dNCompares, 0, sizeof(unsigned long long), h->stream));
    if (!direct) {
        cudaEvent_t evStart;                                            // copies must not overtake earlier work on the handle
        ORBX_CUDA(cudaEventCreateWithFlags(&evStart, cudaEventDisableTiming));
        ORBX_CUDA(cudaEventRecord(evStart, h->stream));
        ORBX_CUDA(cudaStreamWaitEvent(h->copyStream, evStart, 0));
        ORBX_CUDA(cudaEventDestroy(evStart));
    }
    int chunkNo = 0;
    for (int p0 = 0, np = 0; p0 < npairs; p0 += np, ++chunkNo) {
        // host images: ramp the first two chunks and taper the tail (as orbx_extract_batch_to_host does), so that less
        // of the first copy and of the last chunk's kernels is exposed
        int cp = pairsPerChunk;
        if (!direct && npairs > pairsPerChunk) cp = chunkNo == 0 ? std::max(1, cp / 4) : (chunkNo == 1 ? std::max(1, cp / 2) : cp);
        np = std::min(cp, npairs - p0);
        if (!direct && npairs > pairsPerChunk && np == npairs - p0 && np > std::max(1, pairsPerChunk / 4))
            np = std::max(std::max(1, pairsPerChunk / 4), (np + 1) / 2);
        Level0 l0;
        l0.stereo = 1;
        if (direct) {
            l0.base = left + (size_t)p0 * frame_stride; l0.base1 = right + (size_t)p0 * frame_stride;
            l0.frameStride = (long long)frame_stride; l0.pitch = (int)row_stride;
        } else {
            // chunk i+1 is copied on the copy stream into the other half of dLevel0 while chunk i computes
            const int b = chunkNo & 1;
            uint8_t* dl = h->dLevel0 + (size_t)b * h->chunkFrames * h->l0FrameBytes;
            uint8_t* dr = dl + (size_t)pairsPerChunk * h->l0FrameBytes;
            if (h->evFreeValid[b]) ORBX_CUDA(cudaStreamWaitEvent(h->copyStream, h->evFree[b], 0));
            uint8_t* sl = h->dStageLin + (size_t)b * h->chunkFrames * h->l0FrameBytes;
            uint8_t* sr = sl + (size_t)pairsPerChunk * h->l0FrameBytes;
            rc = orbx_copy_frames(h, dl, h->l0FrameBytes, sl, (size_t)pairsPerChunk * h->l0FrameBytes, left + (size_t)p0 * frame_stride, frame_stride, row_stride, width, height, np,
                                  images_on_device != 0, h->copyStream, &l0.frameStride);
            if (rc) return rc;
            rc = orbx_copy_frames(h, dr, h->l0FrameBytes, sr, (size_t)pairsPerChunk * h->l0FrameBytes, right + (size_t)p0 * frame_stride, frame_stride, row_stride, width, height, np,
                                  images_on_device != 0, h->copyStream, &l0.frameStride);
            if (rc) return rc;
            ORBX_CUDA(cudaEventRecord(h->evCopied[b], h->copyStream));
            ORBX_CUDA(cudaStreamWaitEvent(h->stream, h->evCopied[b], 0));
            l0.base = dl; l0.base1 = dr; l0.pitch = h->l0Pitch;
        }
        const int slot0 = first_slot + 2 * p0;
        rc = orbx_run_chunk(h, l0, 2 * np, slot0); if (rc) return rc;
        PyrView pv;
        memset(&pv, 0, sizeof(pv));
        pv.nlevels = g.nlevels;
        for (int l = 0; l < g.nlevels; ++l) {
            const LevelGeom& L = g.L[l];
            pv.w[l] = L.w; pv.h[l] = L.h; pv.scale[l] = h->scale[l]; pv.invScale[l] = h->invScale[l];
            for (int e = 0; e < 2; ++e) {
                if (l == 0) { pv.base[e][0] = e ? l0.base1 : l0.base; pv.pairStride[e][0] = l0.frameStride; pv.pitch[e][0] = l0.pitch; }
                else { pv.base[e][l] = h->dRaw + (size_t)e * g.pyrFrameBytes + L.off; pv.pairStride[e][l] = 2 * g.pyrFrameBytes; pv.pitch[e][l] = L.pitch; }
            }
        }
        orbx_prof_mark(h, -1);
        // outputs indexed by pair: pair p of this call lives at row (first_slot/2 + p) of dURight/dDepth
        const size_t row0 = (size_t)(first_slot / 2 + p0) * h->kpCap;
        FrameStore fs = store_of(h);
        // the kernel addresses outputs by blockIdx (pair inside the chunk): pass offset pointers
        rc = launch_stereo(fs, slot0, np, pv, bf, fx, h->kpCap, h->dURight + row0, h->dDepth + row0,
                           reinterpret_cast<int*>(h->dMatches12) + row0, ncompares ? h->dNCompares : nullptr, h->stream);
        if (rc) return rc;
        ORBX_COUNT(h, ORBX_STAGE_STEREO, 2);
        orbx_prof_mark(h, ORBX_STAGE_STEREO);
        if (!direct) { ORBX_CUDA(cudaEventRecord(h->evFree[chunkNo & 1], h->stream)); h->evFreeValid[chunkNo & 1] = true; }
    }
    ORBX_CUDA(cudaGetLastError());
    const size_t row0 = (size_t)(first_slot / 2) * h->kpCap;
    if (u_right) ORBX_CUDA(cudaMemcpy2DAsync(u_right, (size_t)capacity * 4, h->dURight + row0, (size_t)h->kpCap * 4, (size_t)h->kpCap * 4, npairs, cudaMemcpyDeviceToHost, h->stream));
    if (depth) ORBX_CUDA(cudaMemcpy2DAsync(depth, (size_t)capacity * 4, h->dDepth + row0, (size_t)h->kpCap * 4, (size_t)h->kpCap * 4, npairs, cudaMemcpyDeviceToHost, h->stream));
    if (ncompares) ORBX_CUDA(cudaMemcpyAsync(ncompares, h->dNCompares, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    if (u_right || depth || ncompares) return orbx_sync(h);
    return ORBX_OK;
}


} // extern "C"
