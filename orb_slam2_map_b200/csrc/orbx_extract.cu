// orbx_extract.cu — sm_100a kernels for ORBextractor::operator() (reference src/ORBextractor.cc).
//
//   K1a pyr_resize_kernel   cv::resize INTER_LINEAR 8U fixed point      (ORBextractor.cc:1120)
//   K1b blur7_kernel        GaussianBlur 7x7 s=2 REFLECT_101            (ORBextractor.cc:1085-1086)
//   K2  fast_cells_kernel   per-cell FAST 9/16 + NMS + ini/min retry    (ORBextractor.cc:789-829)
//   K3  octree_kernel       DistributeOctTree                           (ORBextractor.cc:539-763)
//   K4  describe_kernel     IC_Angle + computeOrbDescriptor + output    (ORBextractor.cc:77-147,1076-1103)
//
// All kernels are batched over the frames of a chunk; every intermediate lives in the handle's
// device workspace.  Integer arithmetic reproduces OpenCV 4.13 bit for bit;
// float arithmetic uses explicit round-to-nearest intrinsics so nothing is contracted to FMA.

#include "orbx_internal.h"
#include "../../include/orbx_pattern.h"
#include "../../include/orbx_sincos.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <cstdlib>

typedef uint8_t u8;

// =============================================================================================
// K1a: one level of the pyramid = bilinear 8U resize of the previous level (OpenCV fixed point:
// 11-bit coefficients, horizontal sums kept as int, vertical blend ((b*(S>>4))>>16 ... +2)>>2).
// One thread produces 4 horizontally adjacent pixels and stores them as one 32-bit word.
// =============================================================================================
__global__ void __launch_bounds__(256)
pyr_resize_kernel(Level0 l0, int fromL0, const u8* __restrict__ srcBase, long long srcFrameStride, int srcPitch,
                  u8* __restrict__ dstBase, long long dstFrameStride, int dstPitch,
                  int dw, int dh, const int2* __restrict__ cx, const int2* __restrict__ cy)
{
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x4 >= dw || y >= dh) return;
    const u8* src = fromL0 ? l0_frame(l0, blockIdx.z) : srcBase + (long long)blockIdx.z * srcFrameStride;
    u8* dst = dstBase + (long long)blockIdx.z * dstFrameStride;
    const int2 cyv = cy[y];
    const u8* r0 = src + (long long)(cyv.x & 0xffff) * srcPitch;
    const u8* r1 = src + (long long)(cyv.x >> 16) * srcPitch;
    const int b0 = (short)(cyv.y & 0xffff), b1 = cyv.y >> 16;
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = min(x4 + i, dw - 1);
        const int2 c = cx[x];
        const int s0 = c.x & 0xffff, s1 = c.x >> 16;
        const int a0 = (short)(c.y & 0xffff), a1 = c.y >> 16;
        const int S0 = r0[s0] * a0 + r0[s1] * a1;
        const int S1 = r1[s0] * a0 + r1[s1] * a1;
        const int v = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2;
        out |= (uint32_t)(v & 0xff) << (8 * i);
    }
    *reinterpret_cast<uint32_t*>(dst + (long long)y * dstPitch + x4) = out;
}

// =============================================================================================
// K1a (fast path, scale factor <= 2): one warp owns a 128 x th strip of the destination level;
// a lane owns 4 adjacent output columns and walks down the output rows.  Everything that depends
// only on the column (source column, 11-bit coefficients, byte selectors) is hoisted out of the
// row loop.  A source row contributes 3 aligned 32-bit words per lane; the two taps of a pixel are
// picked with byte-permutes and combined with one IDP.2A (u16 x u8 dot product).  The horizontal
// sums of a source row are computed once per strip (consecutive output rows share a source row
// when the scale is < 2).  Vertical blend = OpenCV's
// ((b*(S>>4))>>16 ... +2)>>2 with the >>16 folded into a high multiply.
// =============================================================================================
#define RS_TARGET 12                    // strips are about this tall, balanced over the level (measured 8 .. 48: taller
                                        // strips need more shared memory per warp and lose more to occupancy than the
                                        // amortised column setup gains)
#define RS_WARPS 4

// per-lane column constants: the two taps of output pixel i lie at bytes t, t+1 (0 <= t <= 11) of the lane's
// three-word window.  sel1 picks them out of words 0-1, sel2 replaces those that live in word 2: two PRMTs
// per pixel, no predicates.
struct RsCol { uint32_t coef[4]; uint32_t sel1[4], sel2[4]; };

// The row walk used to be bound by load latency (one source row = 3 words in flight per lane, 60 % of the stall
// samples on the scoreboard).  Now a strip runs in two phases: (1) the horizontal sums of ALL source rows of the strip,
// four rows (12 independent loads) at a time, parked in shared memory as 4 x u16 per lane; (2) the vertical blends
// out of shared memory.  Every source row is still summed once.
__global__ void __launch_bounds__(RS_WARPS * 32, 12)
pyr_resize4_kernel(Level0 l0, int fromL0, const u8* __restrict__ srcBase, long long srcFrameStride, int srcPitch, int sw, int sh,
                   u8* __restrict__ dstBase, long long dstFrameStride, int dstPitch,
                   int dw, int dh, int th, int nsrcMax, const int2* __restrict__ cx, const int2* __restrict__ cy)
{
    extern __shared__ uint2 rsH[];              // [warp][source row of the strip][lane]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int x = (blockIdx.x * 32 + lane) * 4;
    const int y0 = (blockIdx.y * RS_WARPS + wid) * th;
    if (y0 >= dh) return;
    const u8* src = fromL0 ? l0_frame(l0, blockIdx.z) : srcBase + (long long)blockIdx.z * srcFrameStride;
    u8* dst = dstBase + (long long)blockIdx.z * dstFrameStride;
    // ---- per-column constants ------------------------------------------------------------------
    RsCol c;
    const uint32_t* colp;                       // word wb of source row 0: the lane's window is colp[r*rowWords + 0..2]
    {
        const int lastWord = (sw - 1) >> 2;     // >= 2: levels are at least 12 pixels wide (host check)
        const int2 c0 = cx[min(x, dw - 1)];
        const int wb = max(0, min((c0.x & 0xffff) >> 2, lastWord - 2));
        colp = reinterpret_cast<const uint32_t*>(src) + wb;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int2 ci = cx[min(x + i, dw - 1)];
            const int t0 = (ci.x & 0xffff) - 4 * wb, t1 = min(t0 + 1, 11);       // 0 .. 11 for scale <= 2
            c.coef[i] = (uint32_t)ci.y;                                          // a0 | a1 << 16
            c.sel1[i] = (uint32_t)(t0 & 7) | ((uint32_t)(t1 & 7) << 4);
            c.sel2[i] = (uint32_t)(t0 < 8 ? 0 : t0 - 4) | ((uint32_t)(t1 < 8 ? 1 : t1 - 4) << 4);
        }
    }
    const int rowWords = srcPitch >> 2;
    const int yend = min(y0 + th, dh);
    const int sFirst = cy[y0].x & 0xffff;
    const int nsrc = min((cy[yend - 1].x >> 16) - sFirst + 1, nsrcMax);
    // sum table of this warp: [source row][element 0..3][lane] as 16-bit values: a warp's access to one element of one row is 64
    // contiguous bytes (conflict-free as LDS.U16 / STS.U16; 8-byte records per lane cost two wavefronts per access)
    unsigned short* Hs = reinterpret_cast<unsigned short*>(rsH + (size_t)wid * nsrcMax * 32) + lane;
    // ---- phase 1: horizontal sums of the strip's source rows, four rows (12 loads) in flight.  nsrcMax is a multiple of 4: the
    // rows past nsrc of the last group are computed from a clamped source row and stored, never read -----------------------
    {
        const uint32_t* p = colp + (long long)sFirst * rowWords;
        const int lim = sh - 1 - sFirst;                     // last source row of the image, relative to the strip
        unsigned short* hp = Hs;
        for (int r = 0; r < nsrc; r += 4, hp += 4 * 128) {
            uint32_t w[4][3];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t* q = p + min(r + u, lim) * rowWords;
                w[u][0] = __ldg(q); w[u][1] = __ldg(q + 1); w[u][2] = __ldg(q + 2);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t taps = __byte_perm(__byte_perm(w[u][0], w[u][1], c.sel1[i]), w[u][2], c.sel2[i]);
                    hp[u * 128 + i * 32] = (unsigned short)(__dp2a_lo(c.coef[i], taps, 0u) >> 4);      // (p0*a0 + p1*a1) >> 4 < 2^15
                }
            }
        }
    }
    __syncwarp();
    // ---- phase 2: vertical blends ((b0*SA >> 16) + (b1*SB >> 16) + 2) >> 2.  The row constants (offsets of the two source rows
    // in the sum table, coefficients) are computed lane-parallel, one output row per lane, and broadcast by shuffle; the sums
    // are read back as 16-bit values (no masks); each product is one IMAD.HI that also carries the running sum (the +2 rides
    // in as the first addend); two results are packed by an IMAD, shifted as a pair and byte-selected: 3 ALU-pipe
    // instructions per 4 pixels, the rest on the FMA and load/store pipes ---------------------------------------------------
    const unsigned short* Hb = Hs;
    u8* dp = dst + (long long)y0 * dstPitch + x;
    for (int yb = y0; yb < yend; yb += 32) {
        uint32_t offs = 0, coefs = 0;
        if (yb + lane < yend) {
            const int2 cyv = cy[yb + lane];
            const int ia = min((cyv.x & 0xffff) - sFirst, nsrc - 1), ib = min((cyv.x >> 16) - sFirst, nsrc - 1);
            offs = (uint32_t)(ia * 128) | ((uint32_t)(ib * 128) << 16);             // in 16-bit elements: row * 32 lanes * 4
            coefs = (uint32_t)cyv.y;                                                // b0 | b1 << 16
        }
        const int rows = min(32, yend - yb);
        for (int j = 0; j < rows; ++j, dp += dstPitch) {
            const uint32_t o = __shfl_sync(0xffffffffu, offs, j), cf = __shfl_sync(0xffffffffu, coefs, j);
            const unsigned short* A = Hb + (o & 0xffffu);
            const unsigned short* B = Hb + (o >> 16);
            const uint32_t b0 = cf << 16, b1 = cf & 0xffff0000u;                    // coefficient << 16: umulhi gives (b * S) >> 16
            const uint32_t s0 = __umulhi(b1, B[0]) + (__umulhi(b0, A[0]) + 2u);
            const uint32_t s1 = __umulhi(b1, B[32]) + (__umulhi(b0, A[32]) + 2u);
            const uint32_t s2 = __umulhi(b1, B[64]) + (__umulhi(b0, A[64]) + 2u);
            const uint32_t s3 = __umulhi(b1, B[96]) + (__umulhi(b0, A[96]) + 2u);
            // s <= 1022: (s >> 2) fits a byte, so the pair shift cannot leak between the halves' result bytes
            const uint32_t p01 = (s1 * 65536u + s0) >> 2, p23 = (s3 * 65536u + s2) >> 2;
            if (x < dw) *reinterpret_cast<uint32_t*>(dp) = __byte_perm(p01, p23, 0x6420);
        }
    }
}

// =============================================================================================
// K1b: 7x7 Gaussian blur, kernel [18 34 48 56 48 34 18]/256 per axis, exact integer, reflect-101
// of the level itself.  One warp owns a 128 x 32 strip: a lane owns 4 adjacent columns and walks
// down the rows.  Per row it loads ONE aligned 32-bit word (the warp reads 128 contiguous bytes),
// takes the neighbouring words from the adjacent lanes by shuffle, forms the horizontal 7-tap sums
// with byte-permutes + IDP.4A dot products, and keeps the last 7 rows of sums in registers for the
// vertical pass — no shared memory, every source byte is fetched once per strip (+6 halo rows).
// =============================================================================================
#define BL_TW 120                       // output columns per strip: lanes 1..30 (lanes 0, 31 are halo)
#define BL_TARGET 120                   // strips are about this tall (LevelGeom::blurTH = 7g-6 rows, balanced over the level)
#define BL_WARPS 4
__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * (n - 1) - i;
    return i;
}

__device__ __forceinline__ int reflect_row(int y, int h)
{
    y = y < 0 ? -y : (y >= h ? 2 * (h - 1) - y : y);
    return max(0, min(y, h - 1));
}

// BORDER_REFLECT_101 of the columns is done in registers: the lane whose word holds the last
// column e = w-1 (at byte be) and the lane after it rebuild their words from their left
// neighbours' words with one byte-permute (selectors below); the lane at x = -4 mirrors columns
// 4,3,2,1 out of its right neighbours.  mode: 0 none, 1 = holds e, 2 = lane after, 3 = x < 0.
__device__ __forceinline__ void edge_mode(int x, int w, int& mode, uint32_t& sel)
{
    const int e = w - 1;
    mode = 0; sel = 0x3210;
    if (x < 0) { mode = 3; sel = 0x1234; }
    else if (e >= x && e <= x + 3) {
        const int be = e - x;
        mode = be == 3 ? 0 : 1;
        sel = be == 2 ? 0x5654u : (be == 1 ? 0x3454u : 0x1234u);       // source pair (left word, own word)
    } else if (e >= x - 4 && e < x) {
        const int be = e - (x - 4);
        mode = be == 0 ? 0 : 2;
        sel = be == 3 ? 0x3456u : (be == 2 ? 0x1234u : 0x0012u);       // source pair (2nd left word, left word)
    }
}

// INTERIOR: every source row of the strip (y0-3 .. y0-4+7*ngroups) exists, no row reflection needed.
// EDGE: some lane of the warp holds (or follows) the last column or sits at x = -4 and rebuilds its word from its
// neighbours; strips that lie inside the row skip those four shuffles.
template <bool INTERIOR>
__device__ __forceinline__ void blur_strip(const u8* __restrict__ src, int sp, u8* __restrict__ dst, int dp,
                                           int w, int h, int x, int y0, int th, int lane, int mode, uint32_t sel, bool EDGE)
{
    const unsigned FULL = 0xffffffffu;
    const uint32_t c0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), c1 = 48u | (34u << 8) | (18u << 16);
    const int rowsOut = min(th, h - y0);
    const int ngroups = (rowsOut + 6 + 6) / 7;
    const int xl = min(max(x, 0), (w - 1) & ~3);  // every lane loads an aligned word that lies inside the row
    const bool store = lane >= 1 && lane <= 30 && x < w;
    // Q[k][j]: horizontal sums of rows (r-1, r) of column j packed as two u16 (each < 2^16), r = the row filed in slot k
    uint32_t Q[7][4], Hprev[4] = {0u, 0u, 0u, 0u};
    uint32_t cur[7], nxt[7];
    const u8* colp = src + (long long)(y0 - 3) * sp + xl;         // row y0-3 of the lane's column (INTERIOR addressing)
    u8* dcol = dst + x;                                           // kept in registers: the row address is one IMAD.WIDE
    asm volatile("" : "+l"(dcol));
#pragma unroll
    for (int k = 0; k < 7; ++k)
        cur[k] = INTERIOR ? *reinterpret_cast<const uint32_t*>(colp + k * sp)
                          : *reinterpret_cast<const uint32_t*>(src + (long long)reflect_row(y0 - 3 + k, h) * sp + xl);
#pragma unroll 1
    for (int gi = 0; gi < ngroups; ++gi) {
        if (gi + 1 < ngroups) {
#pragma unroll
            for (int k = 0; k < 7; ++k)
                nxt[k] = INTERIOR ? *reinterpret_cast<const uint32_t*>(colp + ((gi + 1) * 7 + k) * sp)
                                  : *reinterpret_cast<const uint32_t*>(src + (long long)reflect_row(y0 - 3 + (gi + 1) * 7 + k, h) * sp + xl);
        }
        const bool later = gi > 0;                                // rows r >= 7: a full window exists
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const int r = gi * 7 + k;
            uint32_t w1 = cur[k];
            if (EDGE) {
                const uint32_t d1 = __shfl_down_sync(FULL, w1, 1), d2 = __shfl_down_sync(FULL, w1, 2);
                const uint32_t u1 = __shfl_up_sync(FULL, w1, 1), u2 = __shfl_up_sync(FULL, w1, 2);
                const uint32_t pa = mode == 3 ? d1 : (mode == 2 ? u2 : u1), pb = mode == 3 ? d2 : (mode == 2 ? u1 : w1);
                if (mode != 0) w1 = __byte_perm(pa, pb, sel);
            }
            const uint32_t w0 = __shfl_up_sync(FULL, w1, 1), w2 = __shfl_down_sync(FULL, w1, 1);
            uint32_t Hn[4];
            Hn[0] = __dp4a(__byte_perm(w0, w1, 0x4321), c0, __dp4a(__byte_perm(w1, w2, 0x4321), c1, 0u));
            Hn[1] = __dp4a(__byte_perm(w0, w1, 0x5432), c0, __dp4a(__byte_perm(w1, w2, 0x5432), c1, 0u));
            Hn[2] = __dp4a(__byte_perm(w0, w1, 0x6543), c0, __dp4a(__byte_perm(w1, w2, 0x6543), c1, 0u));
            Hn[3] = __dp4a(w1, c0, __dp4a(w2, c1, 0u));
#pragma unroll
            for (int j = 0; j < 4; ++j) { Q[k][j] = __byte_perm(Hprev[j], Hn[j], 0x5410); Hprev[j] = Hn[j]; }
            // window rows r-6 .. r: the pairs (r-6,r-5), (r-4,r-3), (r-2,r-1) are in slots (k+2)%7, (k+4)%7, (k+6)%7, row r
            // is Hn; output row y = y0 + r - 6 (r >= 6 <=> gi > 0 or k == 6).  Three IDP.2A + one IMAD per pixel.
            const int y = y0 + r - 6;
            if ((later || k == 6) && (INTERIOR || y < h)) {
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    o[j] = __dp2a_lo(Q[(k + 2) % 7][j], 18u | (34u << 8),
                           __dp2a_lo(Q[(k + 4) % 7][j], 48u | (56u << 8),
                           __dp2a_lo(Q[(k + 6) % 7][j], 48u | (34u << 8), 18u * Hn[j] + 32768u)));       // result in bits 16..23
                const uint32_t lo = __byte_perm(o[0], o[1], 0x0062), hi = __byte_perm(o[2], o[3], 0x0062);
                if (store) *reinterpret_cast<uint32_t*>(dcol + (unsigned long long)(unsigned)y * (unsigned)dp) = __byte_perm(lo, hi, 0x5410);
            }
        }
#pragma unroll
        for (int k = 0; k < 7; ++k) cur[k] = nxt[k];
    }
}

__global__ void __launch_bounds__(BL_WARPS * 32)
blur7_kernel(Geom g, Level0 l0, const u8* __restrict__ rawBase, u8* __restrict__ blurBase,
             const int* __restrict__ tiles, int ntiles)
{
    const int lane = threadIdx.x & 31;
    const int ti = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
    if (ti >= ntiles) return;
    const int t = tiles[ti];
    const int level = t >> 24, ty = (t >> 12) & 0xfff, tx = t & 0xfff;
    const LevelGeom& L = g.L[level];
    const int frame = blockIdx.y;
    const u8* src;
    int sp;
    if (level == 0) { src = l0_frame(l0, frame); sp = l0.pitch; }
    else { src = rawBase + (long long)frame * g.pyrFrameBytes + L.off; sp = L.pitch; }
    u8* dst = blurBase + (long long)frame * g.pyrFrameBytes + L.off;
    const int y0 = ty * L.blurTH, ngroups = (min(L.blurTH, L.h - y0) + 6 + 6) / 7;
    const int x = tx * BL_TW - 4 + lane * 4;
    int mode; uint32_t sel;
    edge_mode(x, L.w, mode, sel);
    const bool interior = y0 >= 3 && y0 - 3 + 7 * ngroups <= L.h;
    const bool edge = __any_sync(0xffffffffu, mode != 0);         // warp-uniform
    if (interior) blur_strip<true>(src, sp, dst, L.pitch, L.w, L.h, x, y0, L.blurTH, lane, mode, sel, edge);
    else blur_strip<false>(src, sp, dst, L.pitch, L.w, L.h, x, y0, L.blurTH, lane, mode, sel, edge);
}

// =============================================================================================
// K2: FAST per cell.  One warp = one cell = one cv::FAST(sub-image) call of the reference,
// including its private 3x3 non-max suppression and the iniThFAST -> minThFAST retry.
// A corner's strength S (largest t with a 9-arc all beyond t) does not depend on the threshold,
// and NMS(th=ini) == NMS(th=min) restricted to S>ini, so one pass at minTh
// serves both calls.  Candidates are packed x:12 | y:12 | (S-1):8 in window coordinates.
// =============================================================================================
#define FAST_WARPS 1

__device__ __forceinline__ uint32_t vmaxu2(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
__device__ __forceinline__ uint32_t vminu2(uint32_t a, uint32_t b) { return __vminu2(a, b); }
// three-input packed min / max: one VIMNMX3.U16x2 each on sm_100a
__device__ __forceinline__ uint32_t vmax3u2(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t vmin3u2(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }

// S for the pixel at p (row pitch tp): max over the 16 arcs of 9 contiguous ring pixels of
// max(min(v - r), min(r - v)) = max(v - min_arcs(max r), max_arcs(min r) - v).
// Ring elements are packed two per register (u16x2: elements 2j, 2j+1); window maxima / minima grow
// 2 -> 6 -> 9 elements with three-input operations.
__device__ __forceinline__ int fast_strength(const u8* p, int tp)
{
    const int v = p[0];
    uint32_t P[8];
    P[0] = p[3 * tp]      | (p[3 * tp + 1] << 16);
    P[1] = p[2 * tp + 2]  | (p[tp + 3] << 16);
    P[2] = p[3]           | (p[-tp + 3] << 16);
    P[3] = p[-2 * tp + 2] | (p[-3 * tp + 1] << 16);
    P[4] = p[-3 * tp]     | (p[-3 * tp - 1] << 16);
    P[5] = p[-2 * tp - 2] | (p[-tp - 3] << 16);
    P[6] = p[-3]          | (p[tp - 3] << 16);
    P[7] = p[2 * tp - 2]  | (p[3 * tp - 1] << 16);
    uint32_t mx2[8], mn2[8];                      // windows of 2 starting at elements 2j | 2j+1
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t q = __byte_perm(P[j], P[(j + 1) & 7], 0x5432);   // elements 2j+1, 2j+2
        mx2[j] = vmaxu2(P[j], q);
        mn2[j] = vminu2(P[j], q);
    }
    uint32_t mx6[8], mn6[8];                      // windows of 6
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mx6[j] = vmax3u2(mx2[j], mx2[(j + 1) & 7], mx2[(j + 2) & 7]);
        mn6[j] = vmin3u2(mn2[j], mn2[(j + 1) & 7], mn2[(j + 2) & 7]);
    }
    uint32_t mx9[8], mn9[8];                      // windows of 9: six + two + one
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mx9[j] = vmax3u2(mx6[j], mx2[(j + 3) & 7], P[(j + 4) & 7]);
        mn9[j] = vmin3u2(mn6[j], mn2[(j + 3) & 7], P[(j + 4) & 7]);
    }
    const uint32_t lo = vmin3u2(vmin3u2(mx9[0], mx9[1], mx9[2]), vmin3u2(mx9[3], mx9[4], mx9[5]), vminu2(mx9[6], mx9[7]));   // min over arcs of (max over arc)
    const uint32_t hi = vmax3u2(vmax3u2(mn9[0], mn9[1], mn9[2]), vmax3u2(mn9[3], mn9[4], mn9[5]), vmaxu2(mn9[6], mn9[7]));   // max over arcs of (min over arc)
    const int minArcMax = min(lo & 0xffff, lo >> 16);
    const int maxArcMin = max(hi & 0xffff, hi >> 16);
    return max(v - minArcMax, maxArcMin - v);
}

__global__ void __launch_bounds__(FAST_WARPS * 32)
fast_cells_kernel(Geom g, const Cell* __restrict__ cells, Level0 l0, const u8* __restrict__ rawBase,
                  int nframes, int minThArg, int iniThArg, uint32_t* __restrict__ cand,
                  int2* __restrict__ cellSeg,
                  int tilePitch, int scorePitch, int warpSmem, int cellBeg, int cellEnd)
{
    extern __shared__ __align__(16) u8 smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int frame = blockIdx.y, ci = cellBeg + blockIdx.x * FAST_WARPS + wid;
    if (ci >= cellEnd) return;
    const Cell c = cells[ci];
    const LevelGeom& L = g.L[c.level];
    const u8* img;
    int pitch;
    if (c.level == 0) { img = l0_frame(l0, frame); pitch = l0.pitch; }
    else { img = rawBase + (long long)frame * g.pyrFrameBytes + L.off; pitch = L.pitch; }

    u8* tile = smem + (size_t)wid * warpSmem;
    u8* score = tile + (((size_t)tilePitch * g.chMax + 15) & ~(size_t)15);
    const int scoreBytes = (scorePitch * (g.chMax - 4) + 15) & ~15;
    uint32_t* rowMask = reinterpret_cast<uint32_t*>(score + scoreBytes);    // 64 bits per detectable row
    const int zeroVecs = (scoreBytes + 8 * (g.chMax - 6) + 15) >> 4;
    unsigned short* list = reinterpret_cast<unsigned short*>(score + 16 * zeroVecs);

    const int cw = c.cw, ch = c.ch;
    const int dw = cw - 6, dh = ch - 6;               // detectable interior
    const unsigned FULL = 0xffffffffu;
    int nOut = 0;
    unsigned base = 0;
    if (dw > 0 && dh > 0) {
        // ---- load the sub-image so that its column 0 lands on byte 0 of a tile row: two aligned 32-bit loads +
        // funnel shift per word.  A cell starts at an arbitrary byte (x0 = 16 + j*wCell) and a u8 tensor-map box must
        // start on a 16-byte boundary (scripts/probes/probe_tma_u8.cu: any other inner coordinate is an illegal
        // instruction), so TMA cannot deliver the re-aligned tile.  Lanes = (row of a group, word column) with the
        // split done once; the loop only advances two pointers, four row groups (8 loads per lane) in flight. -----
        const int tw = tilePitch >> 2;
        {
            const int xa = c.x0 & ~3, sh8 = 8 * (c.x0 - xa);
            const int nw = (cw + 3) >> 2;                               // words per tile row (<= 17)
            const int rpg = nw <= 8 ? 4 : (nw <= 10 ? 3 : (nw <= 16 ? 2 : 1));      // rows per group of 32 lanes
            const int lr = lane / nw, lw = lane - lr * nw;
            const bool act = lr < rpg;
            const uint32_t* q = reinterpret_cast<const uint32_t*>(img + (long long)(c.y0 + lr) * pitch + xa) + lw;
            uint32_t* d = reinterpret_cast<uint32_t*>(tile) + lr * tw + lw;
            const int qstep = rpg * (pitch >> 2), dstep = rpg * tw;
            for (int y0 = 0; y0 < ch; y0 += 4 * rpg) {
                uint32_t lo[4], hi[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (act && y0 + k * rpg + lr < ch) { lo[k] = __ldg(q + k * qstep); hi[k] = __ldg(q + k * qstep + 1); }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (act && y0 + k * rpg + lr < ch) d[k * dstep] = __funnelshift_r(lo[k], hi[k], sh8);
                q += 4 * qstep; d += 4 * dstep;
            }
        }
        // zero the score tile ((dh+2) x scorePitch)
        // (score and rowMask are contiguous and 16-byte aligned: one run of 16-byte stores covers both)
        for (int i = lane; i < zeroVecs; i += 32) reinterpret_cast<uint4*>(score)[i] = make_uint4(0, 0, 0, 0);
        __syncwarp();
        // The reference calls cv::FAST(iniThFAST) and, only when that returns nothing, cv::FAST(minThFAST)
        // (ORBextractor.cc:809-816).  Same here: most cells are done after the cheap high-threshold pass.
        int n3 = 0;
#pragma unroll 1
        for (int attempt = 0; attempt < 2 && n3 == 0; ++attempt) {
        const int minTh = attempt == 0 ? iniThArg : minThArg;
        // ---- stage 1a: compass pre-test on 8 pixels per lane (byte-packed).  A 9-arc contains two
        // adjacent compass points, so |v - r| > th must hold for one of {N,S} and one of {E,W}.
        // |v-r| per byte = VABSDIFF4; "> t" per byte via the 7-bit carry trick (t < 127).  Each lane
        // drops its 8 flags as ONE byte of the row's 64-bit mask; the ordered candidate list is then
        // expanded from the masks once per cell (lanes = rows), not once per iteration.
        int n1 = 0;
        {
            const int ng = (dw + 7) >> 3;                       // 8-pixel groups per row (<= 8: dw < 64)
            const int gshift = ng <= 4 ? 2 : 3;
            const int lg = lane & ((1 << gshift) - 1), ly = lane >> gshift, rstep = 32 >> gshift;
            const uint32_t T = (uint32_t)(0x7f - min(minTh, 0x7e)) * 0x01010101u;
            const bool packedOk = minTh < 0x7f;
            // low thresholds let a quarter of the pixels through the compass test: two more diameters in the packed test halve that
            // (an arc of 9 holds one end of EVERY diameter) and with it the per-candidate work behind; at high thresholds the
            // handful of candidates does not pay for it
            const bool diag = packedOk && minTh <= 10;
            uint32_t vmask0 = 0, vmask1 = 0;                    // valid pixels of this group
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (8 * lg + j < dw) vmask0 |= 0x80u << (8 * j);
                if (8 * lg + 4 + j < dw) vmask1 |= 0x80u << (8 * j);
            }
            u8* maskBytes = reinterpret_cast<u8*>(rowMask);
            for (int y0 = 0; y0 < dh; y0 += rstep) {
                const int y = y0 + ly;
                if (y < dh) {
                    // pixel x of the group has its centre at byte 8*lg + 3 + x of tile row y+3
                    const uint32_t* rc = reinterpret_cast<const uint32_t*>(tile) + (y + 3) * tw + 2 * lg;
                    const uint2 ta = *reinterpret_cast<const uint2*>(rc), tb = *reinterpret_cast<const uint2*>(rc + 2);
                    const uint2 na = *reinterpret_cast<const uint2*>(rc + 3 * tw), sa = *reinterpret_cast<const uint2*>(rc - 3 * tw);
                    const uint32_t nb2 = rc[3 * tw + 2], sb2 = rc[-3 * tw + 2];
                    const uint32_t V0 = __byte_perm(ta.x, ta.y, 0x6543), V1 = __byte_perm(ta.y, tb.x, 0x6543);
                    const uint32_t E0 = __byte_perm(ta.y, tb.x, 0x5432), E1 = __byte_perm(tb.x, tb.y, 0x5432);
                    const uint32_t N0 = __byte_perm(na.x, na.y, 0x6543), N1 = __byte_perm(na.y, nb2, 0x6543);
                    const uint32_t S0 = __byte_perm(sa.x, sa.y, 0x6543), S1 = __byte_perm(sa.y, sb2, 0x6543);
                    uint32_t p0 = vmask0, p1 = vmask1;
                    if (packedOk) {
                        const uint32_t dn0 = __vabsdiffu4(N0, V0), de0 = __vabsdiffu4(E0, V0), ds0 = __vabsdiffu4(S0, V0), dw0 = __vabsdiffu4(ta.x, V0);
                        const uint32_t dn1 = __vabsdiffu4(N1, V1), de1 = __vabsdiffu4(E1, V1), ds1 = __vabsdiffu4(S1, V1), dw1 = __vabsdiffu4(ta.y, V1);
                        const uint32_t en0 = ((dn0 & 0x7f7f7f7fu) + T) | dn0, ee0 = ((de0 & 0x7f7f7f7fu) + T) | de0;
                        const uint32_t es0 = ((ds0 & 0x7f7f7f7fu) + T) | ds0, ew0 = ((dw0 & 0x7f7f7f7fu) + T) | dw0;
                        const uint32_t en1 = ((dn1 & 0x7f7f7f7fu) + T) | dn1, ee1 = ((de1 & 0x7f7f7f7fu) + T) | de1;
                        const uint32_t es1 = ((ds1 & 0x7f7f7f7fu) + T) | ds1, ew1 = ((dw1 & 0x7f7f7f7fu) + T) | dw1;
                        p0 &= (en0 | es0) & (ee0 | ew0);
                        p1 &= (en1 | es1) & (ee1 | ew1);
                        if (diag) {
                            // the two diagonal diameters (+-2, +-2) of the ring, same test: rows y+3+-2, bytes centre +- 2.
                            // (the "-2" word of pixels 4..7 is the "+2" word of pixels 0..3)
                            const uint2 ua = *reinterpret_cast<const uint2*>(rc + 2 * tw), ub = *reinterpret_cast<const uint2*>(rc + 2 * tw + 2);
                            const uint2 la = *reinterpret_cast<const uint2*>(rc - 2 * tw), lb = *reinterpret_cast<const uint2*>(rc - 2 * tw + 2);
                            const uint32_t uM0 = __byte_perm(ua.x, ua.y, 0x4321), uP0 = __byte_perm(ua.y, ub.x, 0x4321), uP1 = __byte_perm(ub.x, ub.y, 0x4321);
                            const uint32_t lM0 = __byte_perm(la.x, la.y, 0x4321), lP0 = __byte_perm(la.y, lb.x, 0x4321), lP1 = __byte_perm(lb.x, lb.y, 0x4321);
                            auto far = [T](uint32_t r, uint32_t v) { const uint32_t d = __vabsdiffu4(r, v); return ((d & 0x7f7f7f7fu) + T) | d; };
                            p0 &= (far(uP0, V0) | far(lM0, V0)) & (far(lP0, V0) | far(uM0, V0));
                            p1 &= (far(uP1, V1) | far(lP0, V1)) & (far(lP1, V1) | far(uP0, V1));
                        }
                    }
                    // gather the four 0x80 flags of each word into a nibble (multiply trick: no carries collide)
                    const uint32_t m8 = (((p0 >> 7) * 0x01020408u) >> 24) | ((((p1 >> 7) * 0x01020408u) >> 20) & 0xf0u);
                    maskBytes[y * 8 + lg] = (u8)m8;
                }
            }
            __syncwarp();
            // expand: lane = row (two rows per lane when the cell is taller than 32), row-major order
            for (int yb = 0; yb < dh; yb += 32) {
                const int y = yb + lane;
                uint32_t m0 = 0, m1 = 0;
                if (y < dh) { m0 = rowMask[2 * y]; m1 = rowMask[2 * y + 1]; }
                const int cnt = __popc(m0) + __popc(m1);
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
                int pos = n1 + incl - cnt;
                while (m0) { const int b = __ffs(m0) - 1; m0 &= m0 - 1; list[pos++] = (unsigned short)((y << 8) | b); }
                while (m1) { const int b = __ffs(m1) - 1; m1 &= m1 - 1; list[pos++] = (unsigned short)((y << 8) | (32 + b)); }
                n1 += __shfl_sync(FULL, incl, 31);
            }
        }
        __syncwarp();
        // ---- stage 1b: signed opposite-pair bound.  A 9-arc holds at least one pixel of every
        // diameter (k, k+8), so S <= max(min_k max(r_k, r_k+8) - v, v - max_k min(r_k, r_k+8)).
        {
            int nb = 0;
            for (int i0 = 0; i0 < n1; i0 += 32) {
                const int i = i0 + lane;
                bool keep = false;
                unsigned short e = 0;
                if (i < n1) {
                    e = list[i];
                    const int y = e >> 8, x = e & 0xff;
                    const u8* p = tile + (y + 3) * tilePitch + x + 3;
                    const int tp = tilePitch;
                    const int v = p[0];
                    const uint32_t a0 = p[3 * tp] | (p[3 * tp + 1] << 16), b0 = p[-3 * tp] | (p[-3 * tp - 1] << 16);
                    const uint32_t a1 = p[2 * tp + 2] | (p[tp + 3] << 16), b1 = p[-2 * tp - 2] | (p[-tp - 3] << 16);
                    const uint32_t a2 = p[3] | (p[-tp + 3] << 16), b2 = p[-3] | (p[tp - 3] << 16);
                    const uint32_t a3 = p[-2 * tp + 2] | (p[-3 * tp + 1] << 16), b3 = p[2 * tp - 2] | (p[3 * tp - 1] << 16);
                    const uint32_t mx = vminu2(vmin3u2(vmaxu2(a0, b0), vmaxu2(a1, b1), vmaxu2(a2, b2)), vmaxu2(a3, b3));
                    const uint32_t mn = vmaxu2(vmax3u2(vminu2(a0, b0), vminu2(a1, b1), vminu2(a2, b2)), vminu2(a3, b3));
                    const int A = min(mx & 0xffff, mx >> 16), B = max(mn & 0xffff, mn >> 16);
                    keep = max(A - v, v - B) > minTh;
                }
                const unsigned m = __ballot_sync(FULL, keep);
                __syncwarp();
                if (keep) list[nb + __popc(m & ((1u << lane) - 1))] = e;
                nb += __popc(m);
                __syncwarp();
            }
            n1 = nb;
        }
        __syncwarp();
        // ---- stage 2: exact strength, keep corners (S > minTh) -------------------------------
        int n2 = 0;
        for (int i0 = 0; i0 < n1; i0 += 32) {
            const int i = i0 + lane;
            bool corner = false;
            unsigned short e = 0;
            if (i < n1) {
                e = list[i];
                const int y = e >> 8, x = e & 0xff;
                const int S = fast_strength(tile + (y + 3) * tilePitch + x + 3, tilePitch);
                if (S > minTh) { corner = true; score[(y + 1) * scorePitch + x + 1] = (u8)S; }
            }
            __syncwarp();
            const unsigned m = __ballot_sync(FULL, corner);
            if (corner) list[n2 + __popc(m & ((1u << lane) - 1))] = e;
            n2 += __popc(m);
        }
        __syncwarp();
        // ---- stage 3: 3x3 non-max suppression inside the cell --------------------------------
        for (int i0 = 0; i0 < n2; i0 += 32) {
            const int i = i0 + lane;
            bool keep = false;
            unsigned short e = 0;
            if (i < n2) {
                e = list[i];
                const int y = e >> 8, x = e & 0xff;
                const u8* s = score + (y + 1) * scorePitch + x + 1;
                const int v = s[0];
                keep = v > s[-1] && v > s[1] && v > s[-scorePitch - 1] && v > s[-scorePitch] &&
                       v > s[-scorePitch + 1] && v > s[scorePitch - 1] && v > s[scorePitch] &&
                       v > s[scorePitch + 1];
            }
            __syncwarp();
            const unsigned m = __ballot_sync(FULL, keep);
            if (keep) list[n3 + __popc(m & ((1u << lane) - 1))] = e;
            n3 += __popc(m);
        }
        __syncwarp();
        }   // attempt
        // ---- stage 4: emit, order = row-major --------------------------------------------------
        nOut = n3;
        if (nOut > 0) {
            // every cell owns a fixed segment of its level's candidate region (cellCap entries: no two NMS survivors are
            // adjacent, so at most ceil(dw/2)*ceil(dh/2) of them exist): no counter, no atomic, no round trip
            base = (unsigned)(ci - L.cellBeg) * (unsigned)L.cellCap;
            uint32_t* out = cand + (size_t)frame * g.candTotal + L.candOff + base;
            int pos = 0;
            for (int i0 = 0; i0 < n3; i0 += 32) {
                const int i = i0 + lane;
                bool emit = false;
                uint32_t packed = 0;
                if (i < n3) {
                    const unsigned short e = list[i];
                    const int y = e >> 8, x = e & 0xff;
                    const int S = score[(y + 1) * scorePitch + x + 1];
                    emit = true;
                    packed = (uint32_t)(x + 3 + c.offx) | ((uint32_t)(y + 3 + c.offy) << 12) | ((uint32_t)(S - 1) << 24);
                }
                const unsigned m = __ballot_sync(FULL, emit);
                if (emit) out[pos + __popc(m & ((1u << lane) - 1))] = packed;
                pos += __popc(m);
            }
        }
    }
    if (lane == 0) cellSeg[(size_t)frame * g.ncells + ci] = make_int2((int)base, nOut);
}

// =============================================================================================
// K3: DistributeOctTree, re-expressed as deterministic parallel quadtree splitting.  One CTA per
// (frame, level).  The reference's std::list is kept as a REVERSED array of alive nodes
// (push_front == append), rebuilt once per pass:  [survivors in order | children in split order].
// Keys never move: every key carries the index of its node; a pass is
//   count   key-parallel: quadrant of the key inside its (splitting) node -> 4 counters per node
//   order   node-parallel: which nodes split and where their children go.  Breadth-first pass:
//           every node with > 1 key, list order (= descending array index).  "Careful" pass
//           (ORBextractor.cc:673-738): nodes sorted by (nKeys, creation index) descending, split
//           until the node count reaches N — the cut-off is a prefix sum over the sorted order.
//   rebuild node-parallel: write the new array (bounds as DivideNode computes them :481-537)
//   relabel key-parallel: new node index of every key
// Equal-size ties in the careful phase are ordered by creation (array index) — DEVIATION-1 of the
// oracle.  The retained key of a node is its max-response key, first in reference order on ties:
// a 32-bit atomicMax over (response << 24 | ~original index).
// Keys and labels live in shared memory when the level has at most `keyCap` candidates, otherwise
// in the (L2-resident) key buffers.
// =============================================================================================
#define OT_NOSPLIT 0xffffffffu

__device__ __forceinline__ int key_x(uint32_t k) { return k & 0xfff; }
__device__ __forceinline__ int key_y(uint32_t k) { return (k >> 12) & 0xfff; }

// inclusive scan over the block (all threads call it); total = block sum
template <int OT_THREADS>
__device__ __forceinline__ uint32_t block_scan_incl(uint32_t v, uint32_t* wsum, uint32_t& total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += u; }
    if (lane == 31) wsum[wid] = v;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < OT_THREADS / 32; ++w) { const uint32_t sx = wsum[w]; if (w < wid) off += sx; tot += sx; }
    __syncthreads();
    total = tot;
    return v + off;
}

__device__ __forceinline__ int ot_quadrant(uint2 b, uint32_t key, int& mx, int& my)
{
    const int ulx = b.x & 0xffff, uly = b.x >> 16, urx = b.y & 0xffff, bry = b.y >> 16;
    mx = ulx + ((urx - ulx + 1) >> 1);          // UL.x + ceil((UR.x-UL.x)/2), ORBextractor.cc:483
    my = uly + ((bry - uly + 1) >> 1);
    return (key_x(key) >= mx ? 1 : 0) | (key_y(key) >= my ? 2 : 0);
}

template <int OT_THREADS>
__global__ void __launch_bounds__(OT_THREADS)
octree_kernel(Geom g, int nframes, const uint32_t* __restrict__ cand, const int2* __restrict__ cellSeg,
              uint32_t* __restrict__ keysAbase, uint32_t* __restrict__ keysBbase, uint32_t* __restrict__ levelKps,
              int* __restrict__ levelKpCount, int NC, int P2, int cellWords, int keyCap, int* __restrict__ status, int levelBase)
{
    extern __shared__ __align__(16) u8 smem[];
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31;
    // level-major block order: the long level-0 CTAs are dispatched first, the kernel's tail is made of short ones
    const int level = levelBase + blockIdx.x / nframes, frame = blockIdx.x % nframes;
    const LevelGeom& L = g.L[level];
    // ---- shared memory carve-up (the cell tables of the gather alias everything up to cntB) ----
    unsigned long long* sortbuf = reinterpret_cast<unsigned long long*>(smem);
    uint2* bndA = reinterpret_cast<uint2*>(sortbuf + P2);
    uint2* bndB = bndA + NC;
    uint32_t* cc = reinterpret_cast<uint32_t*>(bndB + NC);
    uint32_t* rel = cc + 4 * NC;
    uint32_t* tmp = rel + NC;
    uint32_t* cntA = tmp + NC;
    uint32_t* aliasEnd = cntA + NC;
    uint32_t* segStart = reinterpret_cast<uint32_t*>(smem);
    {
        const int have = (int)(aliasEnd - segStart);
        if (cellWords > have) aliasEnd = segStart + cellWords;
    }
    uint32_t* cntB = aliasEnd;
    uint32_t* wsum = cntB + NC;
    uint32_t* misc = wsum + 32;
    uint32_t* keysS = misc + 8;
    unsigned short* nidS = reinterpret_cast<unsigned short*>(keysS + keyCap);

    uint32_t* outKps = levelKps + (size_t)frame * g.kpTotal + L.kpOff;
    int* outCount = levelKpCount + frame * g.nlevels + level;
    if (L.nIni <= 0 || L.nCells <= 0) { if (tid == 0) *outCount = 0; return; }

    // ---- gather the per-cell segments into reference order (cell-row-major) --------------------
    uint32_t* segSrc = segStart + L.nCells + 1;
    uint32_t carry = 0;
    {
        const int2* seg = cellSeg + (size_t)frame * g.ncells + L.cellBeg;
        for (int c0 = 0; c0 < L.nCells; c0 += OT_THREADS) {
            const int c = c0 + tid;
            int2 sg = make_int2(0, 0);
            if (c < L.nCells) sg = seg[c];
            uint32_t tot;
            const uint32_t incl = block_scan_incl<OT_THREADS>((uint32_t)sg.y, wsum, tot);
            if (c < L.nCells) { segStart[c] = carry + incl - (uint32_t)sg.y; segSrc[c] = (uint32_t)sg.x; }
            carry += tot;
        }
        if (tid == 0) segStart[L.nCells] = carry;
        for (int r = tid; r < L.nIni; r += OT_THREADS) cntB[r] = 0;
    }
    const int n = (int)carry;
    if (n == 0) { if (tid == 0) *outCount = 0; return; }
    __syncthreads();
    const bool inShared = n <= keyCap;
    uint32_t* keys = inShared ? keysS : keysAbase + (size_t)frame * g.candTotal + L.candOff;
    unsigned short* nid = inShared ? nidS : reinterpret_cast<unsigned short*>(keysBbase + (size_t)frame * g.candTotal + L.candOff);
    {
        const uint32_t* csrc = cand + (size_t)frame * g.candTotal + L.candOff;
        for (int k0 = 0; k0 < n; k0 += OT_THREADS) {
            const int k = k0 + tid;
            uint32_t r = 0xffffffffu;
            if (k < n) {
                int lo = 0, hi = L.nCells;                  // last cell with segStart <= k
                while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (segStart[mid] <= (uint32_t)k) lo = mid; else hi = mid; }
                const uint32_t key = csrc[segSrc[lo] + ((uint32_t)k - segStart[lo])];
                keys[k] = key;
                r = (uint32_t)__float2int_rz(__fdiv_rn((float)key_x(key), L.hX));     // vpIniNodes[kp.pt.x / hX], :557
                nid[k] = (unsigned short)r;
            }
            const unsigned peers = __match_any_sync(FULL, r);
            if (r != 0xffffffffu && lane == __ffs(peers) - 1) atomicAdd(&cntB[r], (uint32_t)__popc(peers));
        }
    }
    __syncthreads();
    // ---- roots (reversed list: root nIni-1 first), empty roots are erased (:565-578) ------------
    if (tid == 0) {
        int top = 0;
        for (int r = L.nIni - 1; r >= 0; --r) {
            const uint32_t c = cntB[r];
            if (c == 0) continue;
            const int ulx = __float2int_rz(__fmul_rn(L.hX, (float)r)), urx = __float2int_rz(__fmul_rn(L.hX, (float)(r + 1)));
            bndA[top] = make_uint2((uint32_t)ulx, (uint32_t)urx | ((uint32_t)L.winH << 16));
            cntA[top] = c;
            rel[r] = (uint32_t)top;
            ++top;
        }
        misc[5] = (uint32_t)top;
    }
    __syncthreads();
    int len = (int)misc[5];
    for (int k = tid; k < n; k += OT_THREADS) nid[k] = (unsigned short)rel[nid[k]];
    __syncthreads();

    const int N = L.N;
    bool finish = false, careful = false, overflow = false;
    while (!finish) {
        const int prev = len;
        // ---- count: keys per quadrant of every node that may split ------------------------------
        for (int i = tid; i < 4 * len; i += OT_THREADS) cc[i] = 0;
        if (tid == 0) { misc[0] = 0; misc[2] = 0; misc[3] = 0; misc[4] = 0; }
        __syncthreads();
        for (int k0 = 0; k0 < n; k0 += OT_THREADS) {
            const int k = k0 + tid;
            uint32_t idx = 0xffffffffu;
            if (k < n) {
                const int ni = nid[k];
                if (cntA[ni] > 1) { int mx, my; idx = 4u * ni + ot_quadrant(bndA[ni], keys[k], mx, my); }
            }
            const unsigned peers = __match_any_sync(FULL, idx);
            if (idx != 0xffffffffu && lane == __ffs(peers) - 1) atomicAdd(&cc[idx], (uint32_t)__popc(peers));
        }
        __syncthreads();
        // ---- order ------------------------------------------------------------------------------
        uint32_t sTotal, cTotal;
        if (!careful) {
            // every node with > 1 key splits; children of node len-1 come first
            uint32_t run = 0;
            for (int i0 = 0; i0 < len; i0 += OT_THREADS) {
                const int i = i0 + tid;
                uint32_t v = 0;
                if (i < len) {
                    if (cntA[i] > 1) v = (uint32_t)((cc[4 * i] != 0) + (cc[4 * i + 1] != 0) + (cc[4 * i + 2] != 0) + (cc[4 * i + 3] != 0)) << 16;
                    else v = 1;
                }
                uint32_t tot;
                const uint32_t incl = block_scan_incl<OT_THREADS>(v, wsum, tot);
                if (i < len) tmp[i] = run + incl;
                run += tot;
            }
            sTotal = run & 0xffff; cTotal = run >> 16;
            for (int i = tid; i < len; i += OT_THREADS) {
                const uint32_t t = tmp[i];
                if (cntA[i] > 1) rel[i] = cTotal - (t >> 16);
                else { rel[i] = OT_NOSPLIT; tmp[i] = (t & 0xffff) - 1; }
            }
        } else {
            // candidates = nodes with > 1 key, sorted ascending by (nKeys, array index), processed from the back
            for (int i = tid; i < len; i += OT_THREADS) {
                rel[i] = OT_NOSPLIT;
                if (cntA[i] > 1) { const uint32_t p = atomicAdd(&misc[0], 1u); sortbuf[p] = ((unsigned long long)cntA[i] << 32) | (uint32_t)i; }
            }
            __syncthreads();
            const int nc = (int)misc[0];
            int cap2 = 1; while (cap2 < nc) cap2 <<= 1;
            for (int i = nc + tid; i < cap2; i += OT_THREADS) sortbuf[i] = ~0ull;
            __syncthreads();
            for (int kk = 2; kk <= cap2; kk <<= 1)
                for (int j = kk >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < cap2; i += OT_THREADS) {
                        const int l = i ^ j;
                        if (l > i) {
                            const unsigned long long x = sortbuf[i], y = sortbuf[l];
                            if ((x > y) == ((i & kk) == 0)) { sortbuf[i] = y; sortbuf[l] = x; }
                        }
                    }
                    __syncthreads();
                }
            // prefix sums in processing order: node p is split iff the node count is still < N before it (:736)
            uint32_t run = 0;
            for (int p0 = 0; p0 < nc; p0 += OT_THREADS) {
                const int p = p0 + tid;
                uint32_t v = 0, ni = 0;
                if (p < nc) {
                    ni = (uint32_t)sortbuf[nc - 1 - p];
                    const uint32_t nch = (uint32_t)((cc[4 * ni] != 0) + (cc[4 * ni + 1] != 0) + (cc[4 * ni + 2] != 0) + (cc[4 * ni + 3] != 0));
                    v = (nch - 1) | (nch << 16);
                }
                uint32_t tot;
                const uint32_t incl = block_scan_incl<OT_THREADS>(v, wsum, tot);
                if (p < nc) {
                    const uint32_t inc = run + incl, exc = inc - v;
                    if (len + (int)(exc & 0xffff) < N) {
                        rel[ni] = exc >> 16;
                        if (p == nc - 1 || len + (int)(inc & 0xffff) >= N) { misc[2] = inc & 0xffff; misc[3] = inc >> 16; }
                    }
                }
                run += tot;
            }
            __syncthreads();
            uint32_t srun = 0;
            for (int i0 = 0; i0 < len; i0 += OT_THREADS) {
                const int i = i0 + tid;
                const uint32_t v = (i < len && rel[i] == OT_NOSPLIT) ? 1u : 0u;
                uint32_t tot;
                const uint32_t incl = block_scan_incl<OT_THREADS>(v, wsum, tot);
                if (i < len) tmp[i] = srun + incl - 1;
                srun += tot;
            }
            sTotal = srun; cTotal = misc[3];
        }
        if ((int)(sTotal + cTotal) > NC) { overflow = true; break; }
        // ---- rebuild: survivors keep their order, children follow in split order ----------------
        {
            uint32_t nExp = 0;
            for (int i = tid; i < len; i += OT_THREADS) {
                const uint2 b = bndA[i];
                if (rel[i] == OT_NOSPLIT) {
                    const uint32_t j = tmp[i];
                    bndB[j] = b; cntB[j] = cntA[i]; cc[4 * i] = j;
                } else {
                    uint32_t w = sTotal + rel[i];
                    const int ulx = b.x & 0xffff, uly = b.x >> 16, urx = b.y & 0xffff, bry = b.y >> 16;
                    const int mx = ulx + ((urx - ulx + 1) >> 1), my = uly + ((bry - uly + 1) >> 1);
                    // creation order n1 (UL), n2 (UR), n3 (BL), n4 (BR), :618-661
                    const uint32_t c0 = cc[4 * i], c1 = cc[4 * i + 1], c2 = cc[4 * i + 2], c3 = cc[4 * i + 3];
                    if (c0) { bndB[w] = make_uint2(ulx | (uly << 16), mx | (my << 16));  cntB[w] = c0; cc[4 * i] = w;     nExp += c0 > 1; ++w; }
                    if (c1) { bndB[w] = make_uint2(mx | (uly << 16), urx | (my << 16));  cntB[w] = c1; cc[4 * i + 1] = w; nExp += c1 > 1; ++w; }
                    if (c2) { bndB[w] = make_uint2(ulx | (my << 16), mx | (bry << 16));  cntB[w] = c2; cc[4 * i + 2] = w; nExp += c2 > 1; ++w; }
                    if (c3) { bndB[w] = make_uint2(mx | (my << 16), urx | (bry << 16));  cntB[w] = c3; cc[4 * i + 3] = w; nExp += c3 > 1; ++w; }
                }
            }
            if (nExp) atomicAdd(&misc[4], nExp);
        }
        __syncthreads();
        // ---- relabel the keys ---------------------------------------------------------------------
        for (int k = tid; k < n; k += OT_THREADS) {
            const int ni = nid[k];
            uint32_t j;
            if (rel[ni] == OT_NOSPLIT) j = cc[4 * ni];
            else { int mx, my; j = cc[4 * ni + ot_quadrant(bndA[ni], keys[k], mx, my)]; }
            nid[k] = (unsigned short)j;
        }
        const int nToExpand = (int)misc[4];
        __syncthreads();
        { uint2* tb = bndA; bndA = bndB; bndB = tb; uint32_t* tc = cntA; cntA = cntB; cntB = tc; }
        len = (int)(sTotal + cTotal);
        if (len >= N || len == prev) finish = true;
        else if (!careful && len + nToExpand * 3 > N) careful = true;      // :667-673
    }
    if (overflow) { if (tid == 0) { atomicOr(status, 1); *outCount = 0; } return; }
    if (len > L.kpCap) { if (tid == 0) { atomicOr(status, 2); *outCount = 0; } return; }
    // ---- retain the best key of every node (:741-760); list order = reversed array ----------------
    for (int i = tid; i < len; i += OT_THREADS) cc[i] = 0;
    __syncthreads();
    for (int k = tid; k < n; k += OT_THREADS) atomicMax(&cc[nid[k]], (keys[k] & 0xff000000u) | (0xffffffu - (uint32_t)k));
    __syncthreads();
    for (int i = tid; i < len; i += OT_THREADS) outKps[len - 1 - i] = keys[0xffffffu - (cc[i] & 0xffffffu)];
    if (tid == 0) *outCount = len;
}

// =============================================================================================
// K4: orientation (IC_Angle) + rotated BRIEF + final keypoint record.  One warp per keypoint.
// =============================================================================================

__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    // cv::fastAtan2 scalar path; every op individually rounded
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

#define DESC_WARPS 8
__global__ void __launch_bounds__(DESC_WARPS * 32)
describe_kernel(Geom g, Level0 l0, const u8* __restrict__ rawBase, const u8* __restrict__ blurBase,
                int nframes, const uint32_t* __restrict__ levelKps, const int* __restrict__ levelKpCount,
                const int* __restrict__ kpMap, const signed char* __restrict__ pattern, const uint32_t* __restrict__ discMask,
                orbx_keypoint* __restrict__ outKps, u8* __restrict__ outDesc, int* __restrict__ outCount,
                int* __restrict__ outLevelCount, int firstSlot, int kpCapOut)
{
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int frame = blockIdx.y, slotk = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);
    if (slotk >= g.kpTotal) return;
    // the three loads below do not depend on each other: issue them together (the warp is short-lived, its
    // chain of dependent global loads is what it mostly waits for)
    const int level = kpMap[slotk];
    const uint32_t packed = levelKps[(size_t)frame * g.kpTotal + slotk];
    const int* counts = levelKpCount + frame * g.nlevels;
    // lane l holds the count of level l: output offset = keypoints of the lower levels
    const int myCount = lane < g.nlevels ? counts[lane] : 0;
    const LevelGeom& L = g.L[level];
    const int k = slotk - L.kpOff;
    int offset = lane < level ? myCount : 0, total = myCount;
    for (int d = 16; d > 0; d >>= 1) { offset += __shfl_xor_sync(FULL, offset, d); total += __shfl_xor_sync(FULL, total, d); }
    if (slotk == 0) {
        if (lane == 0) outCount[firstSlot + frame] = total;
        if (lane < g.nlevels) outLevelCount[(firstSlot + frame) * ORBX_MAX_LEVELS + lane] = myCount;
    }
    if (k >= __shfl_sync(FULL, myCount, level)) return;
    const int px = key_x(packed) + 16, py = key_y(packed) + 16;     // + minBorderX/Y (EDGE_THRESHOLD-3)
    const u8* raw;
    int rp;
    if (level == 0) { raw = l0_frame(l0, frame); rp = l0.pitch; }
    else { raw = rawBase + (long long)frame * g.pyrFrameBytes + L.off; rp = L.pitch; }
    // ---- IC_Angle: the 31x31 window as 8 words x 32 rows; lane = (row r of a group of 4, word w), 8 groups.
    // The window is re-aligned to the keypoint with a funnel shift, so byte j of lane w is column
    // u = 4w + j - 15 for every keypoint: the disc mask of (lane, group) is a table word and the column
    // weights are a per-lane constant.  m10 = sum u*I and the row sums come from two IDP.4A per word.
    int m10 = 0, m01 = 0;
    {
        const int r = lane >> 3, wq = lane & 7;
        const u8* win = raw + (long long)(py - 15) * rp + (px - 15);
        const int a = (int)(reinterpret_cast<uintptr_t>(win) & 3);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(win - a) + wq;
        const int rpw = rp >> 2, a8 = 8 * a;
        const int u0 = 4 * wq - 15;
        const uint32_t Uw = (uint32_t)(u0 & 0xff) | ((uint32_t)((u0 + 1) & 0xff) << 8) | ((uint32_t)((u0 + 2) & 0xff) << 16) | ((uint32_t)((u0 + 3) & 0xff) << 24);
        uint32_t lo[8], hi[8], M[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = min(4 * i + r, 30);                 // row 31 does not exist (its mask is 0)
            const uint32_t* q = wp + row * rpw;
            lo[i] = q[0]; hi[i] = q[1];
            M[i] = __ldg(discMask + i * 32 + lane);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t Iw = __funnelshift_r(lo[i], hi[i], a8) & M[i];
            const int cs = (int)__dp4a(Iw, 0x01010101u, 0u);
            asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(Iw), "r"(Uw));
            m01 += (4 * i + r - 15) * cs;
        }
    }
    for (int d = 16; d > 0; d >>= 1) { m10 += __shfl_xor_sync(FULL, m10, d); m01 += __shfl_xor_sync(FULL, m01, d); }
    const float angle = fast_atan2_deg((float)m01, (float)m10);
    // ---- rotated BRIEF on the blurred level: lane i -> descriptor byte i -----------------------
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    float sa, ca;
    orbx_sincosf(__fmul_rn(angle, factorPI), &sa, &ca);
    const u8* bl = blurBase + (long long)frame * g.pyrFrameBytes + L.off;
    const u8* bc = bl + (long long)py * L.pitch + px;
    // lane i evaluates the 8 point pairs of descriptor byte i: 32 bytes of the pattern table
    const int4 patA = __ldg(reinterpret_cast<const int4*>(pattern) + lane * 2), patB = __ldg(reinterpret_cast<const int4*>(pattern) + lane * 2 + 1);
    const int patw[8] = {patA.x, patA.y, patA.z, patA.w, patB.x, patB.y, patB.z, patB.w};
    int val = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int pw = patw[j];                                       // x0, y0, x1, y1 as signed bytes
        const float x0 = (float)(signed char)(pw & 0xff), y0 = (float)(signed char)((pw >> 8) & 0xff);
        const float x1 = (float)(signed char)((pw >> 16) & 0xff), y1 = (float)(signed char)(pw >> 24);
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sa), __fmul_rn(y0, ca)));
        const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sa)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
        const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
        const int t0 = bc[r0 * L.pitch + q0], t1 = bc[r1 * L.pitch + q1];
        val |= (t0 < t1) << j;
    }
    const size_t o = (size_t)(firstSlot + frame) * kpCapOut + offset + k;
    outDesc[o * 32 + lane] = (u8)val;
    if (lane == 0) {
        orbx_keypoint kp;
        kp.x = (float)px; kp.y = (float)py;
        if (level != 0) { kp.x = __fmul_rn(kp.x, L.scale); kp.y = __fmul_rn(kp.y, L.scale); }
        kp.size = L.kpSize;
        kp.angle = angle;
        kp.response = (float)(packed >> 24);
        kp.octave = level;
        kp.class_id = -1;
        outKps[o] = kp;
    }
}

// =============================================================================================
// host side: geometry for an image shape, and the per-chunk launch sequence
// =============================================================================================
static inline int cv_roundf(float v) { return (int)lrintf(v); }

static void linear_coeffs(int sn, int dn, std::vector<int2>& out)
{
    // cv::resize INTER_LINEAR coefficient rule
    const double scale = (double)sn / dn;
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= sn - 1) { s = sn - 1; f = 0.f; }
        const int s1 = std::min(s + 1, sn - 1);
        const int a0 = cv_roundf((1.f - f) * 2048.f), a1 = cv_roundf(f * 2048.f);
        out.push_back(make_int2(s | (s1 << 16), (a0 & 0xffff) | (a1 << 16)));
    }
}

// Geometry, tables and workspace of one image shape.  Everything is computed and validated in locals first; the
// handle forgets its current shape before the first buffer is touched and learns the new one only after every
// allocation and upload succeeded, so a failed call can never leave a shape whose buffers are stale.
int orbx_setup_shape(orbx_handle* h, int w, int hgt)
{
    if (h->curW == w && h->curH == hgt) return ORBX_OK;
    if (w > h->maxW || hgt > h->maxH) { orbx_set_error("image %dx%d exceeds handle maximum %dx%d", w, hgt, h->maxW, h->maxH); return ORBX_E_CAPACITY; }
    if (w > 4000 || hgt > 4000) { orbx_set_error("image larger than 4000 px per side is unsupported"); return ORBX_E_INVALID; }
    const int nl = h->params.nlevels;
    Geom g;
    memset(&g, 0, sizeof(g));
    g.nlevels = nl;
    std::vector<Cell> cells;
    std::vector<int2> coef;
    std::vector<int> tiles, kpMap;
    int blurTileBeg[ORBX_MAX_LEVELS + 1] = {0};
    long long off = 0;
    int candOff = 0, kpOff = 0;
    for (int l = 0; l < nl; ++l) {
        LevelGeom& L = g.L[l];
        L.w = cv_roundf((float)w * h->invScale[l]);
        L.h = cv_roundf((float)hgt * h->invScale[l]);
        if (L.w < 12 || L.h < 1) { orbx_set_error("level %d of a %dx%d image is %dx%d: too small for the pyramid kernels", l, w, hgt, L.w, L.h); return ORBX_E_INVALID; }
        L.pitch = (L.w + 15) & ~15;
        L.off = (int)off;
        off += (long long)L.pitch * L.h;
        off = (off + 255) & ~255LL;
        L.scale = h->scale[l];
        L.kpSize = (float)(int)(31 * h->scale[l]);
        L.N = h->perLevel[l];
        if (l > 0) {
            L.coefX = (int)coef.size(); linear_coeffs(g.L[l - 1].w, L.w, coef);
            L.coefY = (int)coef.size(); linear_coeffs(g.L[l - 1].h, L.h, coef);
        }
        {
            const int nstrips = std::max(1, (L.h + BL_TARGET - 1) / BL_TARGET);
            const int g7 = ((L.h + nstrips - 1) / nstrips + 6 + 6) / 7;        // 7-row groups covering rows + 6 halo rows
            L.blurTH = std::max(1, 7 * g7 - 6);
        }
        blurTileBeg[l] = (int)tiles.size();
        for (int ty = 0; ty * L.blurTH < L.h; ++ty)
            for (int tx = 0; tx * BL_TW < L.w; ++tx) tiles.push_back((l << 24) | (ty << 12) | tx);
        blurTileBeg[l + 1] = (int)tiles.size();
        // FAST cell grid (src/ORBextractor.cc:773-806)
        const int minBX = 16, minBY = 16, maxBX = L.w - 16, maxBY = L.h - 16;
        L.winW = maxBX - minBX; L.winH = maxBY - minBY;
        L.cellBeg = (int)cells.size();
        L.candOff = candOff;
        int cap = 0;
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
        if (L.winW > 0 && L.winH > 0 && nCols > 0 && nRows > 0) {
            const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows);
            L.wCell = wCell; L.hCell = hCell;
            for (int i = 0; i < nRows; ++i) {
                const float iniY = (float)(minBY + i * hCell);
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBY - 3) continue;
                if (maxY > maxBY) maxY = (float)maxBY;
                for (int j = 0; j < nCols; ++j) {
                    const float iniX = (float)(minBX + j * wCell);
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBX - 6) continue;
                    if (maxX > maxBX) maxX = (float)maxBX;
                    Cell c;
                    c.x0 = (unsigned short)iniX; c.y0 = (unsigned short)iniY;
                    c.cw = (unsigned short)((int)maxX - (int)iniX); c.ch = (unsigned short)((int)maxY - (int)iniY);
                    c.offx = (unsigned short)(j * wCell); c.offy = (unsigned short)(i * hCell);
                    c.level = (unsigned short)l; c.pad = 0;
                    if (c.cw > 66 || c.ch > 66) { orbx_set_error("FAST cell too large (%dx%d)", c.cw, c.ch); return ORBX_E_INVALID; }
                    cells.push_back(c);
                    g.cwMax = std::max<int>(g.cwMax, c.cw); g.chMax = std::max<int>(g.chMax, c.ch);
                    const int dw = c.cw - 6, dh = c.ch - 6;
                    if (dw > 0 && dh > 0) cap = std::max(cap, ((dw + 1) / 2) * ((dh + 1) / 2));
                }
            }
            // quadtree roots (src/ORBextractor.cc:543-546)
            L.nIni = (int)roundf((float)(maxBX - minBX) / (maxBY - minBY));
            L.hX = L.nIni > 0 ? (float)(maxBX - minBX) / L.nIni : 0.f;
        }
        L.nCells = (int)cells.size() - L.cellBeg;
        L.cellCap = cap;                                        // candidate segment per cell (the largest cell's bound)
        L.candCap = (L.nCells * cap + 31) & ~31;
        candOff += L.candCap;
        L.kpOff = kpOff;
        L.kpCap = std::max(L.N + 3, 4 * std::max(L.nIni, 1)) + 1;
        for (int k = 0; k < L.kpCap; ++k) kpMap.push_back(l);
        kpOff += L.kpCap;
    }
    g.ncells = (int)cells.size();
    g.candTotal = candOff;
    g.kpTotal = kpOff;
    g.pyrFrameBytes = off;
    if (g.cwMax < 7) g.cwMax = 7;
    if (g.chMax < 7) g.chMax = 7;
    // ---- from here on the handle is touched: forget the current shape first -----------------------------------
    cudaStream_t st = h->stream;
    ORBX_CUDA(cudaStreamSynchronize(st));
    orbx_drop_graph(h);
    h->curW = h->curH = 0;
    h->lastChunkFrames = 0;
    // the per-frame keypoint capacity was sized at creation for the maximum shape; a smaller image with a wider
    // aspect ratio has more quadtree roots per level (nIni) and may need more: the result store grows (stored slots
    // are discarded; orbx_keypoint_capacity / orbx_reserve_shape report the new value)
    if (g.kpTotal > h->kpCap) { const int rc = orbx_grow_result_store(h, (g.kpTotal + 7) & ~7); if (rc) return rc; }
    auto re = [&](void** p, size_t bytes) -> cudaError_t { if (*p) cudaFree(*p); *p = nullptr; return cudaMalloc(p, std::max<size_t>(bytes, 256)); };
    const size_t C = (size_t)h->chunkFrames;
    ORBX_CUDA(re((void**)&h->dCells, cells.size() * sizeof(Cell)));
    ORBX_CUDA(re((void**)&h->dCoef, coef.size() * sizeof(int2)));
    ORBX_CUDA(re((void**)&h->dBlurTiles, tiles.size() * sizeof(int)));
    ORBX_CUDA(re((void**)&h->dKpMap, kpMap.size() * sizeof(int)));
    h->nBlurTiles = (int)tiles.size();
    memcpy(h->blurTileBeg, blurTileBeg, sizeof(blurTileBeg));
    h->l0Pitch = (w + 15) & ~15;
    h->l0FrameBytes = (((long long)h->l0Pitch * hgt) + 255) & ~255LL;
    ORBX_CUDA(re((void**)&h->dLevel0, ORBX_L0_BUFS * C * h->l0FrameBytes));
    ORBX_CUDA(re((void**)&h->dStageLin, ORBX_L0_BUFS * C * h->l0FrameBytes));
    for (int b = 0; b < ORBX_L0_BUFS; ++b) h->evFreeValid[b] = false;
    ORBX_CUDA(re((void**)&h->dRaw, C * g.pyrFrameBytes));
    ORBX_CUDA(re((void**)&h->dBlur, C * g.pyrFrameBytes));
    ORBX_CUDA(re((void**)&h->dCand, C * g.candTotal * sizeof(uint32_t)));
    ORBX_CUDA(re((void**)&h->dKeysA, C * g.candTotal * sizeof(uint32_t)));
    ORBX_CUDA(re((void**)&h->dKeysB, C * g.candTotal * sizeof(uint32_t)));
    ORBX_CUDA(re((void**)&h->dCellSeg, C * g.ncells * sizeof(int2)));
    ORBX_CUDA(re((void**)&h->dLevelKps, C * g.kpTotal * sizeof(uint32_t)));
    ORBX_CUDA(re((void**)&h->dLevelKpCount, C * nl * sizeof(int)));
    ORBX_CUDA(cudaMemcpyAsync(h->dCells, cells.data(), cells.size() * sizeof(Cell), cudaMemcpyHostToDevice, st));
    if (!coef.empty()) ORBX_CUDA(cudaMemcpyAsync(h->dCoef, coef.data(), coef.size() * sizeof(int2), cudaMemcpyHostToDevice, st));
    ORBX_CUDA(cudaMemcpyAsync(h->dBlurTiles, tiles.data(), tiles.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    ORBX_CUDA(cudaMemcpyAsync(h->dKpMap, kpMap.data(), kpMap.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    h->geom = g;
    h->curW = w; h->curH = hgt;
    return ORBX_OK;
}

static signed char* g_dPattern[16] = {nullptr};
static uint32_t* g_dDiscMask[16] = {nullptr};

// device copies of the BRIEF pattern and of the IC_Angle disc masks (per device, shared by all handles:
// both depend only on PATCH_SIZE = 31)
static int get_pattern(int device, const int* umax, const signed char** out, const uint32_t** maskOut)
{
    if (device < 0 || device >= 16) { orbx_set_error("device index %d unsupported", device); return ORBX_E_INVALID; }
    if (!g_dPattern[device]) {
        static const signed char pat[ORBX_PATTERN_LEN] = {ORBX_PATTERN_VALUES};
        ORBX_CUDA(cudaMalloc((void**)&g_dPattern[device], ORBX_PATTERN_LEN));
        ORBX_CUDA(cudaMemcpy(g_dPattern[device], pat, ORBX_PATTERN_LEN, cudaMemcpyHostToDevice));
    }
    if (!g_dDiscMask[device]) {
        // word (group i, lane): rows v = 4i + (lane>>3) - 15, columns u = 4*(lane&7) + j - 15;
        // a pixel is inside the disc iff |v| <= umax[|u|] (IC_Angle, src/ORBextractor.cc:86-100)
        uint32_t tab[256];
        for (int i = 0; i < 8; ++i)
            for (int lane = 0; lane < 32; ++lane) {
                const int v = 4 * i + (lane >> 3) - 15;
                uint32_t m = 0;
                for (int j = 0; j < 4; ++j) {
                    const int u = 4 * (lane & 7) + j - 15;
                    if (abs(u) <= 15 && abs(v) <= 15 && abs(v) <= umax[abs(u)]) m |= 0xffu << (8 * j);
                }
                tab[i * 32 + lane] = m;
            }
        ORBX_CUDA(cudaMalloc((void**)&g_dDiscMask[device], sizeof(tab)));
        ORBX_CUDA(cudaMemcpy(g_dDiscMask[device], tab, sizeof(tab), cudaMemcpyHostToDevice));
    }
    *out = g_dPattern[device];
    *maskOut = g_dDiscMask[device];
    return ORBX_OK;
}

// Launches of one chunk.  forked = false: one stream, kernels in stage order over all levels (throughput form: every
// kernel fills the GPU).  forked = true (the single-frame latency path, orbx_extract): a frame leaves most SMs idle, so
// the dependency graph is what matters: level l's blur and its FAST -> quadtree chain start on side streams as soon as
// resize has produced level l (level 0 at once), run next to the rest of the resize chain, and join before describe.
// Under stream capture this becomes the forked CUDA graph that orbx_extract replays.
static int run_chunk_impl(orbx_handle* h, Level0 l0, int nframes, int firstSlot, bool forked)
{
    const Geom& g = h->geom;
    cudaStream_t st = h->stream;
    const int nl = g.nlevels;
    const signed char* dPat = nullptr;
    const uint32_t* dDisc = nullptr;
    int rc = get_pattern(h->device, h->umax, &dPat, &dDisc);
    if (rc) return rc;
    if (forked) {
        while ((int)h->sideStreams.size() < 2 * nl) {
            cudaStream_t s; ORBX_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)); h->sideStreams.push_back(s);
            cudaEvent_t e; ORBX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->sideEvents.push_back(e);
        }
        while ((int)h->levelEvents.size() < nl) { cudaEvent_t e; ORBX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->levelEvents.push_back(e); }
    }
    // FAST launch geometry (same for every level)
    const int tilePitch = (g.cwMax + 9 + 7) & ~7;          // rows 8-byte aligned; the packed pre-test reads whole words up to byte cw+8
    const int scorePitch = (g.cwMax - 4 + 3) & ~3;
    const int maskBytes = 8 * (g.chMax - 6);                // one 64-bit pre-test mask per detectable row
    int warpSmem = ((tilePitch * g.chMax + 15) & ~15) + ((((scorePitch * (g.chMax - 4) + 15) & ~15) + maskBytes + 15) & ~15) +
                   2 * (g.cwMax - 6) * (g.chMax - 6);
    warpSmem = (warpSmem + 15) & ~15;
    const int fastSmem = warpSmem * FAST_WARPS;
    if (fastSmem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, fastSmem));
    auto launch_fast = [&](cudaStream_t s, int cellBeg, int cellEnd) {
        fast_cells_kernel<<<dim3((cellEnd - cellBeg + FAST_WARPS - 1) / FAST_WARPS, nframes), FAST_WARPS * 32, fastSmem, s>>>(
            g, h->dCells, l0, h->dRaw, nframes, h->params.min_th_fast, h->params.ini_th_fast, h->dCand, h->dCellSeg,
            tilePitch, scorePitch, warpSmem, cellBeg, cellEnd);
    };
    int NC = 8, cellsMax = 1, candMax = 1;
    for (int l = 0; l < nl; ++l) { NC = std::max(NC, g.L[l].kpCap); cellsMax = std::max(cellsMax, g.L[l].nCells); candMax = std::max(candMax, g.L[l].candCap); }
    if (NC > 16383 || candMax >= (1 << 24)) { orbx_set_error("quadtree: %d nodes / %d candidates per level exceed the kernel's index width", NC, candMax); return ORBX_E_CAPACITY; }
    int P2 = 32; while (P2 < NC) P2 <<= 1;
    const int cellWords = 2 * cellsMax + 2;
    const int aliasWords = std::max(2 * P2 + 11 * NC, cellWords);
    const int fixedBytes = 4 * (aliasWords + NC + 32 + 8);
    // keys + labels of a level stay in shared memory up to keyCap candidates (6 bytes each); beyond that
    // the level runs out of the global key buffers
    // CTA width and shared-memory key capacity follow the per-level quota: more features per level means
    // more candidates per level (640x480/1000: ~2.5 k at level 0; 1920x1080/5000: ~18 k)
    int otThreads = NC <= 256 ? 128 : (NC <= 512 ? 256 : 512);
    if (forked) { static const int ot = getenv("ORBX_LAT_OT") ? atoi(getenv("ORBX_LAT_OT")) : 0; if (ot == 128 || ot == 256 || ot == 512) otThreads = ot; }
    int keyCap = std::min((candMax + 63) & ~63, otThreads == 128 ? 4096 : (otThreads == 256 ? 8192 : 24576));
    if (fixedBytes + 6 * keyCap > 200 * 1024) keyCap = std::max(0, (200 * 1024 - fixedBytes) / 6) & ~63;
    if (const char* e = getenv("ORBX_OCTREE_KEYCAP")) keyCap = std::max(0, std::min(keyCap, atoi(e)));   // tests: force the global-memory path
    const int otSmem = fixedBytes + 6 * keyCap;
    auto launch_octree = [&](cudaStream_t s, int levelBeg, int nLevels) -> int {
        auto launch = [&](auto kern) -> int {
            if (otSmem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, otSmem));
            kern<<<nframes * nLevels, otThreads, otSmem, s>>>(g, nframes, h->dCand, h->dCellSeg, h->dKeysA, h->dKeysB, h->dLevelKps,
                                                              h->dLevelKpCount, NC, P2, cellWords, keyCap, h->dStatus, levelBeg);
            return ORBX_OK;
        };
        return otThreads == 128 ? launch(octree_kernel<128>) : (otThreads == 256 ? launch(octree_kernel<256>) : launch(octree_kernel<512>));
    };
    auto launch_blur = [&](cudaStream_t s, int tileBeg, int tileEnd) {
        blur7_kernel<<<dim3((tileEnd - tileBeg + BL_WARPS - 1) / BL_WARPS, nframes), BL_WARPS * 32, 0, s>>>(g, l0, h->dRaw, h->dBlur, h->dBlurTiles + tileBeg, tileEnd - tileBeg);
    };
    // level l's consumers on their side streams (forked form)
    auto fork_level = [&](int l) -> int {
        const LevelGeom& L = g.L[l];
        cudaStream_t sA = h->sideStreams[2 * l], sB = h->sideStreams[2 * l + 1];
        ORBX_CUDA(cudaEventRecord(h->levelEvents[l], st));
        ORBX_CUDA(cudaStreamWaitEvent(sA, h->levelEvents[l], 0));
        ORBX_CUDA(cudaStreamWaitEvent(sB, h->levelEvents[l], 0));
        if (L.nCells > 0) launch_fast(sA, L.cellBeg, L.cellBeg + L.nCells);
        int r = launch_octree(sA, l, 1);
        if (r) return r;
        ORBX_CUDA(cudaEventRecord(h->sideEvents[2 * l], sA));
        // blur: per level (8), or levels grouped into fewer launches that start when their last level exists
        static const int blurGroups = getenv("ORBX_LAT_BLUR") ? atoi(getenv("ORBX_LAT_BLUR")) : 2;      // 8 / 4 / 2 / 1 groups measured 124 / 121 / 121 / 120 us
        const int per = std::max(1, (nl + std::max(1, blurGroups) - 1) / std::max(1, blurGroups));
        if ((l + 1) % per == 0 || l == nl - 1) {
            const int first = (l / per) * per;
            launch_blur(sB, h->blurTileBeg[first], h->blurTileBeg[l + 1]);
        }
        ORBX_CUDA(cudaEventRecord(h->sideEvents[2 * l + 1], sB));
        return ORBX_OK;
    };
    ORBX_NVTX(forked ? "orbx:frame(forked)" : "orbx:chunk");
    orbx_prof_mark(h, -1);
    if (forked) {
        rc = fork_level(0);
        if (rc) return rc;
    }
    // K1a: levels 1..n-1, each from the previous one
    nvtxRangePushA("orbx:resize");
    for (int l = 1; l < nl; ++l) {
        const LevelGeom& S = g.L[l - 1];
        const LevelGeom& D = g.L[l];
        const u8* src; long long sfs; int sp;
        if (l == 1) { src = l0.base; sfs = l0.frameStride; sp = l0.pitch; }
        else { src = h->dRaw + S.off; sfs = g.pyrFrameBytes; sp = S.pitch; }
        if (h->params.scale_factor <= 2.0f) {
            static const int rsTargetEnv = getenv("ORBX_RS_TARGET") ? std::max(4, atoi(getenv("ORBX_RS_TARGET"))) : RS_TARGET;
            static const int rsTargetLat = getenv("ORBX_LAT_RS") ? std::max(2, atoi(getenv("ORBX_LAT_RS"))) : 2;     // latency form: 2-row strips (120 us vs 144 us at 12 rows)
            const int rsTarget = forked ? rsTargetLat : rsTargetEnv;
            const int nstrips = std::max(1, (D.h + rsTarget - 1) / rsTarget), th = (D.h + nstrips - 1) / nstrips;
            dim3 grid((D.w + 127) / 128, (nstrips + RS_WARPS - 1) / RS_WARPS, nframes);
            // source rows a strip of th output rows can touch: th * (S.h / D.h) + 2, with slack for the table's rounding
            const int nsrcMax = ((int)((long long)th * S.h / D.h) + 4 + 3) & ~3;      // a multiple of 4: phase 1 stores whole groups
            const size_t rsSmem = (size_t)RS_WARPS * nsrcMax * 32 * sizeof(uint2);
            if (rsSmem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(pyr_resize4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsSmem));
            pyr_resize4_kernel<<<grid, RS_WARPS * 32, rsSmem, st>>>(l0, l == 1 ? 1 : 0, src, sfs, sp, S.w, S.h, h->dRaw + D.off, g.pyrFrameBytes,
                                                                    D.pitch, D.w, D.h, th, nsrcMax, h->dCoef + D.coefX, h->dCoef + D.coefY);
        } else {
            dim3 grid((D.w + 127) / 128, (D.h + 7) / 8, nframes), block(32, 8);
            pyr_resize_kernel<<<grid, block, 0, st>>>(l0, l == 1 ? 1 : 0, src, sfs, sp, h->dRaw + D.off, g.pyrFrameBytes, D.pitch, D.w, D.h,
                                                      h->dCoef + D.coefX, h->dCoef + D.coefY);
        }
        if (forked) { rc = fork_level(l); if (rc) return rc; }
    }
    nvtxRangePop();
    ORBX_COUNT(h, ORBX_STAGE_RESIZE, nl - 1);
    orbx_prof_mark(h, ORBX_STAGE_RESIZE);
    if (forked) {
        for (int i = 0; i < 2 * nl; ++i) ORBX_CUDA(cudaStreamWaitEvent(st, h->sideEvents[i], 0));
        ORBX_COUNT(h, ORBX_STAGE_BLUR, nl);
        ORBX_COUNT(h, ORBX_STAGE_FAST, nl);
        ORBX_COUNT(h, ORBX_STAGE_OCTREE, nl);
    } else {
        // K1b: blur of all levels
        { ORBX_NVTX("orbx:blur"); launch_blur(st, 0, h->nBlurTiles); }
        ORBX_COUNT(h, ORBX_STAGE_BLUR, 1);
        orbx_prof_mark(h, ORBX_STAGE_BLUR);
        // K2: FAST
        { ORBX_NVTX("orbx:fast"); launch_fast(st, 0, g.ncells); }
        ORBX_COUNT(h, ORBX_STAGE_FAST, 1);
        orbx_prof_mark(h, ORBX_STAGE_FAST);
        // K3: quadtree
        { ORBX_NVTX("orbx:quadtree"); rc = launch_octree(st, 0, nl); }
        if (rc) return rc;
        ORBX_COUNT(h, ORBX_STAGE_OCTREE, 1);
        orbx_prof_mark(h, ORBX_STAGE_OCTREE);
    }
    // K4: orientation + descriptors + output records
    {
        ORBX_NVTX("orbx:describe");
        describe_kernel<<<dim3((g.kpTotal + DESC_WARPS - 1) / DESC_WARPS, nframes), DESC_WARPS * 32, 0, st>>>(
            g, l0, h->dRaw, h->dBlur, nframes, h->dLevelKps, h->dLevelKpCount, h->dKpMap, dPat, dDisc, h->dKps, h->dDesc, h->dCount,
            h->dSlotLevelCount, firstSlot, h->kpCap);
        ORBX_COUNT(h, ORBX_STAGE_DESCRIBE, 1);
        orbx_prof_mark(h, ORBX_STAGE_DESCRIBE);
    }
    ORBX_CUDA(cudaGetLastError());
    h->lastL0 = l0;
    h->lastChunkFrames = nframes;
    return ORBX_OK;
}

int orbx_run_chunk(orbx_handle* h, Level0 l0, int nframes, int firstSlot) { return run_chunk_impl(h, l0, nframes, firstSlot, false); }
int orbx_run_frame_forked(orbx_handle* h, Level0 l0, int firstSlot) { return run_chunk_impl(h, l0, 1, firstSlot, true); }
