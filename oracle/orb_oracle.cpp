// orb_oracle.cpp — CPU oracle for the ORB front end.  TEST INFRASTRUCTURE ONLY (see orb_oracle.h).
//
// Restates, function by function, the reference's CPU path (paths relative to the reference repo
// carry4985/ORB_SLAM2_MAP) with OpenCV's primitives replaced by models pinned to cv2 4.13.0:
//   resize_linear_u8      <- cv::resize INTER_LINEAR 8UC1        (call site src/ORBextractor.cc:1120)
//   gaussian_blur7_u8     <- cv::GaussianBlur 7x7 s=2 REFLECT_101 (src/ORBextractor.cc:1086)
//   fast9_16              <- cv::FAST(img,kps,th,true)            (src/ORBextractor.cc:809,814)
//   fast_atan2            <- cv::fastAtan2                        (src/ORBextractor.cc:103)
// Parity status: primitives pinned against cv2 4.13.0 golden vectors (tests/golden/*.npz, made by
// tests/golden/make_golden.py); the composed extractor is cross-checked against the UNMODIFIED
// reference src/ORBextractor.cc compiled over a cv shim (oracle/_ref, oracle/Makefile).  The
// reference itself ships no tests or golden vectors.
// Determinisations: (1) DistributeOctTree sorts equal-size nodes by creation order
// instead of heap address; (2) cos/sin = include/orbx_sincos.h (libm switch available);
// (3) no FMA contraction (build with -ffp-contract=off); (4) OpenCV arithmetic = 4.13.0.
// Build: g++ -O2 -ffp-contract=off -std=c++17 -shared -fPIC (oracle/Makefile).

#include "orb_oracle.h"
#include "../include/orbx_pattern.h"
#include "../include/orbx_sincos.h"

#include <algorithm>
#include <climits>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <list>
#include <thread>
#include <atomic>
#include <vector>

namespace {

typedef uint8_t u8;

inline int cv_round(double v) { return (int)lrint(v); }
inline int cv_roundf(float v) { return (int)lrintf(v); }

// ------------------------------------------------------------------------------------------
// cv::resize, INTER_LINEAR, 8UC1
// ------------------------------------------------------------------------------------------
struct LinCoef { int s0, s1; short a0, a1; };

void linear_coeffs(int sn, int dn, std::vector<LinCoef>& out)
{
    out.resize(dn);
    const double scale = (double)sn / dn;
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= sn - 1) { s = sn - 1; f = 0.f; }
        LinCoef c;
        c.s0 = s;
        c.s1 = std::min(s + 1, sn - 1);
        c.a0 = (short)cv_roundf((1.f - f) * 2048.f);
        c.a1 = (short)cv_roundf(f * 2048.f);
        out[d] = c;
    }
}

void resize_linear_u8(const u8* src, int sw, int sh, size_t sstride, u8* dst, int dw, int dh,
                      size_t dstride)
{
    std::vector<LinCoef> cx, cy;
    linear_coeffs(sw, dw, cx);
    linear_coeffs(sh, dh, cy);
    std::vector<int> r0(dw), r1(dw);
    int have0 = -1, have1 = -1;
    for (int y = 0; y < dh; ++y) {
        const LinCoef& c = cy[y];
        // horizontal pass of the two source rows (cached between output rows)
        if (have0 != c.s0) {
            if (have1 == c.s0) { r0.swap(r1); std::swap(have0, have1); }
            else {
                const u8* s = src + (size_t)c.s0 * sstride;
                for (int x = 0; x < dw; ++x) r0[x] = s[cx[x].s0] * cx[x].a0 + s[cx[x].s1] * cx[x].a1;
                have0 = c.s0;
            }
        }
        if (have1 != c.s1) {
            const u8* s = src + (size_t)c.s1 * sstride;
            for (int x = 0; x < dw; ++x) r1[x] = s[cx[x].s0] * cx[x].a0 + s[cx[x].s1] * cx[x].a1;
            have1 = c.s1;
        }
        const int* S0 = r0.data();
        const int* S1 = (c.s1 == c.s0) ? r0.data() : r1.data();
        u8* d = dst + (size_t)y * dstride;
        const int b0 = c.a0, b1 = c.a1;
        for (int x = 0; x < dw; ++x) {
            int v = (((b0 * (S0[x] >> 4)) >> 16) + ((b1 * (S1[x] >> 4)) >> 16) + 2) >> 2;
            d[x] = (u8)v;
        }
    }
}

inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * (n - 1) - i;
    }
    return i;
}

// ------------------------------------------------------------------------------------------
// cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101), 8UC1
// ------------------------------------------------------------------------------------------
void gaussian_blur7_u8(const u8* src, int w, int h, size_t sstride, u8* dst, size_t dstride)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<int> hbuf((size_t)w * h);
    std::vector<int> xi(w + 6);
    for (int x = -3; x < w + 3; ++x) xi[x + 3] = reflect101(x, w);
    for (int y = 0; y < h; ++y) {
        const u8* s = src + (size_t)y * sstride;
        int* hb = hbuf.data() + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int i = 0; i < 7; ++i) acc += K[i] * s[xi[x + i]];
            hb[x] = acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        const int* rows[7];
        for (int j = 0; j < 7; ++j) rows[j] = hbuf.data() + (size_t)reflect101(y + j - 3, h) * w;
        u8* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int j = 0; j < 7; ++j) acc += K[j] * rows[j][x];
            d[x] = (u8)((acc + 32768) >> 16);
        }
    }
}

// ------------------------------------------------------------------------------------------
// cv::FAST 9/16 with optional 3x3 non-max suppression
// ------------------------------------------------------------------------------------------
struct Key { float x, y, response; };

const int RING_DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
const int RING_DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// S = largest t such that 9 contiguous ring pixels are all darker than v-t+1 .. or brighter;
// i.e. max over arcs of max(min d, -max d), d_k = v - ring_k.
inline int fast_strength(const u8* p, const int off[16])
{
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = INT_MIN;
    for (int k = 0; k < 16; ++k) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; ++j) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

inline bool fast_quick_reject(const u8* p, const int off[16], int th)
{
    // a 9-arc always contains one pixel of every opposite pair; if for some pair neither end is
    // beyond the threshold on the same side, no arc exists.
    const int v = p[0];
    unsigned bright_all = 3, m;   // bit0: darker-ring possible, bit1: brighter-ring possible
    for (int k = 0; k < 8 && bright_all; ++k) {
        const int a = v - p[off[k]], b = v - p[off[k + 8]];
        m = 0;
        if (a > th || b > th) m |= 1;
        if (a < -th || b < -th) m |= 2;
        bright_all &= m;
    }
    return bright_all == 0;
}

void fast9_16(const u8* img, int w, int h, size_t stride, int th, bool nms, std::vector<Key>& out)
{
    out.clear();
    if (w < 7 || h < 7) return;
    int off[16];
    for (int k = 0; k < 16; ++k) off[k] = RING_DY[k] * (int)stride + RING_DX[k];
    // score rows: three rolling rows as OpenCV keeps them; here a full score image is simpler and
    // gives the same result (non-corners and borders hold 0).
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y) {
        const u8* row = img + (size_t)y * stride;
        int* sr = score.data() + (size_t)y * w;
        for (int x = 3; x < w - 3; ++x) {
            if (fast_quick_reject(row + x, off, th)) continue;
            const int S = fast_strength(row + x, off);
            if (S > th) sr[x] = nms ? (S - 1) : -1;   // -1 marks "corner" when scores are not needed
        }
    }
    for (int y = 3; y < h - 3; ++y) {
        const int* sr = score.data() + (size_t)y * w;
        for (int x = 3; x < w - 3; ++x) {
            const int s = sr[x];
            if (s == 0) continue;
            if (!nms) { out.push_back(Key{(float)x, (float)y, 0.f}); continue; }
            const int* up = sr - w;
            const int* dn = sr + w;
            if (s > sr[x - 1] && s > sr[x + 1] && s > up[x - 1] && s > up[x] && s > up[x + 1] &&
                s > dn[x - 1] && s > dn[x] && s > dn[x + 1])
                out.push_back(Key{(float)x, (float)y, (float)s});
        }
    }
}

// ------------------------------------------------------------------------------------------
// cv::fastAtan2 (degrees), scalar float path; no FMA contraction.
// ------------------------------------------------------------------------------------------
float fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ------------------------------------------------------------------------------------------
// ORBextractor (reference src/ORBextractor.cc)
// ------------------------------------------------------------------------------------------
const int PATCH_SIZE = 31, HALF_PATCH = 15, EDGE_TH = 19;   // src/ORBextractor.cc:72-74

struct Image {
    int w = 0, h = 0;
    std::vector<u8> px;
    u8* row(int y) { return px.data() + (size_t)y * w; }
    const u8* row(int y) const { return px.data() + (size_t)y * w; }
    void alloc(int W, int H) { w = W; h = H; px.assign((size_t)W * H, 0); }
};

// quadtree node of DistributeOctTree (src/ORBextractor.cc:481-537, include/ORBextractor.h:35-47)
struct QNode {
    int ulx, uly, urx, bry;        // UL.x, UL.y, UR.x, BR.y (BL/BR follow)
    std::vector<int> keys;         // indices into the candidate array, order preserved
    bool no_more = false;
    long stamp = 0;                // creation order: replaces the heap address in the sort key
    std::list<QNode>::iterator self;
};

void divide_node(const QNode& p, const std::vector<Key>& cand, QNode ch[4])
{
    const int halfX = (int)std::ceil(static_cast<float>(p.urx - p.ulx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(p.bry - p.uly) / 2);
    const int mx = p.ulx + halfX, my = p.uly + halfY;
    ch[0].ulx = p.ulx; ch[0].uly = p.uly; ch[0].urx = mx;    ch[0].bry = my;      // n1: UL quadrant
    ch[1].ulx = mx;    ch[1].uly = p.uly; ch[1].urx = p.urx; ch[1].bry = my;      // n2: UR
    ch[2].ulx = p.ulx; ch[2].uly = my;    ch[2].urx = mx;    ch[2].bry = p.bry;   // n3: BL
    ch[3].ulx = mx;    ch[3].uly = my;    ch[3].urx = p.urx; ch[3].bry = p.bry;   // n4: BR
    for (int c = 0; c < 4; ++c) { ch[c].keys.clear(); ch[c].no_more = false; }
    for (int idx : p.keys) {
        const Key& k = cand[idx];
        if (k.x < mx) { if (k.y < my) ch[0].keys.push_back(idx); else ch[2].keys.push_back(idx); }
        else if (k.y < my) ch[1].keys.push_back(idx);
        else ch[3].keys.push_back(idx);
    }
    for (int c = 0; c < 4; ++c) if (ch[c].keys.size() == 1) ch[c].no_more = true;
}

// src/ORBextractor.cc:539-763.  Returns indices of the retained candidates in list order.
std::vector<int> distribute_octtree(const std::vector<Key>& cand, int minX, int maxX, int minY,
                                    int maxY, int N)
{
    std::vector<int> result;
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    if (nIni <= 0) return result;     // reference divides by zero here; out of domain
    const float hX = static_cast<float>(maxX - minX) / nIni;

    std::list<QNode> nodes;
    long stamp = 0;
    std::vector<QNode*> roots(nIni);
    for (int i = 0; i < nIni; ++i) {
        QNode n;
        n.ulx = (int)(hX * static_cast<float>(i));
        n.urx = (int)(hX * static_cast<float>(i + 1));
        n.uly = 0;
        n.bry = maxY - minY;
        n.stamp = stamp++;
        nodes.push_back(n);
        roots[i] = &nodes.back();
    }
    for (size_t i = 0; i < cand.size(); ++i) {
        int r = (int)(cand[i].x / hX);
        roots[r]->keys.push_back((int)i);
    }
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->keys.size() == 1) { it->no_more = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }

    typedef std::pair<int, long> SizeStamp;            // (nKeys, creation stamp)
    struct Pending { SizeStamp key; QNode* node; };
    auto by_key = [](const Pending& a, const Pending& b) { return a.key < b.key; };
    std::vector<Pending> expandable;

    auto push_children = [&](QNode ch[4], int& nToExpand) {
        for (int c = 0; c < 4; ++c) {
            if (ch[c].keys.empty()) continue;
            ch[c].stamp = stamp++;
            nodes.push_front(ch[c]);
            nodes.front().self = nodes.begin();
            if (ch[c].keys.size() > 1) {
                ++nToExpand;
                expandable.push_back(Pending{SizeStamp((int)ch[c].keys.size(), nodes.front().stamp),
                                             &nodes.front()});
            }
        }
    };

    bool finish = false;
    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        expandable.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->no_more) { ++it; continue; }
            QNode ch[4];
            divide_node(*it, cand, ch);
            push_children(ch, nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<Pending> prev = expandable;
                expandable.clear();
                std::sort(prev.begin(), prev.end(), by_key);
                for (int j = (int)prev.size() - 1; j >= 0; --j) {
                    QNode ch[4];
                    divide_node(*prev[j].node, cand, ch);
                    int dummy = 0;
                    push_children(ch, dummy);
                    nodes.erase(prev[j].node->self);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }
    result.reserve(nodes.size());
    for (const QNode& n : nodes) {
        int best = n.keys[0];
        float maxResponse = cand[best].response;
        for (size_t k = 1; k < n.keys.size(); ++k)
            if (cand[n.keys[k]].response > maxResponse) { best = n.keys[k]; maxResponse = cand[best].response; }
        result.push_back(best);
    }
    return result;
}

struct Extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;     // the reference member is a double initialised from a float
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> perLevel, umax;
    signed char pattern[ORBX_PATTERN_LEN];
    // state of the last call
    std::vector<Image> raw, blur;
    std::vector<std::vector<Key>> candidates;
    std::vector<std::vector<orbx_keypoint>> levelKeys;

    Extractor(const orbx_params& p)
        : nfeatures(p.nfeatures), nlevels(p.nlevels), iniTh(p.ini_th_fast), minTh(p.min_th_fast),
          scaleFactor(p.scale_factor)
    {
        // src/ORBextractor.cc:410-470
        scale.resize(nlevels); sigma2.resize(nlevels);
        scale[0] = 1.0f; sigma2[0] = 1.0f;
        for (int i = 1; i < nlevels; ++i) {
            scale[i] = (float)(scale[i - 1] * scaleFactor);
            sigma2[i] = scale[i] * scale[i];
        }
        invScale.resize(nlevels); invSigma2.resize(nlevels);
        for (int i = 0; i < nlevels; ++i) { invScale[i] = 1.0f / scale[i]; invSigma2[i] = 1.0f / sigma2[i]; }
        perLevel.resize(nlevels);
        float factor = (float)(1.0f / scaleFactor);
        float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
        int sum = 0;
        for (int l = 0; l < nlevels - 1; ++l) {
            perLevel[l] = cv_roundf(nDesired);
            sum += perLevel[l];
            nDesired *= factor;
        }
        perLevel[nlevels - 1] = std::max(nfeatures - sum, 0);
        static const signed char pat[ORBX_PATTERN_LEN] = {ORBX_PATTERN_VALUES};
        std::memcpy(pattern, pat, sizeof(pat));
        umax.assign(HALF_PATCH + 1, 0);
        int v, v0;
        const int vmax = (int)std::floor(HALF_PATCH * std::sqrt(2.f) / 2 + 1);
        const int vmin = (int)std::ceil(HALF_PATCH * std::sqrt(2.f) / 2);
        const double hp2 = HALF_PATCH * HALF_PATCH;
        for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
        for (v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }

    // src/ORBextractor.cc:1107-1132 (the 19-px border is materialised on demand, see get_level)
    void compute_pyramid(const u8* img, int w, int h, size_t stride)
    {
        raw.resize(nlevels);
        for (int l = 0; l < nlevels; ++l) {
            const float s = invScale[l];
            const int lw = cv_roundf((float)w * s), lh = cv_roundf((float)h * s);
            raw[l].alloc(lw, lh);
            if (l == 0) {
                for (int y = 0; y < h; ++y) std::memcpy(raw[0].row(y), img + (size_t)y * stride, w);
            } else {
                resize_linear_u8(raw[l - 1].px.data(), raw[l - 1].w, raw[l - 1].h, raw[l - 1].w,
                                 raw[l].px.data(), lw, lh, lw);
            }
        }
    }

    // src/ORBextractor.cc:77-104
    float ic_angle(const Image& im, float px, float py) const
    {
        int m_01 = 0, m_10 = 0;
        const int cx = cv_roundf(px), cy = cv_roundf(py);
        const u8* center = im.row(cy) + cx;
        for (int u = -HALF_PATCH; u <= HALF_PATCH; ++u) m_10 += u * center[u];
        const int step = im.w;
        for (int v = 1; v <= HALF_PATCH; ++v) {
            int v_sum = 0;
            const int d = umax[v];
            for (int u = -d; u <= d; ++u) {
                const int val_plus = center[u + v * step], val_minus = center[u - v * step];
                v_sum += (val_plus - val_minus);
                m_10 += u * (val_plus + val_minus);
            }
            m_01 += v * v_sum;
        }
        return fast_atan2((float)m_01, (float)m_10);
    }

    // src/ORBextractor.cc:107-147
    void orb_descriptor(const orbx_keypoint& kp, const Image& im, u8* desc, bool libm) const
    {
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        const float angle = (float)kp.angle * factorPI;
        float a, b;
        if (libm) { a = (float)cosf(angle); b = (float)sinf(angle); }
        else orbx_sincosf(angle, &b, &a);
        const u8* center = im.row(cv_roundf(kp.y)) + cv_roundf(kp.x);
        const int step = im.w;
        const signed char* pat = pattern;
        for (int i = 0; i < 32; ++i, pat += 32) {
            int val = 0;
            for (int j = 0; j < 8; ++j) {
                const int x0 = pat[4 * j], y0 = pat[4 * j + 1], x1 = pat[4 * j + 2], y1 = pat[4 * j + 3];
                const int t0 = center[cv_roundf(x0 * b + y0 * a) * step + cv_roundf(x0 * a - y0 * b)];
                const int t1 = center[cv_roundf(x1 * b + y1 * a) * step + cv_roundf(x1 * a - y1 * b)];
                val |= (t0 < t1) << j;
            }
            desc[i] = (u8)val;
        }
    }

    // src/ORBextractor.cc:765-853
    void compute_keypoints_octtree()
    {
        candidates.assign(nlevels, {});
        levelKeys.assign(nlevels, {});
        const float W = 30;
        std::vector<Key> cell;
        for (int level = 0; level < nlevels; ++level) {
            const Image& im = raw[level];
            const int minBorderX = EDGE_TH - 3, minBorderY = minBorderX;
            const int maxBorderX = im.w - EDGE_TH + 3, maxBorderY = im.h - EDGE_TH + 3;
            std::vector<Key>& toDistribute = candidates[level];
            const float width = (float)(maxBorderX - minBorderX);
            const float height = (float)(maxBorderY - minBorderY);
            const int nCols = (int)(width / W), nRows = (int)(height / W);
            if (nCols <= 0 || nRows <= 0) continue;   // reference divides by zero; out of domain
            const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
            for (int i = 0; i < nRows; ++i) {
                const float iniY = (float)(minBorderY + i * hCell);
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBorderY - 3) continue;
                if (maxY > maxBorderY) maxY = (float)maxBorderY;
                for (int j = 0; j < nCols; ++j) {
                    const float iniX = (float)(minBorderX + j * wCell);
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBorderX - 6) continue;
                    if (maxX > maxBorderX) maxX = (float)maxBorderX;
                    const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
                    fast9_16(im.row(y0) + x0, cw, ch, im.w, iniTh, true, cell);
                    if (cell.empty()) fast9_16(im.row(y0) + x0, cw, ch, im.w, minTh, true, cell);
                    for (Key k : cell) {
                        k.x += j * wCell;
                        k.y += i * hCell;
                        toDistribute.push_back(k);
                    }
                }
            }
            std::vector<int> keep = distribute_octtree(toDistribute, minBorderX, maxBorderX, minBorderY,
                                                       maxBorderY, perLevel[level]);
            const int scaledPatchSize = (int)(PATCH_SIZE * scale[level]);
            std::vector<orbx_keypoint>& out = levelKeys[level];
            out.reserve(keep.size());
            for (int idx : keep) {
                orbx_keypoint kp;
                kp.x = toDistribute[idx].x + minBorderX;
                kp.y = toDistribute[idx].y + minBorderY;
                kp.size = (float)scaledPatchSize;
                kp.angle = -1.f;
                kp.response = toDistribute[idx].response;
                kp.octave = level;
                kp.class_id = -1;
                out.push_back(kp);
            }
        }
        for (int level = 0; level < nlevels; ++level)
            for (orbx_keypoint& kp : levelKeys[level]) kp.angle = ic_angle(raw[level], kp.x, kp.y);
    }

    // src/ORBextractor.cc:1043-1105
    int extract(const u8* img, int w, int h, size_t stride, orbx_keypoint* kps, u8* desc, int cap,
                bool libm)
    {
        if (w <= 0 || h <= 0) return 0;
        compute_pyramid(img, w, h, stride);
        compute_keypoints_octtree();
        int total = 0;
        for (int l = 0; l < nlevels; ++l) total += (int)levelKeys[l].size();
        if (total > cap) return -1;
        blur.assign(nlevels, Image());
        int offset = 0;
        for (int l = 0; l < nlevels; ++l) {
            std::vector<orbx_keypoint>& keys = levelKeys[l];
            blur[l].alloc(raw[l].w, raw[l].h);
            gaussian_blur7_u8(raw[l].px.data(), raw[l].w, raw[l].h, raw[l].w, blur[l].px.data(), raw[l].w);
            if (keys.empty()) continue;
            for (size_t i = 0; i < keys.size(); ++i)
                orb_descriptor(keys[i], blur[l], desc + (size_t)(offset + i) * 32, libm);
            for (size_t i = 0; i < keys.size(); ++i) {
                orbx_keypoint kp = keys[i];
                if (l != 0) { kp.x *= scale[l]; kp.y *= scale[l]; }
                kps[offset + i] = kp;
            }
            offset += (int)keys.size();
        }
        return total;
    }
};

// ------------------------------------------------------------------------------------------
// Frame grid (src/Frame.cc:230-245, 327-392) on a flat frame
// ------------------------------------------------------------------------------------------
const int GRID_COLS = 64, GRID_ROWS = 48;   // include/Frame.h:37-38
const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // src/ORBmatcher.cc:37-39

struct Grid {
    float minX, minY, invW, invH;
    std::vector<int> cell[GRID_COLS][GRID_ROWS];
    const orbm_frame* f;
    explicit Grid(const orbm_frame* fr) : f(fr)
    {
        minX = fr->min_x; minY = fr->min_y;
        invW = static_cast<float>(GRID_COLS) / (fr->max_x - fr->min_x);
        invH = static_cast<float>(GRID_ROWS) / (fr->max_y - fr->min_y);
        for (int i = 0; i < fr->n; ++i) {
            const orbx_keypoint& kp = fr->keys_un[i];
            const int px = (int)std::round((kp.x - minX) * invW);
            const int py = (int)std::round((kp.y - minY) * invH);
            if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
            cell[px][py].push_back(i);
        }
    }
    void in_area(float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) const
    {
        out.clear();
        const int nMinCellX = std::max(0, (int)std::floor((x - minX - r) * invW));
        if (nMinCellX >= GRID_COLS) return;
        const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - minX + r) * invW));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)std::floor((y - minY - r) * invH));
        if (nMinCellY >= GRID_ROWS) return;
        const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - minY + r) * invH));
        if (nMaxCellY < 0) return;
        const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ++ix)
            for (int iy = nMinCellY; iy <= nMaxCellY; ++iy)
                for (int idx : cell[ix][iy]) {
                    const orbx_keypoint& kp = f->keys_un[idx];
                    if (checkLevels) {
                        if (kp.octave < minLevel) continue;
                        if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                    }
                    const float dx = kp.x - x, dy = kp.y - y;
                    if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(idx);
                }
    }
};

// src/ORBmatcher.cc:1647-1663
inline int descriptor_distance(const u8* a, const u8* b)
{
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t pa, pb;
        std::memcpy(&pa, a + 4 * i, 4);
        std::memcpy(&pb, b + 4 * i, 4);
        unsigned v = pa ^ pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// src/ORBmatcher.cc:1601-1642
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; ++i) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

inline int rot_bin(float a1, float a2)
{
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

} // namespace

// ==========================================================================================
// C interface
// ==========================================================================================
struct oracle_extractor { Extractor ex; explicit oracle_extractor(const orbx_params& p) : ex(p) {} };

extern "C" {

void oracle_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst,
                             int dw, int dh, size_t dstride)
{ resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride); }

void oracle_gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst,
                              size_t dstride)
{ gaussian_blur7_u8(src, w, h, sstride, dst, dstride); }

void oracle_copy_make_border101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst,
                                size_t dstride, int border)
{
    for (int y = -border; y < h + border; ++y) {
        const u8* s = src + (size_t)reflect101(y, h) * sstride;
        u8* d = dst + (size_t)(y + border) * dstride;
        for (int x = -border; x < w + border; ++x) d[x + border] = s[reflect101(x, w)];
    }
}

int oracle_fast9_16(const uint8_t* img, int w, int h, size_t stride, int th, int nms, float* xyr,
                    int cap)
{
    std::vector<Key> k;
    fast9_16(img, w, h, stride, th, nms != 0, k);
    for (size_t i = 0; i < k.size() && (int)i < cap; ++i) {
        xyr[3 * i] = k[i].x; xyr[3 * i + 1] = k[i].y; xyr[3 * i + 2] = k[i].response;
    }
    return (int)k.size();
}

float oracle_fast_atan2(float y, float x) { return fast_atan2(y, x); }

void oracle_sincos(float x, float* s, float* c, int use_libm)
{
    if (use_libm) { *s = sinf(x); *c = cosf(x); }
    else orbx_sincosf(x, s, c);
}

int oracle_cv_round(double v) { return cv_round(v); }

oracle_extractor* oracle_extractor_create(const orbx_params* p) { return new oracle_extractor(*p); }
void oracle_extractor_destroy(oracle_extractor* e) { delete e; }

void oracle_extractor_tables(const oracle_extractor* e, float* scale, float* inv_scale,
                             float* sigma2, float* inv_sigma2, int32_t* features_per_level,
                             int32_t* umax16)
{
    const Extractor& x = e->ex;
    for (int i = 0; i < x.nlevels; ++i) {
        if (scale) scale[i] = x.scale[i];
        if (inv_scale) inv_scale[i] = x.invScale[i];
        if (sigma2) sigma2[i] = x.sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = x.invSigma2[i];
        if (features_per_level) features_per_level[i] = x.perLevel[i];
    }
    if (umax16) for (int i = 0; i < 16; ++i) umax16[i] = x.umax[i];
}

int oracle_extract(oracle_extractor* e, const uint8_t* img, int w, int h, size_t stride,
                   orbx_keypoint* kps, uint8_t* desc, int cap, int use_libm_sincos)
{ return e->ex.extract(img, w, h, stride, kps, desc, cap, use_libm_sincos != 0); }

int oracle_level_size(const oracle_extractor* e, int level, int* w, int* h)
{
    if (level < 0 || level >= (int)e->ex.raw.size()) return -1;
    *w = e->ex.raw[level].w; *h = e->ex.raw[level].h;
    return 0;
}

int oracle_get_level(const oracle_extractor* e, int level, int blurred, uint8_t* dst, size_t dstride)
{
    const std::vector<Image>& v = blurred ? e->ex.blur : e->ex.raw;
    if (level < 0 || level >= (int)v.size()) return -1;
    const Image& im = v[level];
    for (int y = 0; y < im.h; ++y) std::memcpy(dst + (size_t)y * dstride, im.row(y), im.w);
    return 0;
}

int oracle_get_candidates(const oracle_extractor* e, int level, float* xyr, int cap)
{
    if (level < 0 || level >= (int)e->ex.candidates.size()) return -1;
    const std::vector<Key>& k = e->ex.candidates[level];
    for (size_t i = 0; i < k.size() && (int)i < cap; ++i) {
        xyr[3 * i] = k[i].x; xyr[3 * i + 1] = k[i].y; xyr[3 * i + 2] = k[i].response;
    }
    return (int)k.size();
}

int oracle_get_level_keypoints(const oracle_extractor* e, int level, orbx_keypoint* kps, int cap)
{
    if (level < 0 || level >= (int)e->ex.levelKeys.size()) return -1;
    const std::vector<orbx_keypoint>& k = e->ex.levelKeys[level];
    for (size_t i = 0; i < k.size() && (int)i < cap; ++i) kps[i] = k[i];
    return (int)k.size();
}

int oracle_distribute_octtree(const float* xyr, int n, int minX, int maxX, int minY, int maxY,
                              int N, int32_t* out_idx, int cap)
{
    std::vector<Key> cand(n);
    for (int i = 0; i < n; ++i) cand[i] = Key{xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
    std::vector<int> keep = distribute_octtree(cand, minX, maxX, minY, maxY, N);
    for (size_t i = 0; i < keep.size() && (int)i < cap; ++i) out_idx[i] = keep[i];
    return (int)keep.size();
}

int oracle_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

int oracle_features_in_area(const orbm_frame* f, float x, float y, float r, int min_level,
                            int max_level, int32_t* out, int cap)
{
    Grid g(f);
    std::vector<int> v;
    g.in_area(x, y, r, min_level, max_level, v);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = v[i];
    return (int)v.size();
}

// src/ORBmatcher.cc:405-520
int oracle_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2,
                                     float* prev_matched, int32_t* matches12, int window_size,
                                     float nnratio, int check_orientation, uint64_t* ncompares)
{
    int nmatches = 0;
    uint64_t ncmp = 0;
    for (int i = 0; i < f1->n; ++i) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> matchedDistance(f2->n, INT_MAX), matches21(f2->n, -1);
    Grid grid2(f2);
    std::vector<int> indices2;
    for (int i1 = 0; i1 < f1->n; ++i1) {
        const orbx_keypoint& kp1 = f1->keys_un[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;
        grid2.in_area(prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window_size, level1, level1, indices2);
        if (indices2.empty()) continue;
        const u8* d1 = f1->descriptors + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : indices2) {
            const int dist = descriptor_distance(d1, f2->descriptors + (size_t)i2 * 32);
            ++ncmp;
            if (matchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                matches21[bestIdx2] = i1;
                matchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation)
                    rotHist[rot_bin(f1->keys_un[i1].angle, f2->keys_un[bestIdx2].angle)].push_back(i1);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < f1->n; ++i1)
        if (matches12[i1] >= 0) {
            prev_matched[2 * i1] = f2->keys_un[matches12[i1]].x;
            prev_matched[2 * i1 + 1] = f2->keys_un[matches12[i1]].y;
        }
    if (ncompares) *ncompares = ncmp;
    return nmatches;
}

// The four SearchByProjection variants on flattened queries (src/ORBmatcher.cc:45-129, 290-403,
// 1328-1470, 1472-1599).  Projection / frustum / distance tests are done by the caller.
int oracle_search_by_projection(int variant, const orbm_frame* frame,
                                const orbm_proj_query* queries, const uint8_t* query_desc, int nq,
                                const uint8_t* query_blocks, uint8_t* occupied, int32_t* assigned,
                                int32_t* best_idx, int32_t* best_dist, int th_dist, float nnratio,
                                int check_orientation)
{
    int nmatches = 0;
    Grid grid(frame);
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> indices;
    for (int i = 0; i < frame->n; ++i) assigned[i] = -1;
    const bool rot = check_orientation && (variant == ORBM_PROJ_LASTFRAME || variant == ORBM_PROJ_KEYFRAME);
    for (int q = 0; q < nq; ++q) {
        const orbm_proj_query& Q = queries[q];
        best_idx[q] = -1; best_dist[q] = 256;
        grid.in_area(Q.u, Q.v, Q.radius, Q.min_level, Q.max_level, indices);
        if (indices.empty()) continue;
        const u8* dq = query_desc + (size_t)q * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : indices) {
            if (occupied[idx]) continue;
            if (variant == ORBM_PROJ_LOCALMAP || variant == ORBM_PROJ_LASTFRAME) {
                const float ur = frame->u_right ? frame->u_right[idx] : -1.f;
                if (ur > 0 && Q.max_right_err >= 0) {
                    const float er = std::fabs(Q.u_right - ur);
                    if (er > Q.max_right_err) continue;
                }
            }
            const int oct = frame->keys_un[idx].octave;
            if (Q.level_lo >= 0 && (oct < Q.level_lo || oct > Q.level_hi)) continue;
            const int dist = descriptor_distance(dq, frame->descriptors + (size_t)idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = oct; bestDist2 = dist;
            }
        }
        if (bestDist <= th_dist) {
            if (variant == ORBM_PROJ_LOCALMAP && bestLevel == bestLevel2 && bestDist > nnratio * bestDist2)
                continue;
            assigned[bestIdx] = q;
            const bool blocks = (variant == ORBM_PROJ_KEYFRAME || variant == ORBM_PROJ_SIM3)
                                    ? true : (query_blocks ? query_blocks[q] != 0 : true);
            occupied[bestIdx] = blocks ? 1 : 0;
            best_idx[q] = bestIdx; best_dist[q] = bestDist;
            nmatches++;
            if (rot) rotHist[rot_bin(Q.angle, frame->keys_un[bestIdx].angle)].push_back(bestIdx);
        }
    }
    if (rot) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { assigned[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// src/Frame.cc:466-638
int oracle_compute_stereo_matches(const orbm_stereo_input* in, float* u_right, float* depth,
                                  uint64_t* ncompares)
{
    const int N = in->n_left, Nr = in->n_right;
    uint64_t ncmp = 0;
    for (int i = 0; i < N; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; }
    const int nRows = in->level_height[0];
    std::vector<std::vector<int>> rowIdx(nRows);
    for (int iR = 0; iR < Nr; ++iR) {
        const orbx_keypoint& kp = in->keys_right[iR];
        const float kpY = kp.y;
        const float r = 2.0f * in->scale_factors[kp.octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);   // (reference has no bounds check)
    }
    const float mb = in->bf / in->fx;
    const float minZ = mb, minD = -3, maxD = in->bf / minZ;
    std::vector<std::pair<int, int>> distIdx;
    auto pix = [&](const u8* const* pyr, const int64_t* strides, int lvl, int x, int y) -> int {
        const int w = in->level_width[lvl], h = in->level_height[lvl];
        return pyr[lvl][(size_t)reflect101(y, h) * strides[lvl] + reflect101(x, w)];
    };
    for (int iL = 0; iL < N; ++iL) {
        const orbx_keypoint& kpL = in->keys_left[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        const std::vector<int>& cands = rowIdx[row];
        if (cands.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        int bestIdxR = 0;
        const u8* dL = in->desc_left + (size_t)iL * 32;
        for (int iR : cands) {
            const orbx_keypoint& kpR = in->keys_right[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, in->desc_right + (size_t)iR * 32);
                ++ncmp;
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < TH_HIGH) {
            const float uR0 = in->keys_right[bestIdxR].x;
            const float sf = in->inv_scale_factors[kpL.octave];
            const float scaleduL = std::round(kpL.x * sf);
            const float scaledvL = std::round(kpL.y * sf);
            const float scaleduR0 = std::round(uR0 * sf);
            const int w = 5, L = 5;
            int bestSad = INT_MAX, bestincR = 0;
            float dists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= in->level_width[kpL.octave]) continue;
            const int cuL = (int)scaleduL, cvL = (int)scaledvL, cuR = (int)scaleduR0;
            const int cL = pix(in->pyr_left, in->pyr_stride_left, kpL.octave, cuL, cvL);
            for (int incR = -L; incR <= L; ++incR) {
                const int cR = pix(in->pyr_right, in->pyr_stride_right, kpL.octave, cuR + incR, cvL);
                float dist = 0.f;
                for (int dy = -w; dy <= w; ++dy)
                    for (int dx = -w; dx <= w; ++dx) {
                        const float a = (float)(pix(in->pyr_left, in->pyr_stride_left, kpL.octave, cuL + dx, cvL + dy) - cL);
                        const float b = (float)(pix(in->pyr_right, in->pyr_stride_right, kpL.octave, cuR + incR + dx, cvL + dy) - cR);
                        dist += std::fabs(a - b);
                    }
                if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
                dists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = dists[L + bestincR - 1], dist2 = dists[L + bestincR], dist3 = dists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = in->scale_factors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= 0 && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = in->bf / disparity;
                u_right[iL] = bestuR;
                distIdx.push_back(std::pair<int, int>(bestSad, iL));
            }
        }
    }
    if (!distIdx.empty()) {     // the reference indexes an empty vector here (UB); guarded
        std::sort(distIdx.begin(), distIdx.end());
        const float median = (float)distIdx[distIdx.size() / 2].first;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = (int)distIdx.size() - 1; i >= 0; --i) {
            if (distIdx[i].first < thDist) break;
            u_right[distIdx[i].second] = -1;
            depth[distIdx[i].second] = -1;
        }
    }
    if (ncompares) *ncompares = ncmp;
    return 0;
}

// brute-force 2-NN (BASELINE.json config 5); scan order = ascending database index, strict '<'
void oracle_knn2(const uint8_t* queries, int64_t nq, const uint8_t* database, int64_t ndb,
                 int32_t db_index_offset, uint16_t* out_dist, int32_t* out_idx, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    auto work = [&](int64_t q0, int64_t q1) {
        for (int64_t q = q0; q < q1; ++q) {
            uint64_t a[4];
            std::memcpy(a, queries + q * 32, 32);
            int d1 = 257, d2 = 257; int64_t i1 = -1, i2 = -1;
            for (int64_t j = 0; j < ndb; ++j) {
                uint64_t b[4];
                std::memcpy(b, database + j * 32, 32);
                const int d = __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
                              __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
                if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = j; }
                else if (d < d2) { d2 = d; i2 = j; }
            }
            out_dist[2 * q] = (uint16_t)d1; out_dist[2 * q + 1] = (uint16_t)d2;
            out_idx[2 * q] = i1 < 0 ? -1 : (int32_t)(i1 + db_index_offset);
            out_idx[2 * q + 1] = i2 < 0 ? -1 : (int32_t)(i2 + db_index_offset);
        }
    };
    if (nthreads == 1) { work(0, nq); return; }
    std::vector<std::thread> th;
    const int64_t per = (nq + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        const int64_t q0 = t * per, q1 = std::min<int64_t>(nq, q0 + per);
        if (q0 < q1) th.emplace_back(work, q0, q1);
    }
    for (auto& t : th) t.join();
}

int64_t oracle_extract_many(const orbx_params* p, const uint8_t* imgs, size_t frame_stride, int w,
                            int h, size_t stride, int nframes, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next(0);
    std::atomic<int64_t> total(0);
    auto work = [&]() {
        Extractor ex(*p);
        const int cap = p->nfeatures + 64 * p->nlevels + 64;
        std::vector<orbx_keypoint> kps(cap);
        std::vector<u8> desc((size_t)cap * 32);
        for (;;) {
            const int f = next.fetch_add(1);
            if (f >= nframes) break;
            const int n = ex.extract(imgs + (size_t)f * frame_stride, w, h, stride, kps.data(), desc.data(), cap, false);
            if (n > 0) total += n;
        }
    };
    if (nthreads == 1) { work(); return total; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work);
    for (auto& t : th) t.join();
    return total;
}

// CPU baseline driver for BASELINE.json config 2: extract every frame (one frame per thread at a
// time) and run SearchForInitialization on consecutive pairs with vbPrevMatched = F1 keypoints.
int64_t oracle_extract_match_many(const orbx_params* p, const uint8_t* imgs, size_t frame_stride,
                                  int w, int h, size_t stride, int nframes, int nthreads,
                                  int window_size, float nnratio, int check_orientation,
                                  int64_t* total_matches, uint64_t* ncompares)
{
    if (nthreads < 1) nthreads = 1;
    const int cap = p->nfeatures + 64 * p->nlevels + 64;
    std::vector<std::vector<orbx_keypoint>> K(nframes);
    std::vector<std::vector<u8>> D(nframes);
    std::atomic<int> next(0);
    auto extract_work = [&]() {
        Extractor ex(*p);
        std::vector<orbx_keypoint> kps(cap);
        std::vector<u8> desc((size_t)cap * 32);
        for (;;) {
            const int f = next.fetch_add(1);
            if (f >= nframes) break;
            const int n = std::max(0, ex.extract(imgs + (size_t)f * frame_stride, w, h, stride, kps.data(), desc.data(), cap, false));
            K[f].assign(kps.begin(), kps.begin() + n);
            D[f].assign(desc.begin(), desc.begin() + (size_t)n * 32);
        }
    };
    std::atomic<int> nextPair(0);
    std::atomic<int64_t> matches(0);
    std::atomic<uint64_t> cmps(0);
    auto match_work = [&]() {
        std::vector<float> prev;
        std::vector<int32_t> m12;
        for (;;) {
            const int i = nextPair.fetch_add(1);
            if (i >= nframes - 1) break;
            orbm_frame f1{(int32_t)K[i].size(), K[i].data(), D[i].data(), 0.f, (float)w, 0.f, (float)h, nullptr, nullptr, 0};
            orbm_frame f2{(int32_t)K[i + 1].size(), K[i + 1].data(), D[i + 1].data(), 0.f, (float)w, 0.f, (float)h, nullptr, nullptr, 0};
            prev.resize(2 * K[i].size() + 2);
            m12.resize(K[i].size() + 1);
            for (size_t k = 0; k < K[i].size(); ++k) { prev[2 * k] = K[i][k].x; prev[2 * k + 1] = K[i][k].y; }
            uint64_t c = 0;
            matches += oracle_search_for_initialization(&f1, &f2, prev.data(), m12.data(), window_size, nnratio, check_orientation, &c);
            cmps += c;
        }
    };
    auto run = [&](auto fn) {
        if (nthreads == 1) { fn(); return; }
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(fn);
        for (auto& t : th) t.join();
    };
    run(extract_work);
    run(match_work);
    int64_t total = 0;
    for (int f = 0; f < nframes; ++f) total += (int64_t)K[f].size();
    if (total_matches) *total_matches = matches;
    if (ncompares) *ncompares = cmps;
    return total;
}

} // extern "C"
