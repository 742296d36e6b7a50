// kfmatch_oracle.cpp — CPU restatement of the keyframe-rate matchers.
// TEST INFRASTRUCTURE ONLY (tests/, smoke, bench CPU legs); liborbx.so never links it.
//
//   oracle_search_for_triangulation   ORBmatcher::SearchForTriangulation   src/ORBmatcher.cc:657-823
//                                     ORBmatcher::CheckDistEpipolarLine    src/ORBmatcher.cc:140-157
//   oracle_fuse                       the per-MapPoint search of ORBmatcher::Fuse x2 (:825-973, :975-1100) and of one
//                                     direction of SearchBySim3, over KeyFrame::GetFeaturesInArea src/KeyFrame.cc:607-646
//   oracle_search_by_sim3             ORBmatcher::SearchBySim3 agreement   src/ORBmatcher.cc:1102-1326
//
// Parity: unpinned by reference execution (ORBmatcher.cc / KeyFrame.cc need the whole SLAM graph, g2o and PCL);
// the control flow follows the reference line by line and tests/test_oracle_kfmatch.py checks it against an
// independent Python statement.  Compile with -ffp-contract=off: every float operation is separately rounded.

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "../include/orbx.h"

namespace {
const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
const int GRID_COLS = 64, GRID_ROWS = 48;                        // include/Frame.h:37-38

int distance256(const uint8_t* a, const uint8_t* b)
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

bool check_dist_epipolar_line(const orbx_keypoint& kp1, const orbx_keypoint& kp2, const float* F12, const float* levelSigma2)
{
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * levelSigma2[kp2.octave];
}

// The keyframe grid is the Frame's (KeyFrame.cc:43-51 copies F.mGrid): Frame::AssignFeaturesToGrid / PosInGrid
// (src/Frame.cc:230-245, 382-392), then KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:607-646).
struct KfGrid {
    float minX, minY, invW, invH;
    std::vector<int> cell[GRID_COLS][GRID_ROWS];
    const orbm_frame* f;
    explicit KfGrid(const orbm_frame* fr) : f(fr)
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
    std::vector<size_t> in_area(const float& x, const float& y, const float& r) const
    {
        std::vector<size_t> vIndices;
        const int nMinCellX = std::max(0, (int)std::floor((x - minX - r) * invW));
        if (nMinCellX >= GRID_COLS) return vIndices;
        const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - minX + r) * invW));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = std::max(0, (int)std::floor((y - minY - r) * invH));
        if (nMinCellY >= GRID_ROWS) return vIndices;
        const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - minY + r) * invH));
        if (nMaxCellY < 0) return vIndices;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                for (int idx : cell[ix][iy]) {
                    const orbx_keypoint& kpUn = f->keys_un[idx];
                    const float distx = kpUn.x - x, disty = kpUn.y - y;
                    if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back((size_t)idx);
                }
        return vIndices;
    }
};

void fuse_queries(int variant, const orbm_keyframe* kf, const orbm_proj_query* queries, const uint8_t* qdesc, int nq,
                  int32_t* best_idx, int32_t* best_dist)
{
    const KfGrid grid(&kf->frame);
    const orbm_frame& F = kf->frame;
    for (int q = 0; q < nq; ++q) {
        const orbm_proj_query& Q = queries[q];
        const float u = Q.u, v = Q.v, ur = Q.u_right;
        const int nPredictedLevel = Q.level_hi;
        best_idx[q] = -1; best_dist[q] = 256;
        const std::vector<size_t> vIndices = grid.in_area(u, v, Q.radius);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = qdesc + (size_t)q * 32;
        int bestDist = 256, bestIdx = -1;           // INT_MAX in the Sim3 variants: same outcome, distances are <= 256
        for (size_t idx : vIndices) {
            const orbx_keypoint& kp = F.keys_un[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (variant == ORBM_FUSE_MAPPOINTS) {
                if (F.u_right && F.u_right[idx] >= 0) {
                    const float ex = u - kp.x, ey = v - kp.y, er = ur - F.u_right[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * kf->inv_level_sigma2[kpLevel] > 7.8) continue;
                } else {
                    const float ex = u - kp.x, ey = v - kp.y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * kf->inv_level_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = distance256(dMP, F.descriptors + idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        best_dist[q] = bestDist;
        if (bestDist <= (variant == ORBM_FUSE_SIM3_LEG ? TH_HIGH : TH_LOW)) best_idx[q] = bestIdx;
    }
}
} // namespace

extern "C" {

void oracle_fuse(int variant, const orbm_keyframe* kf, const orbm_proj_query* queries, const uint8_t* qdesc, int nq,
                 int32_t* best_idx, int32_t* best_dist)
{
    fuse_queries(variant, kf, queries, qdesc, nq, best_idx, best_dist);
}

int oracle_search_by_sim3(const orbm_keyframe* kf1, const orbm_keyframe* kf2, const orbm_proj_query* queries12,
                          const uint8_t* desc12, const int32_t* index1, int nq12, const orbm_proj_query* queries21,
                          const uint8_t* desc21, const int32_t* index2, int nq21, int32_t* match12)
{
    const int N1 = kf1->frame.n, N2 = kf2->frame.n;
    std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
    std::vector<int32_t> bi(std::max(std::max(nq12, nq21), 1)), bd(bi.size());
    fuse_queries(ORBM_FUSE_SIM3_LEG, kf2, queries12, desc12, nq12, bi.data(), bd.data());
    for (int q = 0; q < nq12; ++q) vnMatch1[index1[q]] = bi[q];
    fuse_queries(ORBM_FUSE_SIM3_LEG, kf1, queries21, desc21, nq21, bi.data(), bd.data());
    for (int q = 0; q < nq21; ++q) vnMatch2[index2[q]] = bi[q];
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

int oracle_search_for_triangulation(const orbm_keyframe* kf1, const orbm_keyframe* kf2, const float* F12, float ex, float ey,
                                    int bOnlyStereo, int check_orientation, int32_t* vMatches12)
{
    const orbm_frame &A = kf1->frame, &B = kf2->frame;
    std::map<uint32_t, std::vector<unsigned>> vFeatVec1, vFeatVec2;
    for (int i = 0; i < A.n; ++i) if (kf1->flags[i] & 1) vFeatVec1[kf1->node_id[i]].push_back((unsigned)i);
    for (int i = 0; i < B.n; ++i) if (kf2->flags[i] & 1) vFeatVec2[kf2->node_id[i]].push_back((unsigned)i);
    int nmatches = 0;
    std::vector<bool> vbMatched2(B.n, false);                     // never set by the reference either
    for (int i = 0; i < A.n; ++i) vMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    auto f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    while (f1it != vFeatVec1.end() && f2it != vFeatVec2.end()) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0; i1 < f1it->second.size(); i1++) {
                const size_t idx1 = f1it->second[i1];
                if (kf1->flags[idx1] & 2) continue;                 // pMP1
                const bool bStereo1 = A.u_right && A.u_right[idx1] >= 0;
                if (bOnlyStereo) if (!bStereo1) continue;
                const orbx_keypoint& kp1 = A.keys_un[idx1];
                const uint8_t* d1 = A.descriptors + idx1 * 32;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (size_t i2 = 0; i2 < f2it->second.size(); i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || (kf2->flags[idx2] & 2)) continue;
                    const bool bStereo2 = B.u_right && B.u_right[idx2] >= 0;
                    if (bOnlyStereo) if (!bStereo2) continue;
                    const int dist = distance256(d1, B.descriptors + idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orbx_keypoint& kp2 = B.keys_un[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * B.scale_factors[kp2.octave]) continue;
                    }
                    if (check_dist_epipolar_line(kp1, kp2, F12, kf2->level_sigma2)) { bestIdx2 = (int)idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    const orbx_keypoint& kp2 = B.keys_un[bestIdx2];
                    vMatches12[idx1] = bestIdx2;
                    nmatches++;
                    if (check_orientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) f1it = vFeatVec1.lower_bound(f2it->first);
        else f2it = vFeatVec2.lower_bound(f1it->first);
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { vMatches12[j] = -1; nmatches--; }
        }
    }
    return nmatches;
}

} // extern "C"
