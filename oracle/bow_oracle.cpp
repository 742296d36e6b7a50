// bow_oracle.cpp — CPU restatement of the bag-of-words matcher path.
// TEST INFRASTRUCTURE ONLY (tests/, smoke, bench CPU legs); liborbx.so never links it.
//
//   oracle_bow_transform      DBoW2 TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)
//                             Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1231-1271 with FORB::distance
//                             (FORB.cpp:81-101), per feature, as ::transform(features, v, fv, levelsup) :1140-1218 calls it
//   oracle_distinctive        MapPoint::ComputeDistinctiveDescriptors                src/MapPoint.cc:242-305
//   oracle_search_by_bow      ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...)        src/ORBmatcher.cc:159-288 (variant 0)
//                             ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...)     src/ORBmatcher.cc:522-655 (variant 1)
//
// Parity: unpinned by reference execution (ORBmatcher.cc needs PCL/g2o, the ORB vocabulary file is not shipped:
// .MISSING_LARGE_BLOBS); the control flow follows the reference line by line with std::map in node order exactly
// like DBoW2::FeatureVector, and tests/test_oracle_bow.py checks it against an independent Python statement.

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

namespace {
const int TH_LOW = 50, HISTO_LENGTH = 30;

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
} // namespace

extern "C" {

// Flattened vocabulary tree: node 0 = root; children of node i = children[child_begin[i] .. +child_count[i]);
// child_count == 0 marks a leaf (word).  depth = m_L.
void oracle_bow_transform(int nnodes, const int32_t* child_begin, const int32_t* child_count, const int32_t* children,
                          const uint8_t* node_desc, const uint32_t* node_word, const double* node_weight, int depth,
                          const uint8_t* features, int n, int levelsup, uint32_t* word_id, double* weight, uint32_t* node_id)
{
    (void)nnodes;
    const int nid_level = depth - levelsup;
    for (int f = 0; f < n; ++f) {
        const uint8_t* feat = features + (size_t)f * 32;
        uint32_t nid = 0;                      // nid_level <= 0 -> root
        int final_id = 0, current_level = 0;
        do {
            ++current_level;
            const int32_t* nodes = children + child_begin[final_id];
            const int cnt = child_count[final_id];
            final_id = nodes[0];
            double best_d = distance256(feat, node_desc + (size_t)final_id * 32);
            for (int c = 1; c < cnt; ++c) {
                const int id = nodes[c];
                const double d = distance256(feat, node_desc + (size_t)id * 32);
                if (d < best_d) { best_d = d; final_id = id; }
            }
            if (current_level == nid_level) nid = (uint32_t)final_id;
        } while (child_count[final_id] != 0);
        word_id[f] = node_word[final_id];
        weight[f] = node_weight[final_id];
        node_id[f] = nid;
    }
}

// flags: bit 0 = the feature is in the FeatureVector (its word weight is > 0), bit 1 = it has a good MapPoint.
// variant 0: A = keyframe, B = frame; matchB[iB] = iA (vpMapPointMatches[realIdxF] = pMP of KF feature iA)
// variant 1: A = KF1, B = KF2;        matchA[iA] = iB (vpMatches12[idx1] = MapPoint of KF2 feature iB)
// Both arrays are filled in both variants (-1 = none).  Returns nmatches.
int oracle_search_by_bow(int variant, int nA, const uint8_t* descA, const float* angleA, const uint32_t* nodeA, const uint8_t* flagsA,
                         int nB, const uint8_t* descB, const float* angleB, const uint32_t* nodeB, const uint8_t* flagsB,
                         float nnratio, int check_orientation, int32_t* matchA, int32_t* matchB)
{
    std::map<uint32_t, std::vector<unsigned>> fvA, fvB;          // DBoW2::FeatureVector: node id -> feature indices (ascending)
    for (int i = 0; i < nA; ++i) if (flagsA[i] & 1) fvA[nodeA[i]].push_back((unsigned)i);
    for (int i = 0; i < nB; ++i) if (flagsB[i] & 1) fvB[nodeB[i]].push_back((unsigned)i);
    for (int i = 0; i < nA; ++i) matchA[i] = -1;
    for (int i = 0; i < nB; ++i) matchB[i] = -1;
    std::vector<bool> matchedB(nB, false);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    auto ait = fvA.begin(), bit = fvB.begin();
    while (ait != fvA.end() && bit != fvB.end()) {
        if (ait->first == bit->first) {
            for (size_t ia = 0; ia < ait->second.size(); ++ia) {
                const unsigned idxA = ait->second[ia];
                if (!(flagsA[idxA] & 2)) continue;                       // !pMP || pMP->isBad()
                const uint8_t* dA = descA + (size_t)idxA * 32;
                int bestDist1 = 256, bestIdxB = -1, bestDist2 = 256;
                for (size_t ib = 0; ib < bit->second.size(); ++ib) {
                    const unsigned idxB = bit->second[ib];
                    if (matchedB[idxB]) continue;                        // vpMapPointMatches[realIdxF] / vbMatched2[idx2]
                    if (variant == 1 && !(flagsB[idxB] & 2)) continue;   // !pMP2 || pMP2->isBad()
                    const int dist = distance256(dA, descB + (size_t)idxB * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxB = (int)idxB; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                const bool close = variant == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;       // :229 vs :597
                if (close && (float)bestDist1 < nnratio * (float)bestDist2) {
                    matchA[idxA] = bestIdxB; matchB[bestIdxB] = (int)idxA; matchedB[bestIdxB] = true;
                    if (check_orientation) {
                        float rot = angleA[idxA] - angleB[bestIdxB];
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idxA);
                    }
                    ++nmatches;
                }
            }
            ++ait; ++bit;
        } else if (ait->first < bit->first) ait = fvA.lower_bound(bit->first);
        else bit = fvB.lower_bound(ait->first);
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idxA : rotHist[i]) { matchB[matchA[idxA]] = -1; matchA[idxA] = -1; --nmatches; }
        }
    }
    return nmatches;
}

// MapPoint::ComputeDistinctiveDescriptors for npoints map points: point p owns descriptors offsets[p] .. offsets[p+1]
// (its observations in good keyframes, in the std::map order the caller walked); best_idx[p] = index inside the group of
// the descriptor with the least median distance to the others (-1 for an empty group), best_median[p] = that median.
void oracle_distinctive(const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best_idx, int32_t* best_median)
{
    for (int p = 0; p < npoints; ++p) {
        const uint8_t* d = desc + (size_t)offsets[p] * 32;
        const size_t N = (size_t)(offsets[p + 1] - offsets[p]);
        if (N == 0) { best_idx[p] = -1; best_median[p] = -1; continue; }
        std::vector<float> Distances(N * N);
        for (size_t i = 0; i < N; i++) {
            Distances[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                const int distij = distance256(d + i * 32, d + j * 32);
                Distances[i * N + j] = (float)distij;
                Distances[j * N + i] = (float)distij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + (i + 1) * N);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
        }
        best_idx[p] = BestIdx; best_median[p] = BestMedian;
    }
}

} // extern "C"
