// orbx.hpp — host-side C++ mirror of the reference's ORB front-end classes over the C ABI
// (include/orbx.h -> liborbx.so).  Header-only.
//
// The reference has no plugin/FFI layer: its boundary for this path is the C++ class API
//   ORB_SLAM2::ORBextractor  (reference include/ORBextractor.h:45-112)
//   ORB_SLAM2::ORBmatcher    (reference include/ORBmatcher.h:36-101)
//   Frame::ExtractORB / Frame::ComputeStereoMatches (reference src/Frame.cc:247,466)
// The classes below keep those names, argument meanings and error behaviour (void/int returns, an
// empty image returns silently) so the call sites in Frame.cc / Tracking.cc stay as they are.
//
// Two flavours of the same classes:
//   * default: self-contained POD types (orbx::KeyPoint has cv::KeyPoint's 28-byte layout,
//     orbx::Image is a non-owning 8-bit view).  Used by the tests and the bench here, where no
//     OpenCV C++ headers exist.
//   * -DORBX_WITH_OPENCV (after including the OpenCV headers): adds overloads with the reference's
//     exact signatures on cv::InputArray / cv::OutputArray / cv::KeyPoint / cv::Mat.
//
// There is no CPU fallback: a CUDA failure (including "no device") throws std::runtime_error from
// the constructor or the call — the rest of ORB-SLAM2 cannot continue without features anyway.
#ifndef ORBX_HPP
#define ORBX_HPP

#include <cstdint>
#include <algorithm>
#include <cstring>
#include <utility>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/orbx.h"

namespace orbx {

struct KeyPoint {                      // == cv::KeyPoint, field for field
    float x = 0, y = 0, size = 0, angle = -1, response = 0;
    int32_t octave = 0, class_id = -1;
};
static_assert(sizeof(KeyPoint) == sizeof(orbx_keypoint), "KeyPoint layout");

struct Point2f { float x, y; };

struct Image {                         // non-owning 8UC1 view (rows x cols, row stride in bytes)
    const uint8_t* data = nullptr;
    int rows = 0, cols = 0;
    size_t step = 0;
    bool empty() const { return !data || rows <= 0 || cols <= 0; }
};

struct OwnedImage {                    // one pyramid level copied to the host
    std::vector<uint8_t> buf;
    int rows = 0, cols = 0;
    size_t step = 0;
    Image view() const { return Image{buf.data(), rows, cols, step}; }
};

inline void check(int rc, const char* what)
{
    if (rc != ORBX_OK) throw std::runtime_error(std::string(what) + ": " + orbx_last_error());
}

// ---------------------------------------------------------------------------------------------
// ORBextractor (reference include/ORBextractor.h:45).  One object = one CUDA stream + workspace,
// to be used by one thread at a time, exactly like the reference object (src/Frame.cc:78-79 runs
// two objects on two threads).
// ---------------------------------------------------------------------------------------------
class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                 int maxWidth = 1920, int maxHeight = 1080, int device = 0)
        : nlevels_(nlevels), scaleFactor_(scaleFactor)
    {
        orbx_params p{nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST};
        check(orbx_create(&p, maxWidth, maxHeight, 1, 1, device, nullptr, &h_), "ORBextractor");
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        check(orbx_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                              mvInvLevelSigma2.data(), nullptr), "orbx_get_tables");
        cap_ = orbx_keypoint_capacity(h_);
        mvImagePyramid.resize(nlevels);
    }
    ~ORBextractor() { orbx_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image (reference ORBextractor.h:56-61).
    // Mask is ignored, as in the reference.  descriptors: N x 32 bytes, row-major.
    void operator()(const Image& image, const Image& /*mask*/, std::vector<KeyPoint>& keypoints,
                    std::vector<uint8_t>& descriptors)
    {
        if (image.empty()) return;                        // src/ORBextractor.cc:1046-1047
        if (image.cols != shapeW_ || image.rows != shapeH_) {
            // a new shape: the handle lays out its workspace and reports the capacity this shape needs
            const int cap = orbx_reserve_shape(h_, image.cols, image.rows);
            check(cap < 0 ? cap : ORBX_OK, "ORBextractor::operator()");
            cap_ = cap; shapeW_ = image.cols; shapeH_ = image.rows;
        }
        kbuf_.resize(cap_); dbuf_.resize((size_t)cap_ * 32);
        int n = 0;
        check(orbx_extract(h_, image.data, image.cols, image.rows, image.step,
                           reinterpret_cast<orbx_keypoint*>(kbuf_.data()), dbuf_.data(), cap_, &n), "ORBextractor::operator()");
        keypoints.assign(kbuf_.begin(), kbuf_.begin() + n);             // _keypoints.clear(); insert (:1072,:1103)
        descriptors.assign(dbuf_.begin(), dbuf_.begin() + (size_t)n * 32);   // create(n,32,CV_8U) / release()
        pyramidValid_ = false;
    }

    int GetLevels() { return nlevels_; }
    float GetScaleFactor() { return scaleFactor_; }
    std::vector<float> GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // The reference exposes the pyramid of the last image as a public member
    // (ORBextractor.h:85, read by Frame::ComputeStereoMatches).  The device keeps it; it is copied
    // to the host only when somebody asks (callers that were not moved to the device path).
    // Each level carries the reference's 19-pixel BORDER_REFLECT_101 margin; view = interior.
    const std::vector<OwnedImage>& FetchPyramid()
    {
        if (!pyramidValid_) {
            for (int l = 0; l < nlevels_; ++l) {
                int w = 0, hh = 0;
                check(orbx_get_level(h_, 0, l, 0, 0, nullptr, 0, &w, &hh), "orbx_get_level");
                OwnedImage& m = mvImagePyramid[l];
                m.cols = w; m.rows = hh; m.step = (size_t)w + 38;
                m.buf.assign(m.step * (size_t)(hh + 38), 0);
                check(orbx_get_level(h_, 0, l, 0, 19, m.buf.data(), m.step, nullptr, nullptr), "orbx_get_level");
            }
            pyramidValid_ = true;
        }
        return mvImagePyramid;
    }
    std::vector<OwnedImage> mvImagePyramid;   // filled by FetchPyramid(); interior of level l starts at (19,19)

    orbx_handle* handle() { return h_; }
    int capacity() const { return cap_; }

#ifdef ORBX_WITH_OPENCV
    // the reference's exact signature (include/ORBextractor.h:59)
    void operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                    cv::OutputArray _descriptors)
    {
        (void)_mask;
        if (_image.empty()) return;
        cv::Mat image = _image.getMat();
        if (image.cols != shapeW_ || image.rows != shapeH_) {
            const int cap = orbx_reserve_shape(h_, image.cols, image.rows);
            check(cap < 0 ? cap : ORBX_OK, "ORBextractor::operator()");
            cap_ = cap; shapeW_ = image.cols; shapeH_ = image.rows;
        }
        kbuf_.resize(cap_); dbuf_.resize((size_t)cap_ * 32);
        int n = 0;
        check(orbx_extract(h_, image.data, image.cols, image.rows, (size_t)image.step,
                           reinterpret_cast<orbx_keypoint*>(kbuf_.data()), dbuf_.data(), cap_, &n), "ORBextractor::operator()");
        if (n == 0) _descriptors.release();
        else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), dbuf_.data() + (size_t)i * 32, 32);
        }
        _keypoints.clear();
        _keypoints.reserve(n);
        for (int i = 0; i < n; ++i) {
            const KeyPoint& k = kbuf_[i];
            _keypoints.push_back(cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
        }
        pyramidValid_ = false;
    }
#endif

protected:
    orbx_handle* h_ = nullptr;
    int nlevels_, cap_ = 0, shapeW_ = 0, shapeH_ = 0;
    float scaleFactor_;
    bool pyramidValid_ = false;
    std::vector<KeyPoint> kbuf_;
    std::vector<uint8_t> dbuf_;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
};

// ---------------------------------------------------------------------------------------------
// What the matchers need from a Frame (reference include/Frame.h:132-189): undistorted
// keypoints, descriptors, image bounds of the 64x48 grid, optional right coordinates.
// ---------------------------------------------------------------------------------------------
struct FrameView {
    const std::vector<KeyPoint>* mvKeysUn = nullptr;
    const uint8_t* mDescriptors = nullptr;         // N x 32
    float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    const std::vector<float>* mvuRight = nullptr;
    const std::vector<float>* mvScaleFactors = nullptr;

    orbm_frame c() const
    {
        orbm_frame f;
        f.n = mvKeysUn ? (int32_t)mvKeysUn->size() : 0;
        f.keys_un = mvKeysUn ? reinterpret_cast<const orbx_keypoint*>(mvKeysUn->data()) : nullptr;
        f.descriptors = mDescriptors;
        f.min_x = mnMinX; f.max_x = mnMaxX; f.min_y = mnMinY; f.max_y = mnMaxY;
        f.u_right = mvuRight ? mvuRight->data() : nullptr;
        f.scale_factors = mvScaleFactors ? mvScaleFactors->data() : nullptr;
        f.nlevels = mvScaleFactors ? (int32_t)mvScaleFactors->size() : 0;
        return f;
    }
};

// Map points already projected by the caller (the projection itself is scalar host code behind
// MapPoint mutexes in the reference, src/ORBmatcher.cc:51-76,1340-1376): one query per point.
struct ProjectionQueries {
    std::vector<orbm_proj_query> q;
    std::vector<uint8_t> descriptors;    // q.size() x 32  (MapPoint::GetDescriptor)
    std::vector<uint8_t> blocks;         // pMP->Observations() > 0
};

// The reference's own expressions that turn one projected MapPoint into a query (what the maintainer's loop computes before it
// appends to ProjectionQueries::q), float for float.  scale = F.mvScaleFactors.
struct ProjectionFlatten {
    // src/ORBmatcher.cc:61-77 — r = RadiusByViewingCos(viewCos) [* th when th != 1]; window r * scale[level]; levels [level-1, level];
    // stereo gate |mTrackProjXR - mvuRight| <= the same window.  The cosine is compared with the DOUBLE literal 0.998 (:131-137).
    static orbm_proj_query LocalMap(float projX, float projY, float projXR, int level, float viewCos, float th, const std::vector<float>& scale)
    {
        float r = (double)viewCos > 0.998 ? 2.5f : 4.0f;
        if (th != 1.0f) r *= th;
        orbm_proj_query q = blank();
        q.u = projX; q.v = projY; q.radius = r * scale[level];
        q.min_level = level - 1; q.max_level = level;
        q.u_right = projXR; q.max_right_err = q.radius;
        return q;
    }
    // src/ORBmatcher.cc:1374-1410 — window th * scale[octave of the last frame's keypoint]; levels by the motion direction
    // (forward: [octave, any], backward: [0, octave], else [octave-1, octave+1]); ur = u - mbf * invzc, gate = the window.
    static orbm_proj_query LastFrame(float u, float v, float invzc, int lastOctave, float angle, float th, float mbf, bool forward,
                                     bool backward, const std::vector<float>& scale)
    {
        orbm_proj_query q = blank();
        q.u = u; q.v = v; q.radius = th * scale[lastOctave];
        if (forward) { q.min_level = lastOctave; q.max_level = -1; }
        else if (backward) { q.min_level = 0; q.max_level = lastOctave; }
        else { q.min_level = lastOctave - 1; q.max_level = lastOctave + 1; }
        q.u_right = u - mbf * invzc; q.max_right_err = q.radius;
        q.angle = angle;
        return q;
    }
    // src/ORBmatcher.cc:1524-1529 — window th * scale[nPredictedLevel], levels [level-1, level+1]
    static orbm_proj_query KeyFrame(float u, float v, int level, float angle, float th, const std::vector<float>& scale)
    {
        orbm_proj_query q = blank();
        q.u = u; q.v = v; q.radius = th * scale[level];
        q.min_level = level - 1; q.max_level = level + 1;
        q.angle = angle;
        return q;
    }
    // src/ORBmatcher.cc:357-377, :884-905, :1041-1062, :1196-1215 — KeyFrame::GetFeaturesInArea has no level window; the octave filter
    // kpLevel in [level-1, level] is applied per candidate.  A NEGATIVE level_lo switches that filter off in the flat form, so a
    // predicted level 0 is passed as [0, 0].  u_right = u - bf * invz is read by Fuse's chi-square gate only.
    static orbm_proj_query Sim3OrFuse(float u, float v, float invz, int level, float th, float bf, const std::vector<float>& scale)
    {
        orbm_proj_query q = blank();
        q.u = u; q.v = v; q.radius = th * scale[level];
        q.level_lo = level > 0 ? level - 1 : 0; q.level_hi = level;
        q.u_right = u - bf * invz;
        return q;
    }
private:
    static orbm_proj_query blank()
    {
        orbm_proj_query q;
        std::memset(&q, 0, sizeof(q));
        q.min_level = q.max_level = q.level_lo = q.level_hi = -1;
        q.max_right_err = -1.f;
        return q;
    }
};

struct ProjectionResult {
    int nmatches = 0;
    std::vector<int32_t> assigned;       // per keypoint: query that owns it at the end, or -1
    std::vector<int32_t> bestIdx, bestDist;   // per query
};

// What the keyframe-rate matchers read from a KeyFrame (reference include/KeyFrame.h): the Frame data plus the level
// sigma tables, the FeatureVector node of every feature and two per-feature flags.
struct KeyFrameView {
    FrameView frame;
    const std::vector<float>* mvLevelSigma2 = nullptr;
    const std::vector<float>* mvInvLevelSigma2 = nullptr;
    const std::vector<uint32_t>* featNode = nullptr;     // node id of mFeatVec that holds feature i
    const std::vector<uint8_t>* flags = nullptr;         // bit 0: in mFeatVec, bit 1: GetMapPoint(i) != NULL (SearchByBoW: and not bad)

    orbm_keyframe c() const
    {
        orbm_keyframe k;
        k.frame = frame.c();
        k.level_sigma2 = mvLevelSigma2 ? mvLevelSigma2->data() : nullptr;
        k.inv_level_sigma2 = mvInvLevelSigma2 ? mvInvLevelSigma2->data() : nullptr;
        k.node_id = featNode ? featNode->data() : nullptr;
        k.flags = flags ? flags->data() : nullptr;
        return k;
    }
};

// ORBVocabulary (DBoW2 TemplatedVocabulary) flattened once and kept on the device.
class Vocabulary {
public:
    Vocabulary(const orbm_vocabulary& flat, int device = 0) { check(orbm_vocabulary_upload(&flat, device, &h_), "Vocabulary"); }
    ~Vocabulary() { if (h_) orbm_vocabulary_free(h_); }
    Vocabulary(const Vocabulary&) = delete;
    Vocabulary& operator=(const Vocabulary&) = delete;
    // what Frame::ComputeBoW / KeyFrame::ComputeBoW evaluate per feature (levelsup = 4): the caller accumulates
    // mBowVec[word] += weight and mFeatVec[node].push_back(i) for the features with weight > 0
    void transform(const uint8_t* descriptors, int n, int levelsup, std::vector<uint32_t>& word, std::vector<double>& weight,
                   std::vector<uint32_t>& node)
    {
        word.assign((size_t)n, 0); weight.assign((size_t)n, 0.0); node.assign((size_t)n, 0);
        check(orbm_bow_transform(h_, descriptors, n, levelsup, word.data(), weight.data(), node.data()), "Vocabulary::transform");
    }
private:
    orbm_vocab_handle* h_ = nullptr;
};

struct BowSide {                       // one frame of SearchByBoW
    int n = 0;
    const uint8_t* descriptors = nullptr;
    const float* angles = nullptr;
    const uint32_t* node = nullptr;
    const uint8_t* flags = nullptr;
    orbm_bow_frame c() const { orbm_bow_frame f; f.n = n; f.descriptors = descriptors; f.angles = angles; f.node_id = node; f.flags = flags; return f; }
};

// ---------------------------------------------------------------------------------------------
// ORBmatcher (reference include/ORBmatcher.h:36)
// ---------------------------------------------------------------------------------------------
class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6f, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:37-39

    // Hamming distance of two 32-byte descriptors (src/ORBmatcher.cc:1647).  A single pair is host
    // work (8 popcounts); batches go through DescriptorDistances.
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b)
    {
        int d = 0;
        for (int i = 0; i < 8; ++i) {
            uint32_t x, y;
            std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4);
            d += __builtin_popcount(x ^ y);
        }
        return d;
    }
    static void DescriptorDistances(const uint8_t* a, const uint8_t* b, int count, int32_t* dist)
    {
        check(orbm_descriptor_distance(a, b, count, dist), "DescriptorDistance");
    }

    // src/ORBmatcher.cc:405 — same arguments, same in/out behaviour of vbPrevMatched / vnMatches12
    int SearchForInitialization(const FrameView& F1, const FrameView& F2, std::vector<Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10)
    {
        const orbm_frame f1 = F1.c(), f2 = F2.c();
        vnMatches12.assign((size_t)f1.n, -1);
        if ((int)vbPrevMatched.size() < f1.n) throw std::runtime_error("vbPrevMatched smaller than F1");
        int32_t n = 0;
        check(orbm_search_for_initialization(&f1, &f2, reinterpret_cast<float*>(vbPrevMatched.data()), vnMatches12.data(),
                                             windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, &n), "SearchForInitialization");
        return n;
    }

    // The four SearchByProjection overloads (src/ORBmatcher.cc:45,290,1328,1472) on flattened queries.
    // occupied: in/out, one byte per keypoint of F (it already holds a map point that blocks).
    ProjectionResult SearchByProjection(int variant, const FrameView& F, const ProjectionQueries& Q,
                                        std::vector<uint8_t>& occupied, int thDist)
    {
        const orbm_frame f = F.c();
        ProjectionResult r;
        r.assigned.assign((size_t)f.n, -1);
        r.bestIdx.assign(Q.q.size(), -1); r.bestDist.assign(Q.q.size(), 256);
        occupied.resize((size_t)f.n, 0);
        int32_t n = 0;
        check(orbm_search_by_projection(variant, &f, Q.q.data(), Q.descriptors.data(), (int)Q.q.size(),
                                        Q.blocks.empty() ? nullptr : Q.blocks.data(), occupied.data(), r.assigned.data(),
                                        r.bestIdx.data(), r.bestDist.data(), thDist, mfNNratio, mbCheckOrientation ? 1 : 0, &n),
              "SearchByProjection");
        r.nmatches = n;
        return r;
    }

    // SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:159, variant 0) and
    // SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (:522, variant 1).  matchA[i] / matchB[j]: partner feature or -1.
    int SearchByBoW(int variant, const BowSide& A, const BowSide& B, std::vector<int>& matchA, std::vector<int>& matchB, int device = 0)
    {
        const orbm_bow_frame a = A.c(), b = B.c();
        matchA.assign((size_t)A.n, -1); matchB.assign((size_t)B.n, -1);
        int32_t n = 0;
        check(orbm_search_by_bow(variant, &a, &b, 1, mfNNratio, mbCheckOrientation ? 1 : 0, device, matchA.data(), matchB.data(), &n), "SearchByBoW");
        return n;
    }

    // SearchForTriangulation (src/ORBmatcher.cc:657): F12 row-major, (ex, ey) the epipole of :664-670
    int SearchForTriangulation(const KeyFrameView& KF1, const KeyFrameView& KF2, const float F12[9], float ex, float ey,
                               std::vector<std::pair<size_t, size_t>>& vMatchedPairs, bool bOnlyStereo, int device = 0)
    {
        const orbm_keyframe k1 = KF1.c(), k2 = KF2.c();
        std::vector<int32_t> m((size_t)std::max(k1.frame.n, 1), -1);
        const float epi[2] = {ex, ey};
        int32_t n = 0;
        check(orbm_search_for_triangulation(&k1, &k2, 1, F12, epi, bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0, device, m.data(), &n),
              "SearchForTriangulation");
        vMatchedPairs.clear();
        vMatchedPairs.reserve((size_t)n);
        for (int i = 0; i < k1.frame.n; ++i) if (m[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m[i]));
        return n;
    }

    // the per-MapPoint search of Fuse (variant ORBM_FUSE_MAPPOINTS / ORBM_FUSE_SIM3); the caller replays the side effects
    std::vector<int32_t> Fuse(int variant, const KeyFrameView& KF, const ProjectionQueries& Q, int device = 0)
    {
        const orbm_keyframe k = KF.c();
        std::vector<int32_t> best(Q.q.size(), -1);
        check(orbm_fuse(variant, &k, Q.q.data(), Q.descriptors.data(), (int)Q.q.size(), device, best.data(), nullptr), "Fuse");
        return best;
    }

    // SearchBySim3 (src/ORBmatcher.cc:1102): Q12 = MapPoints of KF1 (features index1) projected into KF2, Q21 the reverse
    int SearchBySim3(const KeyFrameView& KF1, const KeyFrameView& KF2, const ProjectionQueries& Q12, const std::vector<int32_t>& index1,
                     const ProjectionQueries& Q21, const std::vector<int32_t>& index2, std::vector<int>& match12, int device = 0)
    {
        const orbm_keyframe k1 = KF1.c(), k2 = KF2.c();
        match12.assign((size_t)std::max(k1.frame.n, 1), -1);
        int32_t n = 0;
        check(orbm_search_by_sim3(&k1, &k2, Q12.q.data(), Q12.descriptors.data(), index1.data(), (int)Q12.q.size(), Q21.q.data(),
                                  Q21.descriptors.data(), index2.data(), (int)Q21.q.size(), device, match12.data(), &n), "SearchBySim3");
        match12.resize((size_t)k1.frame.n);
        return n;
    }

protected:
    float mfNNratio;
    bool mbCheckOrientation;
};

// Frame::ComputeStereoMatches (src/Frame.cc:466) for callers that hold both eyes on the host.
// (The fused path — extraction of both eyes + matching without the pyramids leaving the device —
// is orbx_extract_stereo_batch.)
inline void ComputeStereoMatches(const std::vector<KeyPoint>& keysLeft, const uint8_t* descLeft,
                                 const std::vector<KeyPoint>& keysRight, const uint8_t* descRight,
                                 const std::vector<Image>& pyrLeft, const std::vector<Image>& pyrRight,
                                 const std::vector<float>& scaleFactors, const std::vector<float>& invScaleFactors,
                                 float mbf, float fx, std::vector<float>& mvuRight, std::vector<float>& mvDepth)
{
    const int nl = (int)pyrLeft.size();
    std::vector<const uint8_t*> pl(nl), pr(nl);
    std::vector<int32_t> lw(nl), lh(nl);
    std::vector<int64_t> sl(nl), sr(nl);
    for (int l = 0; l < nl; ++l) {
        pl[l] = pyrLeft[l].data; pr[l] = pyrRight[l].data;
        lw[l] = pyrLeft[l].cols; lh[l] = pyrLeft[l].rows;
        sl[l] = (int64_t)pyrLeft[l].step; sr[l] = (int64_t)pyrRight[l].step;
    }
    orbm_stereo_input in;
    in.n_left = (int32_t)keysLeft.size(); in.n_right = (int32_t)keysRight.size();
    in.keys_left = reinterpret_cast<const orbx_keypoint*>(keysLeft.data());
    in.keys_right = reinterpret_cast<const orbx_keypoint*>(keysRight.data());
    in.desc_left = descLeft; in.desc_right = descRight;
    in.nlevels = nl;
    in.scale_factors = scaleFactors.data(); in.inv_scale_factors = invScaleFactors.data();
    in.pyr_left = pl.data(); in.pyr_right = pr.data();
    in.level_width = lw.data(); in.level_height = lh.data();
    in.pyr_stride_left = sl.data(); in.pyr_stride_right = sr.data();
    in.bf = mbf; in.fx = fx;
    mvuRight.assign(keysLeft.size(), -1.0f);
    mvDepth.assign(keysLeft.size(), -1.0f);
    check(orbm_compute_stereo_matches(&in, mvuRight.data(), mvDepth.data(), nullptr), "ComputeStereoMatches");
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242) for a batch of map points: descriptors of point p are
// rows offsets[p] .. offsets[p+1]; returns the index (inside its group) of each point's distinctive descriptor.
inline std::vector<int32_t> ComputeDistinctiveDescriptors(const uint8_t* descriptors, const std::vector<int32_t>& offsets, int device = 0)
{
    const int np = (int)offsets.size() - 1;
    std::vector<int32_t> best((size_t)std::max(np, 0), -1);
    if (np > 0) check(orbm_distinctive_descriptors(descriptors, offsets.data(), np, device, best.data(), nullptr), "ComputeDistinctiveDescriptors");
    return best;
}

// PointCloudMapping::generatePointCloud (src/PointCloudMap.cc:78) for one keyframe: depth CV_32F, bgr 8UC3 (or NULL),
// poseInv = toSE3Quat(kf->GetPose()).inverse().matrix() row-major (or NULL: convertToPointCloud)
inline std::vector<orbx_cloud_point> GeneratePointCloud(const float* depth, size_t depthStepFloats, const uint8_t* bgr, size_t bgrStep,
                                                        int cols, int rows, float fx, float fy, float cx, float cy,
                                                        const double* poseInv, int device = 0)
{
    orbx_camera cam{};
    cam.fx = fx; cam.fy = fy; cam.cx = cx; cam.cy = cy;
    const int cap = ((cols + 2) / 3) * ((rows + 2) / 3);
    std::vector<orbx_cloud_point> pts((size_t)cap);
    int32_t n = 0;
    check(orbm_generate_point_clouds(depth, depthStepFloats, depthStepFloats * rows, bgr, bgrStep, bgrStep * rows, cols, rows, 1, 3, &cam,
                                     poseInv, device, pts.data(), cap, &n), "generatePointCloud");
    pts.resize((size_t)n);
    return pts;
}

// voxel.setInputCloud(cloud); voxel.filter(*tmp) with voxel.setLeafSize(resolution, resolution, resolution)
// (src/PointCloudMap.cc:41, 240-241, 259-260, 277-280)
inline std::vector<orbx_cloud_point> VoxelGridFilter(const std::vector<orbx_cloud_point>& cloud, float resolution, int device = 0)
{
    std::vector<orbx_cloud_point> out(cloud.size());
    int32_t n = 0;
    if (!cloud.empty()) check(orbm_voxel_grid(cloud.data(), (int)cloud.size(), resolution, device, out.data(), (int)out.size(), &n), "VoxelGrid::filter");
    out.resize((size_t)n);
    return out;
}

// sor.setMeanK(meanK); sor.setStddevMulThresh(mul); sor.setInputCloud(globalMap); sor.filter(*tmp)
// (src/PointCloudMap.cc:46-47, 284-285)
inline std::vector<orbx_cloud_point> StatisticalOutlierFilter(const std::vector<orbx_cloud_point>& cloud, int meanK = 50, double mul = 1.0, int device = 0)
{
    std::vector<orbx_cloud_point> out(cloud.size());
    int64_t n = 0;
    if (!cloud.empty()) check(orbm_statistical_outlier_removal(cloud.data(), (int64_t)cloud.size(), meanK, mul, device, out.data(), (int64_t)out.size(),
                                                               &n, nullptr, nullptr), "StatisticalOutlierRemoval::filter");
    out.resize((size_t)n);
    return out;
}

} // namespace orbx
#endif
