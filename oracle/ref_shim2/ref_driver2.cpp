/* ref_driver2.cpp — drives the UNMODIFIED reference matcher code (src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc,
 * src/MapPoint.cc, src/Map.cc, src/KeyFrameDatabase.cc, src/ORBextractor.cc and the vendored DBoW2) from flat arrays,
 * so that tests can assert  oracle == reference  for every matcher row of SURVEY.md §8.
 *
 * TEST INFRASTRUCTURE ONLY (oracle/_ref/libref_matchers.so; built by oracle/Makefile target `ref2` when
 * /root/reference is mounted).  Nothing here is reference code: this file only BUILDS the reference's own objects
 * (Frame, KeyFrame, MapPoint, ORBVocabulary) from arrays and calls the reference's own member functions.  It is
 * compiled with -fno-access-control so that private / protected members (Frame::AssignFeaturesToGrid, MapPoint::nObs,
 * mfMaxDistance ...) can be set directly instead of replaying Tracking / LocalMapping to reach the same state.
 *
 * Geometry: the functions that project map points themselves (SearchByProjection from a Frame / KeyFrame / Sim3,
 * Fuse, SearchBySim3) are fed identity poses, fx = fy = 1, cx = cy = 0 and world points (u, v, 1) unless a test says
 * otherwise: every product and sum of the projection is then exact, the point lands on (u, v) bit for bit, and the
 * comparison does not depend on how OpenCV accumulates small float matrix products (cvshim2.h).
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <set>
#include <vector>

#include "cvshim2.h"

#include "Converter.h"
#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"

using namespace ORB_SLAM2;

/* src/Converter.cc is not compiled (it needs Eigen and g2o); this one member is all the matcher path uses.
 * Same behaviour as src/Converter.cc:27-35: one header per descriptor row. */
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors)
{
    std::vector<cv::Mat> v;
    v.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; ++j) v.push_back(Descriptors.row(j));
    return v;
}

/* Map::Save (src/Map.cc:157) names this Eigen-based helper; Map saving is not on the pinned paths */
std::vector<float> Converter::toQuaternion(const cv::Mat&) { cv::shim_abort("Converter::toQuaternion (needs Eigen)"); }

namespace {

std::mutex g_lock;      /* Frame keeps calibration, bounds and grid pitch in static members */

struct Cam { float fx, fy, cx, cy, bf; };
const Cam kUnitCam = {1.f, 1.f, 0.f, 0.f, 0.f};

cv::Mat eye4() { return cv::Mat::eye(4, 4, CV_32F); }

cv::Mat desc_mat(const uint8_t* d, int n)
{
    cv::Mat m(std::max(n, 0), 32, CV_8U);
    if (n > 0) std::memcpy(m.data, d, (size_t)n * 32);
    return m;
}

cv::KeyPoint to_kp(const orbx_keypoint& k) { return cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id); }

/* A Frame in the state its constructor leaves it in (src/Frame.cc:136-189 minus extraction): keypoints, descriptors,
 * scale tables, static calibration / bounds / grid pitch, grid. */
Frame* make_frame(const orbm_frame* f, const Cam& cam, const cv::Mat& Tcw = cv::Mat())
{
    Frame* F = new Frame();
    F->mpORBvocabulary = nullptr; F->mpORBextractorLeft = F->mpORBextractorRight = nullptr;
    F->mTimeStamp = 0; F->mnId = Frame::nNextId++; F->mpReferenceKF = nullptr;
    F->N = f->n;
    F->mvKeys.resize(f->n); F->mvKeysUn.resize(f->n);
    for (int i = 0; i < f->n; ++i) F->mvKeys[i] = F->mvKeysUn[i] = to_kp(f->keys_un[i]);
    F->mDescriptors = desc_mat(f->descriptors, f->n);
    F->mvuRight.assign(f->n, -1.f); F->mvDepth.assign(f->n, -1.f);
    if (f->u_right) for (int i = 0; i < f->n; ++i) F->mvuRight[i] = f->u_right[i];
    F->mvpMapPoints.assign(f->n, static_cast<MapPoint*>(nullptr));
    F->mvbOutlier.assign(f->n, false);
    const int nl = f->scale_factors ? f->nlevels : 8;
    F->mnScaleLevels = nl;
    F->mvScaleFactors.resize(nl); F->mvInvScaleFactors.resize(nl); F->mvLevelSigma2.resize(nl); F->mvInvLevelSigma2.resize(nl);
    for (int l = 0; l < nl; ++l) {
        float s = 1.f;
        if (f->scale_factors) s = f->scale_factors[l];
        else for (int k = 0; k < l; ++k) s = (float)(s * (double)1.2f);
        F->mvScaleFactors[l] = s;
        F->mvLevelSigma2[l] = s * s;                         /* src/ORBextractor.cc:419-431 */
        F->mvInvScaleFactors[l] = 1.0f / s;
        F->mvInvLevelSigma2[l] = 1.0f / F->mvLevelSigma2[l];
    }
    F->mfScaleFactor = nl > 1 ? F->mvScaleFactors[1] : 1.2f;
    F->mfLogScaleFactor = std::log(F->mfScaleFactor);       /* src/Frame.cc:150 */
    F->mbf = cam.bf; F->mThDepth = 0.f;
    F->mK = cv::Mat::eye(3, 3, CV_32F);
    F->mK.at<float>(0, 0) = cam.fx; F->mK.at<float>(1, 1) = cam.fy; F->mK.at<float>(0, 2) = cam.cx; F->mK.at<float>(1, 2) = cam.cy;
    F->mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    /* the statics of src/Frame.cc:170-186 */
    Frame::mnMinX = f->min_x; Frame::mnMaxX = f->max_x; Frame::mnMinY = f->min_y; Frame::mnMaxY = f->max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::fx = cam.fx; Frame::fy = cam.fy; Frame::cx = cam.cx; Frame::cy = cam.cy;
    Frame::invfx = 1.0f / cam.fx; Frame::invfy = 1.0f / cam.fy;
    Frame::mbInitialComputations = false;
    F->mb = F->mbf / Frame::fx;
    F->AssignFeaturesToGrid();
    if (!Tcw.empty()) F->SetPose(Tcw);
    return F;
}

struct World {
    Map map;
    KeyFrameDatabase* db = nullptr;
    std::vector<Frame*> frames;
    std::vector<KeyFrame*> kfs;
    std::vector<MapPoint*> mps;
    KeyFrame* anchor = nullptr;                 /* reference keyframe of every MapPoint built here */
    ~World()
    {
        for (MapPoint* p : mps) delete p;
        for (KeyFrame* k : kfs) delete k;
        for (Frame* f : frames) delete f;
    }
    KeyFrame* make_keyframe(const orbm_frame* f, const Cam& cam, const cv::Mat& Tcw)
    {
        Frame* F = make_frame(f, cam, Tcw.empty() ? eye4() : Tcw);
        frames.push_back(F);
        KeyFrame* kf = new KeyFrame(*F, &map, db);
        kfs.push_back(kf);
        return kf;
    }
    /* a MapPoint as LocalMapping would have left it: position, descriptor, normal, scale-invariance range, observation
     * count (src/MapPoint.cc:34-47 ctor, then the fields ComputeDistinctiveDescriptors / UpdateNormalAndDepth maintain) */
    MapPoint* make_point(float x, float y, float z, const uint8_t* desc, int nobs, float maxDist = 1e30f, float minDist = 0.f)
    {
        if (!anchor) {
            orbm_frame e; std::memset(&e, 0, sizeof(e));
            e.min_x = 0; e.max_x = 640; e.min_y = 0; e.max_y = 480;
            const float savedB[4] = {Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY};
            const float savedG[2] = {Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv};
            anchor = make_keyframe(&e, kUnitCam, eye4());
            Frame::mnMinX = savedB[0]; Frame::mnMaxX = savedB[1]; Frame::mnMinY = savedB[2]; Frame::mnMaxY = savedB[3];
            Frame::mfGridElementWidthInv = savedG[0]; Frame::mfGridElementHeightInv = savedG[1];
        }
        cv::Mat pos = (cv::Mat_<float>(3, 1) << x, y, z);
        MapPoint* p = new MapPoint(pos, anchor, &map);
        if (desc) p->mDescriptor = desc_mat(desc, 1);
        p->nObs = nobs;
        const float n = std::sqrt(x * x + y * y + z * z);
        p->mNormalVector = (cv::Mat_<float>(3, 1) << x / n, y / n, z / n);     /* seen head-on from the origin */
        p->mfMaxDistance = maxDist; p->mfMinDistance = minDist;
        mps.push_back(p);
        return p;
    }
};

/* index of a MapPoint* inside a query list, -2 for a pre-existing point, -1 for NULL */
void report_assignment(const std::vector<MapPoint*>& slots, const std::vector<MapPoint*>& queries, int32_t* assigned)
{
    for (size_t i = 0; i < slots.size(); ++i) {
        int32_t a = -1;
        if (slots[i]) {
            a = -2;
            for (size_t q = 0; q < queries.size(); ++q) if (queries[q] == slots[i]) { a = (int32_t)q; break; }
        }
        assigned[i] = a;
    }
}

} // namespace

extern "C" {

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) */
int ref2_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    cv::Mat ma = desc_mat(a, 1), mb = desc_mat(b, 1);
    return ORBmatcher::DescriptorDistance(ma, mb);
}

/* Frame::GetFeaturesInArea (src/Frame.cc:327-380) on a frame gridded by Frame::AssignFeaturesToGrid (:230-245) */
int ref2_features_in_area(const orbm_frame* f, float x, float y, float r, int min_level, int max_level, int32_t* out, int cap)
{
    std::lock_guard<std::mutex> g(g_lock);
    Frame* F = make_frame(f, kUnitCam);
    const std::vector<size_t> v = F->GetFeaturesInArea(x, y, r, min_level, max_level);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int32_t)v[i];
    delete F;
    return (int)v.size();
}

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) */
int ref2_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_matched, int32_t* matches12,
                                   int window_size, float nnratio, int check_orientation)
{
    std::lock_guard<std::mutex> g(g_lock);
    Frame* F1 = make_frame(f1, kUnitCam);
    Frame* F2 = make_frame(f2, kUnitCam);       /* same bounds for both, as in the reference (statics) */
    std::vector<cv::Point2f> prev(f1->n);
    for (int i = 0; i < f1->n; ++i) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchForInitialization(*F1, *F2, prev, m12, window_size);
    for (int i = 0; i < f1->n; ++i) { matches12[i] = m12[i]; prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y; }
    delete F1; delete F2;
    return n;
}

/* Frame::ComputeStereoMatches (src/Frame.cc:466-638): two ORBextractor objects whose mvImagePyramid members view the
 * caller's levels (the reference reads them at :471, :561-581), a Frame holding both eyes' keypoints. */
int ref2_compute_stereo_matches(const orbm_stereo_input* in, float* u_right, float* depth)
{
    std::lock_guard<std::mutex> g(g_lock);
    const int nl = in->nlevels;
    ORBextractor exL(1000, nl > 1 ? in->scale_factors[1] : 1.2f, nl, 20, 7), exR(1000, nl > 1 ? in->scale_factors[1] : 1.2f, nl, 20, 7);
    for (int l = 0; l < nl; ++l) {
        exL.mvImagePyramid[l] = cv::Mat(in->level_height[l], in->level_width[l], CV_8U, const_cast<uint8_t*>(in->pyr_left[l]), (size_t)in->pyr_stride_left[l]);
        exR.mvImagePyramid[l] = cv::Mat(in->level_height[l], in->level_width[l], CV_8U, const_cast<uint8_t*>(in->pyr_right[l]), (size_t)in->pyr_stride_right[l]);
    }
    orbm_frame fl; std::memset(&fl, 0, sizeof(fl));
    fl.n = in->n_left; fl.keys_un = in->keys_left; fl.descriptors = in->desc_left;
    fl.min_x = 0; fl.max_x = (float)in->level_width[0]; fl.min_y = 0; fl.max_y = (float)in->level_height[0];
    fl.scale_factors = in->scale_factors; fl.nlevels = nl;
    const Cam cam = {in->fx, in->fx, 0.f, 0.f, in->bf};
    Frame* F = make_frame(&fl, cam);
    F->mpORBextractorLeft = &exL; F->mpORBextractorRight = &exR;
    /* the reference takes the inverse scale factors from the extractor (src/Frame.cc:97); they equal 1/scale */
    F->mvInvScaleFactors = exL.GetInverseScaleFactors();
    for (int l = 0; l < nl; ++l) if (F->mvInvScaleFactors[l] != in->inv_scale_factors[l] || exL.GetScaleFactors()[l] != in->scale_factors[l]) { delete F; return -1; }
    F->mvKeysRight.resize(in->n_right);
    for (int i = 0; i < in->n_right; ++i) F->mvKeysRight[i] = to_kp(in->keys_right[i]);
    F->mDescriptorsRight = desc_mat(in->desc_right, in->n_right);
    F->ComputeStereoMatches();
    for (int i = 0; i < in->n_left; ++i) { u_right[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i]; }
    delete F;
    return 0;
}

/* ---- SearchByProjection x4 -------------------------------------------------------------------------------------
 * slot_state[n]: 0 = free, 1 = holds a MapPoint with Observations() > 0, 2 = holds a MapPoint with no observation.
 * assigned[n] (out): index of the query that owns keypoint i at the end, -2 = the pre-existing point, -1 = none. */

/* (Frame&, const vector<MapPoint*>&, th)  src/ORBmatcher.cc:45-129.  Per map point the fields Frame::isInFrustum
 * leaves behind (src/Frame.cc:272-325): mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos, mbTrackInView. */
int ref2_proj_localmap(const orbm_frame* f, const uint8_t* slot_state, int nq, const float* proj_x, const float* proj_y,
                       const float* proj_xr, const int32_t* level, const float* view_cos, const uint8_t* in_view,
                       const uint8_t* qdesc, const int32_t* nobs, float th, float nnratio, int check_orientation, int32_t* assigned)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    Frame* F = make_frame(f, kUnitCam);
    w.frames.push_back(F);
    static const uint8_t zero[32] = {0};
    for (int i = 0; i < f->n; ++i)
        if (slot_state[i]) F->mvpMapPoints[i] = w.make_point(0, 0, 1, zero, slot_state[i] == 1 ? 3 : 0);
    std::vector<MapPoint*> q(nq);
    for (int i = 0; i < nq; ++i) {
        MapPoint* p = w.make_point(proj_x[i], proj_y[i], 1.f, qdesc + (size_t)i * 32, nobs[i]);
        p->mbTrackInView = in_view[i] != 0;
        p->mTrackProjX = proj_x[i]; p->mTrackProjY = proj_y[i]; p->mTrackProjXR = proj_xr[i];
        p->mnTrackScaleLevel = level[i]; p->mTrackViewCos = view_cos[i];
        q[i] = p;
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchByProjection(*F, q, th);
    report_assignment(F->mvpMapPoints, q, assigned);
    return n;
}

/* (Frame& CurrentFrame, const Frame& LastFrame, th, bMono)  src/ORBmatcher.cc:1328-1470.
 * Query i = keypoint i of the last frame (octave[i], angle[i]) carrying a map point that projects to (u[i], v[i]) at
 * depth 1 in the current frame (identity pose, unit camera).  last_tz: z of the last frame's translation: with the
 * current pose at the origin tlc = tlw, so last_tz > mb selects the forward window, -last_tz > mb the backward one. */
int ref2_proj_lastframe(const orbm_frame* cur, const uint8_t* slot_state, float bf, int nq, const float* u, const float* v,
                        const int32_t* octave, const float* angle, const uint8_t* qdesc, const int32_t* nobs, const uint8_t* outlier,
                        float last_tz, float th, int mono, float nnratio, int check_orientation, int32_t* assigned)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    const Cam cam = {1.f, 1.f, 0.f, 0.f, bf};
    std::vector<orbx_keypoint> lk(nq);
    for (int i = 0; i < nq; ++i) { std::memset(&lk[i], 0, sizeof(orbx_keypoint)); lk[i].x = 100.f; lk[i].y = 100.f; lk[i].size = 31.f; lk[i].angle = angle[i]; lk[i].octave = octave[i]; lk[i].class_id = -1; }
    orbm_frame lf = *cur; lf.n = nq; lf.keys_un = lk.data(); lf.descriptors = qdesc; lf.u_right = nullptr;
    cv::Mat Tlw = eye4(); Tlw.at<float>(2, 3) = last_tz;
    Frame* L = make_frame(&lf, cam, Tlw);
    w.frames.push_back(L);
    Frame* F = make_frame(cur, cam, eye4());
    w.frames.push_back(F);
    static const uint8_t zero[32] = {0};
    for (int i = 0; i < cur->n; ++i)
        if (slot_state[i]) F->mvpMapPoints[i] = w.make_point(0, 0, 1, zero, slot_state[i] == 1 ? 3 : 0);
    std::vector<MapPoint*> q(nq);
    for (int i = 0; i < nq; ++i) {
        q[i] = w.make_point(u[i], v[i], 1.f, qdesc + (size_t)i * 32, nobs[i]);
        L->mvpMapPoints[i] = q[i];
        L->mvbOutlier[i] = outlier && outlier[i];
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchByProjection(*F, *L, th, mono != 0);
    report_assignment(F->mvpMapPoints, q, assigned);
    return n;
}

/* (Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist)  src/ORBmatcher.cc:1472-1599.
 * Query i = keypoint i of the keyframe (angle[i]) carrying a map point at (u, v, 1) whose mfMaxDistance makes
 * MapPoint::PredictScale (src/MapPoint.cc:385-394) return level[i]: max_dist[i] (the test computes it). */
int ref2_proj_keyframe(const orbm_frame* cur, const uint8_t* slot_state, int nq, const float* u, const float* v, const float* max_dist,
                       const float* angle, const uint8_t* qdesc, const uint8_t* already_found, float th, int orb_dist, float nnratio,
                       int check_orientation, int32_t* assigned, int32_t* predicted_level)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    std::vector<orbx_keypoint> kk(nq);
    for (int i = 0; i < nq; ++i) { std::memset(&kk[i], 0, sizeof(orbx_keypoint)); kk[i].x = 100.f; kk[i].y = 100.f; kk[i].size = 31.f; kk[i].angle = angle[i]; kk[i].class_id = -1; }
    orbm_frame kf = *cur; kf.n = nq; kf.keys_un = kk.data(); kf.descriptors = qdesc; kf.u_right = nullptr;
    KeyFrame* K = w.make_keyframe(&kf, kUnitCam, eye4());
    Frame* F = make_frame(cur, kUnitCam, eye4());
    w.frames.push_back(F);
    static const uint8_t zero[32] = {0};
    for (int i = 0; i < cur->n; ++i)
        if (slot_state[i]) F->mvpMapPoints[i] = w.make_point(0, 0, 1, zero, 1);
    std::vector<MapPoint*> q(nq);
    std::set<MapPoint*> found;
    for (int i = 0; i < nq; ++i) {
        q[i] = w.make_point(u[i], v[i], 1.f, qdesc + (size_t)i * 32, 1, max_dist[i], 0.f);
        K->mvpMapPoints[i] = q[i];
        if (already_found && already_found[i]) found.insert(q[i]);
        if (predicted_level) {
            const float d = (float)std::sqrt((double)u[i] * u[i] + (double)v[i] * v[i] + 1.0);
            predicted_level[i] = q[i]->PredictScale(d, F->mfLogScaleFactor);
        }
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchByProjection(*F, K, found, th, orb_dist);
    report_assignment(F->mvpMapPoints, q, assigned);
    return n;
}

/* (KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, th)  src/ORBmatcher.cc:290-403.
 * matched_state[n]: 1 = vpMatched[i] already holds some other point.  Scw = identity (scale 1). */
int ref2_proj_sim3(const orbm_frame* kf, const uint8_t* matched_state, int nq, const float* u, const float* v, const float* max_dist,
                   const uint8_t* qdesc, int th, int32_t* assigned, int32_t* predicted_level)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    KeyFrame* K = w.make_keyframe(kf, kUnitCam, eye4());
    static const uint8_t zero[32] = {0};
    std::vector<MapPoint*> matched(kf->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < kf->n; ++i) if (matched_state[i]) matched[i] = w.make_point(0, 0, 1, zero, 1);
    std::vector<MapPoint*> q(nq);
    for (int i = 0; i < nq; ++i) {
        q[i] = w.make_point(u[i], v[i], 1.f, qdesc + (size_t)i * 32, 1, max_dist[i], 0.f);
        if (predicted_level) {
            const float d = (float)std::sqrt((double)u[i] * u[i] + (double)v[i] * v[i] + 1.0);
            predicted_level[i] = q[i]->PredictScale(d, K->mfLogScaleFactor);
        }
    }
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchByProjection(K, eye4(), q, matched, th);
    report_assignment(matched, q, assigned);
    return n;
}

} // extern "C"
