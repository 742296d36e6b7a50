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
#include <cstdio>
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


/* ================================================================================================================
 * DBoW2 transform, SearchByBoW x2, SearchForTriangulation, Fuse x2, SearchBySim3,
 * MapPoint::ComputeDistinctiveDescriptors
 * ================================================================================================================ */

/* An ORBVocabulary built by the reference's own text loader (TemplatedVocabulary::loadFromTextFile, the format of the
 * ORBvoc.txt the reference ships instructions for): header "k L scoring weighting", then one line per node
 * "parent isLeaf d0 .. d31 weight" in node-id order.  Nodes 1 .. nnodes-1 (0 is the root); TF_IDF weighting, L1 scoring. */
void* ref2_voc_create(int k, int L, int nnodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight,
                      const char* tmp_path)
{
    FILE* f = std::fopen(tmp_path, "w");
    if (!f) return nullptr;
    std::fprintf(f, "%d %d 0 0", k, L);
    for (int i = 1; i < nnodes; ++i) {
        std::fprintf(f, "\n%d %d", parent[i], (int)is_leaf[i]);
        for (int b = 0; b < 32; ++b) std::fprintf(f, " %d", (int)desc[(size_t)i * 32 + b]);
        std::fprintf(f, " %.17g", weight[i]);                  /* round-trips the double exactly */
    }
    std::fclose(f);                                            /* no trailing newline: the loader would read a phantom node */
    ORBVocabulary* v = new ORBVocabulary();
    const bool ok = v->loadFromTextFile(tmp_path);
    std::remove(tmp_path);
    if (!ok || (int)v->m_nodes.size() != nnodes) { delete v; return nullptr; }
    return v;
}
void ref2_voc_destroy(void* v) { delete static_cast<ORBVocabulary*>(v); }

/* TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup) (TemplatedVocabulary.h:1231-1271) per feature */
void ref2_voc_transform(void* voc, const uint8_t* feats, int n, int levelsup, uint32_t* word, double* weight, uint32_t* node)
{
    ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
    for (int i = 0; i < n; ++i) {
        cv::Mat d = desc_mat(feats + (size_t)i * 32, 1);
        DBoW2::WordId w; DBoW2::WordValue val; DBoW2::NodeId nid = 0;
        v->transform(d, w, val, &nid, levelsup);
        word[i] = w; weight[i] = val; node[i] = nid;
    }
}

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1140-1218), what Frame::ComputeBoW
 * (src/Frame.cc:394-402) calls.  Outputs in std::map order; returns the BowVector size, *nfv = FeatureVector entries. */
int ref2_voc_transform_all(void* voc, const uint8_t* feats, int n, int levelsup, uint32_t* bow_word, double* bow_value,
                           uint32_t* fv_node, uint32_t* fv_feat, int* nfv)
{
    ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
    std::vector<cv::Mat> vd = Converter::toDescriptorVector(desc_mat(feats, n));
    DBoW2::BowVector bow; DBoW2::FeatureVector fv;
    v->transform(vd, bow, fv, levelsup);
    int i = 0;
    for (auto& e : bow) { bow_word[i] = e.first; bow_value[i] = e.second; ++i; }
    int j = 0;
    for (auto& e : fv) for (unsigned int idx : e.second) { fv_node[j] = e.first; fv_feat[j] = idx; ++j; }
    *nfv = j;
    return i;
}

namespace {
/* mFeatVec as Frame::ComputeBoW leaves it: addFeature(node, i) in feature order for the features whose word has a weight */
void fill_featvec(DBoW2::FeatureVector& fv, const uint32_t* node, const uint8_t* flags, int n)
{
    for (int i = 0; i < n; ++i) if (flags[i] & 1) fv.addFeature(node[i], (unsigned)i);
}
/* flags bit 1: the feature holds a good MapPoint, bit 2: it holds a bad one */
void fill_points(World& w, KeyFrame* K, const orbm_frame* f, const uint8_t* flags, std::vector<MapPoint*>& pts)
{
    pts.assign(f->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < f->n; ++i)
        if (flags[i] & 6) {
            pts[i] = w.make_point(0, 0, 1, f->descriptors + (size_t)i * 32, 1);
            if (flags[i] & 4) pts[i]->mbBad = true;
            K->mvpMapPoints[i] = pts[i];
        }
}
int index_of(const std::vector<MapPoint*>& pts, MapPoint* p)
{
    if (!p) return -1;
    for (size_t i = 0; i < pts.size(); ++i) if (pts[i] == p) return (int)i;
    return -2;
}
}

/* ORBmatcher::SearchByBoW.  variant 0: (KeyFrame* pKF = a, Frame& F = b, vpMapPointMatches) src/ORBmatcher.cc:159-288;
 * variant 1: (KeyFrame* pKF1 = a, KeyFrame* pKF2 = b, vpMatches12) :522-655.  match_a[i] / match_b[j]: partner or -1. */
int ref2_search_by_bow(int variant, const orbm_frame* fa, const uint32_t* node_a, const uint8_t* flags_a, const orbm_frame* fb,
                       const uint32_t* node_b, const uint8_t* flags_b, float nnratio, int check_orientation, int32_t* match_a, int32_t* match_b)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    KeyFrame* A = w.make_keyframe(fa, kUnitCam, eye4());
    fill_featvec(A->mFeatVec, node_a, flags_a, fa->n);
    std::vector<MapPoint*> pa, pb;
    fill_points(w, A, fa, flags_a, pa);
    ORBmatcher matcher(nnratio, check_orientation != 0);
    for (int i = 0; i < fa->n; ++i) match_a[i] = -1;
    for (int j = 0; j < fb->n; ++j) match_b[j] = -1;
    int n;
    if (variant == 0) {
        Frame* F = make_frame(fb, kUnitCam);
        w.frames.push_back(F);
        fill_featvec(F->mFeatVec, node_b, flags_b, fb->n);
        std::vector<MapPoint*> m;
        n = matcher.SearchByBoW(A, *F, m);
        for (int j = 0; j < fb->n; ++j) { const int i = index_of(pa, m[j]); match_b[j] = i; if (i >= 0) match_a[i] = j; }
    } else {
        KeyFrame* B = w.make_keyframe(fb, kUnitCam, eye4());
        fill_featvec(B->mFeatVec, node_b, flags_b, fb->n);
        fill_points(w, B, fb, flags_b, pb);
        std::vector<MapPoint*> m;
        n = matcher.SearchByBoW(A, B, m);
        for (int i = 0; i < fa->n; ++i) { const int j = index_of(pb, m[i]); match_a[i] = j; if (j >= 0) match_b[j] = i; }
    }
    return n;
}

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823).  KF2 sits at the origin and KF1's camera centre at
 * (ex, ey, 1), so the epipole the reference computes (:664-670) is exactly (ex, ey) with the unit camera. */
int ref2_search_for_triangulation(const orbm_frame* f1, const uint32_t* node1, const uint8_t* flags1, const orbm_frame* f2,
                                  const uint32_t* node2, const uint8_t* flags2, const float* F12, float ex, float ey, int only_stereo,
                                  int check_orientation, int32_t* matches12)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    cv::Mat T1 = eye4();
    T1.at<float>(0, 3) = -ex; T1.at<float>(1, 3) = -ey; T1.at<float>(2, 3) = -1.f;
    KeyFrame* K1 = w.make_keyframe(f1, kUnitCam, T1);
    KeyFrame* K2 = w.make_keyframe(f2, kUnitCam, eye4());
    fill_featvec(K1->mFeatVec, node1, flags1, f1->n);
    fill_featvec(K2->mFeatVec, node2, flags2, f2->n);
    std::vector<MapPoint*> p1, p2;
    fill_points(w, K1, f1, flags1, p1);
    fill_points(w, K2, f2, flags2, p2);
    cv::Mat F(3, 3, CV_32F);
    for (int i = 0; i < 9; ++i) F.at<float>(i / 3, i % 3) = F12[i];
    std::vector<std::pair<size_t, size_t>> pairs;
    ORBmatcher matcher(0.6f, check_orientation != 0);
    const int n = matcher.SearchForTriangulation(K1, K2, F, pairs, only_stereo != 0);
    for (int i = 0; i < f1->n; ++i) matches12[i] = -1;
    for (auto& pr : pairs) matches12[pr.first] = (int32_t)pr.second;
    return n;
}

/* ORBmatcher::Fuse.  variant 0: (pKF, vpMapPoints, th) src/ORBmatcher.cc:825-973; variant 1: (pKF, Scw = identity, vpPoints,
 * th, vpReplacePoint) :975-1100.  The keyframe starts without map points; a query's answer is read back from the side
 * effects: the slot it was added to, or the slot of the point it was replaced by / reported against. */
int ref2_fuse(int variant, const orbm_frame* kf, float bf, int nq, const float* u, const float* v, const float* max_dist,
              const uint8_t* qdesc, const int32_t* nobs, float th, int32_t* best_idx, int32_t* predicted_level)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    const Cam cam = {1.f, 1.f, 0.f, 0.f, bf};
    KeyFrame* K = w.make_keyframe(kf, cam, eye4());
    std::vector<MapPoint*> q(nq);
    for (int i = 0; i < nq; ++i) {
        q[i] = w.make_point(u[i], v[i], 1.f, qdesc + (size_t)i * 32, nobs ? nobs[i] : 1, max_dist[i], 0.f);
        if (predicted_level) {
            const float d = (float)std::sqrt((double)u[i] * u[i] + (double)v[i] * v[i] + 1.0);
            predicted_level[i] = q[i]->PredictScale(d, K->mfLogScaleFactor);
        }
    }
    ORBmatcher matcher(0.6f, true);
    int n;
    std::vector<MapPoint*> repl(nq, static_cast<MapPoint*>(nullptr));
    if (variant == 0) n = matcher.Fuse(K, q, th);
    else n = matcher.Fuse(K, eye4(), q, th, repl);
    for (int i = 0; i < nq; ++i) {
        MapPoint* p = variant == 1 && repl[i] ? repl[i] : q[i];
        for (int hop = 0; hop < nq + 2 && p && p->mbBad && p->mpReplaced; ++hop) p = p->mpReplaced;
        best_idx[i] = (p && p->mObservations.count(K)) ? (int32_t)p->mObservations[K] : -1;
    }
    return n;
}

/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326) with both keyframes at the origin, s12 = 1, R12 = I, t12 = 0:
 * map point i of KF1 lies at (u1[i], v1[i], 1) and so projects to (u1[i], v1[i]) in KF2, and the other way round.
 * has1 / has2: the feature holds a map point; pre12[i] >= 0: vpMatches12[i] already holds the point of KF2 feature pre12[i].
 * match12[i] (out): KF2 feature newly matched to KF1 feature i, or -1. */
int ref2_search_by_sim3(const orbm_frame* f1, const orbm_frame* f2, const uint8_t* has1, const float* u1, const float* v1,
                        const float* maxd1, const uint8_t* has2, const float* u2, const float* v2, const float* maxd2,
                        const int32_t* pre12, float th, int32_t* match12, int32_t* level1, int32_t* level2)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    KeyFrame* K1 = w.make_keyframe(f1, kUnitCam, eye4());
    KeyFrame* K2 = w.make_keyframe(f2, kUnitCam, eye4());
    std::vector<MapPoint*> p1(f1->n, static_cast<MapPoint*>(nullptr)), p2(f2->n, static_cast<MapPoint*>(nullptr));
    auto lvl = [](MapPoint* p, float u, float v, KeyFrame* K) {
        const float d = (float)std::sqrt((double)u * u + (double)v * v + 1.0);
        return p->PredictScale(d, K->mfLogScaleFactor);
    };
    for (int i = 0; i < f1->n; ++i) if (has1[i]) {
        p1[i] = w.make_point(u1[i], v1[i], 1.f, f1->descriptors + (size_t)i * 32, 1, maxd1[i], 0.f);
        K1->mvpMapPoints[i] = p1[i];
        p1[i]->mObservations[K1] = i;
        if (level1) level1[i] = lvl(p1[i], u1[i], v1[i], K2);
    }
    for (int i = 0; i < f2->n; ++i) if (has2[i]) {
        p2[i] = w.make_point(u2[i], v2[i], 1.f, f2->descriptors + (size_t)i * 32, 1, maxd2[i], 0.f);
        K2->mvpMapPoints[i] = p2[i];
        p2[i]->mObservations[K2] = i;
        if (level2) level2[i] = lvl(p2[i], u2[i], v2[i], K1);
    }
    std::vector<MapPoint*> m12(f1->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < f1->n; ++i) if (pre12[i] >= 0) m12[i] = p2[pre12[i]];
    cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t12 = cv::Mat::zeros(3, 1, CV_32F);
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchBySim3(K1, K2, m12, 1.0f, R12, t12, th);
    for (int i = 0; i < f1->n; ++i) match12[i] = (pre12[i] >= 0 || !m12[i]) ? -1 : index_of(p2, m12[i]);
    return n;
}

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305): n observations, one keyframe each.  mObservations is
 * a std::map keyed by KeyFrame*: the keyframes are sorted by address and descriptor j goes to the j-th of them, so the
 * reference walks the descriptors in the caller's order.  out32 = the descriptor the reference keeps. */
void ref2_distinctive(const uint8_t* desc, int n, uint8_t* out32)
{
    std::lock_guard<std::mutex> g(g_lock);
    World w;
    static const uint8_t zero[32] = {0};
    orbx_keypoint kp; std::memset(&kp, 0, sizeof(kp)); kp.x = 50; kp.y = 50; kp.size = 31; kp.class_id = -1;
    orbm_frame f; std::memset(&f, 0, sizeof(f));
    f.n = 1; f.keys_un = &kp; f.descriptors = zero; f.min_x = 0; f.max_x = 640; f.min_y = 0; f.max_y = 480;
    MapPoint* p = w.make_point(0, 0, 1, zero, 0);
    std::vector<KeyFrame*> ks;
    for (int j = 0; j < n; ++j) ks.push_back(w.make_keyframe(&f, kUnitCam, eye4()));
    std::sort(ks.begin(), ks.end());
    for (int j = 0; j < n; ++j) {
        std::memcpy(ks[j]->mDescriptors.data, desc + (size_t)j * 32, 32);      /* the member is const: its pixels are not */
        p->mObservations[ks[j]] = 0;
    }
    p->nObs = n;
    p->ComputeDistinctiveDescriptors();
    std::memcpy(out32, p->mDescriptor.data, 32);
}

} // extern "C"
