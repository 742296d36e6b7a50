// frame_oracle.cpp — CPU restatement of the steps of the reference's RGB-D Frame construction that sit
// either side of ORBextractor::operator().  TEST INFRASTRUCTURE ONLY: loaded by
// tests/, __graft_entry__.smoke() and bench.py's CPU legs; the product (liborbx.so) never links it.
//
//   oracle_cvt_gray            cv::cvtColor(RGB/BGR/RGBA/BGRA -> GRAY), Tracking.cc:308-321 (and :265-290 stereo)
//   oracle_depth_to_float      imDepth.convertTo(CV_32F, mDepthMapFactor), Tracking.cc:324-327
//   oracle_undistort_points    cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK), Frame.cc:404-434, 436-464
//   oracle_image_bounds        Frame::ComputeImageBounds, Frame.cc:436-464
//   oracle_stereo_from_rgbd    Frame::ComputeStereoFromRGBD, Frame.cc:641-662
//   oracle_generate_point_cloud  PointCloudMapping::generatePointCloud / convertToPointCloud, PointCloudMap.cc:78-139
//                              with pcl::transformPointCloud<PointT, double> (PCL >= 1.7, not vendored: its published
//                              per-point arithmetic is restated; parity unpinned for this function)
//
// The OpenCV arithmetic is that of 4.13.0, pinned by tests/golden/frame_front.npz (generated with the cv2
// wheel by tests/golden/make_golden_frame.py): 15-bit luma coefficients with round-half-up for 8U cvtColor;
// five fixed-point iterations in double for undistortPoints (default TermCriteria(COUNT, 5)); float(src) * float(alpha)
// for 16U -> 32F convertTo (restated from the OpenCV sources: the cv2 Python module has no Mat::convertTo).
// Compile with -ffp-contract=off: every double operation below is separately rounded.

#include <cfloat>
#include <climits>
#include <cmath>
#include <cstddef>
#include <climits>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <utility>
#include <vector>

extern "C" {

// cv::cvtColor(src, dst, CV_{RGB,BGR,RGBA,BGRA}2GRAY) for 8U: (R*9798 + G*19235 + B*3735 + 2^14) >> 15
void oracle_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int rgb_order,
                     uint8_t* dst, size_t dstride)
{
    const int RY = 9798, GY = 19235, BY = 3735;
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * sstride;
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x, s += channels) {
            const int c0 = s[0], c1 = s[1], c2 = s[2];
            const int r = rgb_order ? c0 : c2, b = rgb_order ? c2 : c0;
            d[x] = (uint8_t)((r * RY + c1 * GY + b * BY + (1 << 14)) >> 15);
        }
    }
}

// imDepth.convertTo(imDepth, CV_32F, factor) for a 16U depth map (a 32F map with factor 1 is left alone by the reference)
void oracle_depth_to_float(const uint16_t* src, int w, int h, size_t sstride_elems, float factor, float* dst, size_t dstride_elems)
{
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) dst[(size_t)y * dstride_elems + x] = (float)src[(size_t)y * sstride_elems + x] * factor;
}

// cv::undistortPoints with R = I and P = K: cam = (fx, fy, cx, cy), dist = (k1, k2, p1, p2, k3) as floats
// (mK / mDistCoef are CV_32F, Tracking.cc:60-80; OpenCV widens them to double)
void oracle_undistort_points(const float* xy_in, int n, const float* cam, const float* dist, float* xy_out)
{
    const double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    const double ifx = 1. / fx, ify = 1. / fy;
    const double k1 = dist[0], k2 = dist[1], p1 = dist[2], p2 = dist[3], k3 = dist[4];
    for (int i = 0; i < n; ++i) {
        const double u = xy_in[2 * i], v = xy_in[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) {
            const double r2 = x * x + y * y;
            // k[5..7] (k4..k6) and the thin-prism terms k[8..11] are zero for a 5-coefficient model
            const double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x) + 0 * r2 + 0 * r2 * r2;
            const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y + 0 * r2 + 0 * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        // P = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1 / (0*x + 0*y + 1)
        const double xx = fx * x + 0 * y + cx, yy = 0 * x + fy * y + cy, ww = 1. / (0 * x + 0 * y + 1);
        xy_out[2 * i] = (float)(xx * ww);
        xy_out[2 * i + 1] = (float)(yy * ww);
    }
}

// Frame::ComputeImageBounds: bounds = (minX, maxX, minY, maxY)
void oracle_image_bounds(int cols, int rows, const float* cam, const float* dist, float* bounds)
{
    if (dist[0] != 0.0f) {
        const float c[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
        float u[8];
        oracle_undistort_points(c, 4, cam, dist, u);
        bounds[0] = std::fmin(u[0], u[4]);
        bounds[1] = std::fmax(u[2], u[6]);
        bounds[2] = std::fmin(u[1], u[3]);
        bounds[3] = std::fmax(u[5], u[7]);
    } else {
        bounds[0] = 0.0f; bounds[1] = (float)cols; bounds[2] = 0.0f; bounds[3] = (float)rows;
    }
}

// Frame::ComputeStereoFromRGBD: keys = distorted keypoint positions, keys_un = undistorted; depth is CV_32F
void oracle_stereo_from_rgbd(const float* keys_xy, const float* keys_un_xy, int n, const float* depth, int w, int h,
                             size_t dstride_elems, float bf, float* u_right, float* z)
{
    (void)w; (void)h;
    for (int i = 0; i < n; ++i) {
        u_right[i] = -1.f; z[i] = -1.f;
        const int v = (int)keys_xy[2 * i + 1], u = (int)keys_xy[2 * i];     // imDepth.at<float>(v, u) with float -> int conversion
        const float d = depth[(size_t)v * dstride_elems + u];
        if (d > 0) { z[i] = d; u_right[i] = keys_un_xy[2 * i] - bf / d; }
    }
}

// One keyframe.  pose_inv = T.inverse().matrix() row-major, or NULL (convertToPointCloud).  out: 16-byte records
// {float x, y, z; u8 b, g, r, a}.  Returns the number of points.
int oracle_generate_point_cloud(const float* depth, size_t dstride_elems, const uint8_t* bgr, size_t cstride, int cols, int rows,
                                int step, float fx, float fy, float cx, float cy, const double* pose_inv, uint8_t* out)
{
    int count = 0;
    for (int m = 0; m < rows; m += step)
        for (int n = 0; n < cols; n += step) {
            const float d = depth[(size_t)m * dstride_elems + n];
            if (d < 0.01 || d > 10) continue;
            float p[3];
            p[2] = d;
            p[0] = (n - cx) * p[2] / fx;
            p[1] = (m - cy) * p[2] / fy;
            uint8_t c[4] = {0, 0, 0, 255};                 // PointXYZRGBA(): r = g = b = 0, a = 255
            if (bgr) { c[0] = bgr[(size_t)m * cstride + n * 3]; c[1] = bgr[(size_t)m * cstride + n * 3 + 1]; c[2] = bgr[(size_t)m * cstride + n * 3 + 2]; }
            float o[3] = {p[0], p[1], p[2]};
            if (pose_inv) {
                // pcl::transformPointCloud, Scalar = double: static_cast<float>(tf(r,0)*x + tf(r,1)*y + tf(r,2)*z + tf(r,3))
                const double* tf = pose_inv;
                for (int r = 0; r < 3; ++r)
                    o[r] = static_cast<float>(tf[4 * r + 0] * p[0] + tf[4 * r + 1] * p[1] + tf[4 * r + 2] * p[2] + tf[4 * r + 3]);
            }
            uint8_t* rec = out + (size_t)count * 16;
            for (int k = 0; k < 3; ++k) { const float v = o[k]; const uint8_t* b = reinterpret_cast<const uint8_t*>(&v); for (int t = 0; t < 4; ++t) rec[4 * k + t] = b[t]; }
            for (int t = 0; t < 4; ++t) rec[12 + t] = c[t];
            ++count;
        }
    return count;
}

// pcl::VoxelGrid<pcl::PointXYZRGBA>::applyFilter with its default parameters (downsample_all_data, no field filter, no minimum
// count), the call at src/PointCloudMap.cc:240-241, 259-260, 277-280 with the leaf size of :41.  PCL (>= 1.7, CMakeLists.txt:39) is
// not vendored under /root/reference: this restates its published algorithm (pcl/filters/impl/voxel_grid.hpp) — PARITY UNPINNED:
//   bounds of the finite points; inverse leaf = 1 / leaf; min_b = floor(min * inv), div_b = floor(max * inv) - min_b + 1;
//   a grid whose cell count would overflow an int is refused (PCL warns and passes the cloud through; here: -2);
//   idx = i + j * div_x + k * div_x * div_y with i = int(floor(x * inv) - float(min_b.x));
//   one output point per occupied voxel in ascending idx order: the mean of its points (x, y, z summed as float, r, g, b as
//   integers then truncated after the float division, a = 255 as PCL >= 1.8 leaves it).
// DETERMINISATION: PCL sorts (idx, point) pairs with an unstable std::sort and adds a voxel's points in that order, so its float
// centroids are not bit-defined; here a voxel's points are added in input order (a stable sort), one legitimate execution.
// points / out: 16-byte records {float x, y, z; u8 b, g, r, a}.  Returns the voxel count (-1: cap too small, -2: index overflow).
int oracle_voxel_grid(const uint8_t* points, int n, float leaf, uint8_t* out, int cap)
{
    struct P { float x, y, z; uint8_t b, g, r, a; };
    const P* p = reinterpret_cast<const P*>(points);
    P* o = reinterpret_cast<P*>(out);
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = 0; i < n; ++i) {
        if (!std::isfinite(p[i].x) || !std::isfinite(p[i].y) || !std::isfinite(p[i].z)) continue;
        const float v[3] = {p[i].x, p[i].y, p[i].z};
        for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], v[k]); mx[k] = std::max(mx[k], v[k]); }
    }
    if (mn[0] > mx[0]) return 0;
    const float inv = 1.0f / leaf;
    long long dchk = 1, div[3];
    int minb[3];
    for (int k = 0; k < 3; ++k) {
        dchk *= (long long)((mx[k] - mn[k]) * inv) + 1;
        if (dchk > INT_MAX) return -2;
        minb[k] = (int)std::floor(mn[k] * inv);
        div[k] = (long long)(int)std::floor(mx[k] * inv) - minb[k] + 1;
    }
    if (div[0] * div[1] * div[2] > INT_MAX) return -2;
    const int mul1 = (int)div[0], mul2 = (int)(div[0] * div[1]);
    std::vector<std::pair<uint32_t, int>> iv;
    iv.reserve(n);
    for (int i = 0; i < n; ++i) {
        if (!std::isfinite(p[i].x) || !std::isfinite(p[i].y) || !std::isfinite(p[i].z)) continue;
        const int a = (int)(std::floor(p[i].x * inv) - (float)minb[0]);
        const int b = (int)(std::floor(p[i].y * inv) - (float)minb[1]);
        const int c = (int)(std::floor(p[i].z * inv) - (float)minb[2]);
        iv.push_back(std::make_pair((uint32_t)(a + b * mul1 + c * mul2), i));
    }
    std::stable_sort(iv.begin(), iv.end(), [](const std::pair<uint32_t, int>& l, const std::pair<uint32_t, int>& r) { return l.first < r.first; });
    int nv = 0;
    for (size_t s = 0; s < iv.size();) {
        size_t e = s;
        float sx = 0.f, sy = 0.f, sz = 0.f;
        unsigned sr = 0, sg = 0, sb = 0;
        while (e < iv.size() && iv[e].first == iv[s].first) {
            const P& q = p[iv[e].second];
            sx += q.x; sy += q.y; sz += q.z; sr += q.r; sg += q.g; sb += q.b;
            ++e;
        }
        if (nv >= cap) return -1;
        const float fc = (float)(e - s);
        P r;
        r.x = sx / fc; r.y = sy / fc; r.z = sz / fc;
        r.r = (uint8_t)(int)((float)sr / fc); r.g = (uint8_t)(int)((float)sg / fc); r.b = (uint8_t)(int)((float)sb / fc); r.a = 255;
        o[nv++] = r;
        s = e;
    }
    return nv;
}

// The feature loop of Map::_WriteKeyFrame (src/Map.cc:160-180) into a byte buffer: kps = 28-byte cv::KeyPoint records.
// has_mp[i] == 0 -> ULONG_MAX.  Returns bytes written.
size_t oracle_write_keyframe_features(const uint8_t* kps, const uint8_t* desc, const uint64_t* mp_index, const uint8_t* has_mp,
                                      int N, uint8_t* out)
{
    uint8_t* f = out;
    for (int i = 0; i < N; i++) {
        const uint8_t* kp = kps + (size_t)i * 28;
        std::memcpy(f, kp + 0, 4); f += 4;       // pt.x
        std::memcpy(f, kp + 4, 4); f += 4;       // pt.y
        std::memcpy(f, kp + 8, 4); f += 4;       // size
        std::memcpy(f, kp + 12, 4); f += 4;      // angle
        std::memcpy(f, kp + 16, 4); f += 4;      // response
        std::memcpy(f, kp + 20, 4); f += 4;      // octave
        for (int j = 0; j < 32; j++) *f++ = desc[(size_t)i * 32 + j];
        unsigned long int mpidx;
        if (!has_mp || !has_mp[i]) mpidx = ULONG_MAX;
        else mpidx = mp_index[i];
        std::memcpy(f, &mpidx, sizeof(mpidx)); f += sizeof(mpidx);
    }
    return (size_t)(f - out);
}

// pcl::StatisticalOutlierRemoval<PointXYZRGBA>::applyFilterIndices as PointCloudMapping uses it on the global map
// (reference src/PointCloudMap.cc:46-47 setMeanK(50), setStddevMulThresh(1.0); :284-285 sor.setInputCloud / sor.filter).
// PARITY UNPINNED: PCL is an external dependency that is absent from /root/reference (no version is pinned there either);
// this restates the published algorithm of PCL 1.7-1.9, filters/impl/statistical_outlier_removal.hpp:
//   searcher_k = mean_k + 1 nearest neighbours of every finite point (the point itself is result 0); squared distances as FLANN's
//   L2_Simple<float> accumulates them (result += diff * diff over x, y, z in float); dist_sum += sqrt(nn_dists[k]) in double for
//   k = 1 .. mean_k in ascending order (plain sqrt on a float argument: the double overload in those versions);
//   distances[i] = (float)(dist_sum / mean_k); non-finite points and points whose search returns fewer neighbours get 0.0 and are
//   not counted; sum += d, sq_sum += d * d (float product) serially in double; mean = sum / valid; variance = (sq_sum - sum * sum /
//   valid) / (valid - 1); threshold = mean + mul * sqrt(variance); a point is removed when distances[i] > threshold.
// The k nearest neighbours are found by brute force (exact, like the kd-tree with eps = 0; ties at the k-th distance change
// nothing because equal distances add equal terms).  points / out: 16-byte records.  stats: threshold, mean, stddev, valid.
long long oracle_statistical_outlier_removal(const uint8_t* points, long long n, int mean_k, double mul, uint8_t* out,
                                             float* mean_dist, double* stats)
{
    struct P { float x, y, z; uint8_t b, g, r, a; };
    const P* p = reinterpret_cast<const P*>(points);
    P* o = reinterpret_cast<P*>(out);
    const int K = mean_k + 1;
    std::vector<float> distances((size_t)n, 0.0f), heap;
    long long finite = 0;
    for (long long i = 0; i < n; ++i) finite += std::isfinite(p[i].x) && std::isfinite(p[i].y) && std::isfinite(p[i].z);
    long long valid = 0;
    for (long long i = 0; i < n; ++i) {
        if (!std::isfinite(p[i].x) || !std::isfinite(p[i].y) || !std::isfinite(p[i].z)) continue;
        if (finite < K) continue;                                  // nearestKSearch returns fewer than searcher_k results
        heap.clear();                                              // max-heap of the K smallest squared distances
        for (long long j = 0; j < n; ++j) {
            if (!std::isfinite(p[j].x) || !std::isfinite(p[j].y) || !std::isfinite(p[j].z)) continue;
            float d2 = 0.0f;
            const float dx = p[i].x - p[j].x; d2 += dx * dx;
            const float dy = p[i].y - p[j].y; d2 += dy * dy;
            const float dz = p[i].z - p[j].z; d2 += dz * dz;
            if ((int)heap.size() < K) { heap.push_back(d2); std::push_heap(heap.begin(), heap.end()); }
            else if (d2 < heap.front()) { std::pop_heap(heap.begin(), heap.end()); heap.back() = d2; std::push_heap(heap.begin(), heap.end()); }
        }
        std::sort(heap.begin(), heap.end());
        double dist_sum = 0.0;
        for (int k = 1; k < K; ++k) dist_sum += std::sqrt((double)heap[k]);
        distances[i] = static_cast<float>(dist_sum / mean_k);
        ++valid;
    }
    double sum = 0, sq_sum = 0;
    for (long long i = 0; i < n; ++i) { sum += distances[i]; sq_sum += distances[i] * distances[i]; }
    const double mean = sum / static_cast<double>(valid);
    const double variance = (sq_sum - sum * sum / static_cast<double>(valid)) / (static_cast<double>(valid) - 1);
    const double stddev = std::sqrt(variance);
    const double thr = mean + mul * stddev;
    long long kept = 0;
    for (long long i = 0; i < n; ++i) {
        if (distances[i] > thr) continue;
        if (o) o[kept] = p[i];
        ++kept;
    }
    if (mean_dist) for (long long i = 0; i < n; ++i) mean_dist[i] = distances[i];
    if (stats) { stats[0] = thr; stats[1] = mean; stats[2] = stddev; stats[3] = (double)valid; }
    return kept;
}

} // extern "C"
