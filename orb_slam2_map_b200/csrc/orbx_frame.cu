// orbx_frame.cu — the steps of the reference's RGB-D Frame construction either side of the extractor
// (SURVEY.md §8(f) rank 1), as sm_100a kernels behind the C ABI:
//
//   K12 cvt_gray_kernel      cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY)            src/Tracking.cc:308-321
//   K13 depth_float_kernel   imDepth.convertTo(CV_32F, mDepthMapFactor)         src/Tracking.cc:324-327
//   K14 undistort_kernel     cv::undistortPoints (K, D, R = I, P = K)           src/Frame.cc:404-434, 436-464
//   K15 frame_front_kernel   UndistortKeyPoints + ComputeStereoFromRGBD         src/Frame.cc:404-434, 641-662
//
// OpenCV 4.13 arithmetic (oracle/frame_oracle.cpp): 15-bit luma coefficients, round half up; five fixed-point
// iterations in double with every operation separately rounded (no FMA contraction).

#include "orbx_internal.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>

typedef uint8_t u8;

struct CamD { double fx, fy, cx, cy, ifx, ify, k1, k2, p1, p2, k3; int distorted; };

static CamD cam_double(const orbx_camera* c)
{
    CamD d;
    d.fx = c->fx; d.fy = c->fy; d.cx = c->cx; d.cy = c->cy;
    d.ifx = 1. / d.fx; d.ify = 1. / d.fy;
    d.k1 = c->k1; d.k2 = c->k2; d.p1 = c->p1; d.p2 = c->p2; d.k3 = c->k3;
    d.distorted = c->k1 != 0.0f;           // mDistCoef.at<float>(0)==0.0 -> mvKeysUn = mvKeys (Frame.cc:406)
    return d;
}

// one 8-bit gray pixel: (R*9798 + G*19235 + B*3735 + 2^14) >> 15
__device__ __forceinline__ uint32_t luma(uint32_t r, uint32_t g, uint32_t b)
{
    return (r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15;
}

// thread = 4 adjacent output pixels of one row
__global__ void __launch_bounds__(256)
cvt_gray_kernel(const u8* __restrict__ src, long long sFrame, long long sRow, int channels, int rgbOrder,
                u8* __restrict__ dst, long long dFrame, long long dRow, int w, int h, int vecOk)
{
    const int x4 = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4, y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x4 >= w || y >= h) return;
    const u8* s = src + (long long)blockIdx.z * sFrame + (long long)y * sRow + (long long)x4 * channels;
    u8* d = dst + (long long)blockIdx.z * dFrame + (long long)y * dRow + x4;
    const int i0 = rgbOrder ? 0 : 2, i2 = 2 - i0;
    if (vecOk && x4 + 4 <= w) {
        uint32_t g[4];
        if (channels == 3) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(s);
            const uint32_t a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);      // 12 bytes = 4 pixels
            const uint32_t px[4] = {a & 0xffffff, (a >> 24) | ((b & 0xffff) << 8), (b >> 16) | ((c & 0xff) << 16), c >> 8};
#pragma unroll
            for (int k = 0; k < 4; ++k) g[k] = luma((px[k] >> (8 * i0)) & 0xff, (px[k] >> 8) & 0xff, (px[k] >> (8 * i2)) & 0xff);
        } else {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(s));
            const uint32_t px[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) g[k] = luma((px[k] >> (8 * i0)) & 0xff, (px[k] >> 8) & 0xff, (px[k] >> (8 * i2)) & 0xff);
        }
        *reinterpret_cast<uint32_t*>(d) = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
    } else {
        for (int k = 0; k < 4 && x4 + k < w; ++k) {
            const u8* p = s + k * channels;
            d[k] = (u8)luma(p[i0], p[1], p[i2]);
        }
    }
}

__global__ void __launch_bounds__(256)
depth_float_kernel(const uint16_t* __restrict__ src, long long sRow, float* __restrict__ dst, long long dRow, int w, int h, float factor)
{
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (x < w && y < h) dst[(long long)y * dRow + x] = __fmul_rn((float)src[(long long)y * sRow + x], factor);
}

// cvUndistortPointsInternal, 5 iterations, R = I, P = K (the terms that are zero for a 5-coefficient model are
// kept out: adding an exact zero changes nothing)
__device__ __forceinline__ void undistort_point(const CamD& c, float uf, float vf, float& xo, float& yo)
{
    const double u = (double)uf, v = (double)vf;
    double x = __dmul_rn(__dsub_rn(u, c.cx), c.ifx), y = __dmul_rn(__dsub_rn(v, c.cy), c.ify);
    const double x0 = x, y0 = y;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
        const double xx = __dmul_rn(x, x), yy = __dmul_rn(y, y);
        const double r2 = __dadd_rn(xx, yy);
        const double den = __dadd_rn(1., __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(c.k3, r2), c.k2), r2), c.k1), r2));
        const double icdist = __ddiv_rn(1., den);
        if (icdist < 0) { x = x0; y = y0; break; }
        // deltaX = 2*p1*x*y + p2*(r2 + 2*x*x);  deltaY = p1*(r2 + 2*y*y) + 2*p2*x*y   (left-to-right products)
        const double dX = __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2., c.p1), x), y), __dmul_rn(c.p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2., x), x))));
        const double dY = __dadd_rn(__dmul_rn(c.p1, __dadd_rn(r2, __dmul_rn(__dmul_rn(2., y), y))), __dmul_rn(__dmul_rn(__dmul_rn(2., c.p2), x), y));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    xo = (float)__dadd_rn(__dmul_rn(c.fx, x), c.cx);
    yo = (float)__dadd_rn(__dmul_rn(c.fy, y), c.cy);
}

__global__ void __launch_bounds__(128)
undistort_kernel(CamD c, const float* __restrict__ in, int n, float* __restrict__ out)
{
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    float x = in[2 * i], y = in[2 * i + 1];
    if (c.distorted) undistort_point(c, x, y, x, y);
    out[2 * i] = x; out[2 * i + 1] = y;
}

// thread = one keypoint of one frame: mvKeysUn record, mvuRight, mvDepth
__global__ void __launch_bounds__(128)
frame_front_kernel(CamD c, float bf, const orbx_keypoint* __restrict__ kps, const int* __restrict__ count, int cap,
                   int firstSlot, const u8* __restrict__ depth, int depthU16, float factor, int scaleDepth,
                   long long dFrame, long long dRow, orbx_keypoint* __restrict__ kpsUn,
                   float* __restrict__ uRight, float* __restrict__ zOut,
                   orbx_keypoint* __restrict__ packUn = nullptr, float* __restrict__ packUR = nullptr, float* __restrict__ packZ = nullptr)
{
    const int slot = firstSlot + blockIdx.y, k = blockIdx.x * 128 + threadIdx.x;
    if (k >= cap) return;
    const size_t o = (size_t)slot * cap + k;
    if (k >= count[slot]) { uRight[o] = -1.f; zOut[o] = -1.f; return; }
    orbx_keypoint kp = kps[o];
    const float kx = kp.x, ky = kp.y;
    if (c.distorted) undistort_point(c, kx, ky, kp.x, kp.y);
    kpsUn[o] = kp;
    // const float d = imDepth.at<float>(v, u) with v = kp.pt.y, u = kp.pt.x converted to int (Frame.cc:651-654)
    const u8* row = depth + (long long)blockIdx.y * dFrame + (long long)(int)ky * dRow;
    float d;
    if (depthU16) d = __fmul_rn((float)reinterpret_cast<const uint16_t*>(row)[(int)kx], factor);
    else { d = reinterpret_cast<const float*>(row)[(int)kx]; if (scaleDepth) d = __fmul_rn(d, factor); }
    float ur = -1.f, z = -1.f;
    if (d > 0) { z = d; ur = __fsub_rn(kp.x, __fdiv_rn(bf, d)); }
    uRight[o] = ur; zOut[o] = z;
    if (packUn) { packUn[k] = kp; packUR[k] = ur; packZ[k] = z; }      // single-frame call: also into the host-mapped result pack
}

// ---- small RAII device buffer for the stateless entry points ---------------------------------------
struct FBuf {
    void* p = nullptr;
    ~FBuf() { orbx_scratch_free(p); }
    int alloc(size_t bytes) { return orbx_scratch_alloc(&p, bytes); }
};

static int check_cam(const orbx_camera* cam)
{
    if (!cam) { orbx_set_error("null camera"); return ORBX_E_INVALID; }
    if (!(cam->fx > 0) || !(cam->fy > 0)) { orbx_set_error("camera focal lengths must be positive"); return ORBX_E_INVALID; }
    return ORBX_OK;
}

static int launch_cvt(const u8* src, size_t sFrame, size_t sRow, int channels, int rgbOrder, u8* dst, size_t dFrame,
                      size_t dRow, int w, int h, int nf, cudaStream_t st)
{
    const int vecOk = ((reinterpret_cast<uintptr_t>(src) | sFrame | sRow) % (channels == 4 ? 16 : 4) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dst) | dFrame | dRow) % 4 == 0);
    for (int f0 = 0; f0 < nf; f0 += 32768) {
        const int n = std::min(32768, nf - f0);
        cvt_gray_kernel<<<dim3((w + 255) / 256, (h + 3) / 4, n), 256, 0, st>>>(src + (size_t)f0 * sFrame, (long long)sFrame, (long long)sRow, channels,
                                                                                 rgbOrder, dst + (size_t)f0 * dFrame, (long long)dFrame, (long long)dRow, w, h, vecOk);
    }
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

extern "C" {

int orbx_cvt_gray(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t frame_stride,
                  size_t row_stride, int width, int height, int nframes, int on_device, uint8_t* gray,
                  size_t gray_frame_stride, size_t gray_row_stride)
{
    if (!h || !color || !gray) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (channels != 3 && channels != 4) { orbx_set_error("cvtColor to gray takes 3 or 4 channels, got %d", channels); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || nframes <= 0) return ORBX_OK;
    if (row_stride < (size_t)width * channels || gray_row_stride < (size_t)width) { orbx_set_error("row stride smaller than a row"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    if (on_device) return launch_cvt(color, frame_stride, row_stride, channels, rgb_order, gray, gray_frame_stride, gray_row_stride, width, height, nframes, st);
    FBuf dc, dg;
    const size_t cRow = (size_t)width * channels, cFrame = cRow * height, gFrame = (size_t)width * height;
    int rc = dc.alloc(cFrame * nframes); if (rc) return rc;
    rc = dg.alloc(gFrame * nframes); if (rc) return rc;
    ORBX_CUDA(cudaMemcpy2DAsync(dc.p, cRow, color, row_stride, cRow, (size_t)height * nframes, cudaMemcpyHostToDevice, st));
    if (frame_stride != row_stride * height)
        for (int f = 0; f < nframes; ++f)
            ORBX_CUDA(cudaMemcpy2DAsync((u8*)dc.p + f * cFrame, cRow, color + f * frame_stride, row_stride, cRow, height, cudaMemcpyHostToDevice, st));
    rc = launch_cvt((const u8*)dc.p, cFrame, cRow, channels, rgb_order, (u8*)dg.p, gFrame, width, width, height, nframes, st);
    if (rc) return rc;
    for (int f = 0; f < nframes; ++f)
        ORBX_CUDA(cudaMemcpy2DAsync(gray + f * gray_frame_stride, gray_row_stride, (u8*)dg.p + f * gFrame, width, width, height, cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

int orbx_depth_to_float(orbx_handle* h, const uint16_t* depth, size_t row_stride, int width, int height, float factor,
                        int on_device, float* out, size_t out_row_stride)
{
    if (!h || !depth || !out) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const dim3 grid((width + 255) / 256, height);
    if (on_device) {
        depth_float_kernel<<<grid, 256, 0, st>>>(depth, (long long)row_stride, out, (long long)out_row_stride, width, height, factor);
        ORBX_CUDA(cudaGetLastError());
        return ORBX_OK;
    }
    FBuf di, dout;
    int rc = di.alloc((size_t)width * height * 2); if (rc) return rc;
    rc = dout.alloc((size_t)width * height * 4); if (rc) return rc;
    ORBX_CUDA(cudaMemcpy2DAsync(di.p, (size_t)width * 2, depth, row_stride * 2, (size_t)width * 2, height, cudaMemcpyHostToDevice, st));
    depth_float_kernel<<<grid, 256, 0, st>>>((const uint16_t*)di.p, width, (float*)dout.p, width, width, height, factor);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpy2DAsync(out, out_row_stride * 4, dout.p, (size_t)width * 4, (size_t)width * 4, height, cudaMemcpyDeviceToHost, st));
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

int orbx_undistort_points(orbx_handle* h, const float* xy, int n, const orbx_camera* cam, float* xy_out)
{
    if (!h || !xy || !xy_out) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    int rc = check_cam(cam); if (rc) return rc;
    if (n <= 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(h->device));
    FBuf di, dout;
    rc = di.alloc((size_t)n * 8); if (rc) return rc;
    rc = dout.alloc((size_t)n * 8); if (rc) return rc;
    ORBX_CUDA(cudaMemcpyAsync(di.p, xy, (size_t)n * 8, cudaMemcpyHostToDevice, h->stream));
    undistort_kernel<<<(n + 127) / 128, 128, 0, h->stream>>>(cam_double(cam), (const float*)di.p, n, (float*)dout.p);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(xy_out, dout.p, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_image_bounds(orbx_handle* h, int cols, int rows, const orbx_camera* cam, float bounds[4])
{
    if (!h || !bounds) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    int rc = check_cam(cam); if (rc) return rc;
    if (cam->k1 == 0.0f) { bounds[0] = 0.0f; bounds[1] = (float)cols; bounds[2] = 0.0f; bounds[3] = (float)rows; return ORBX_OK; }
    const float c[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
    float u[8];
    rc = orbx_undistort_points(h, c, 4, cam, u); if (rc) return rc;
    bounds[0] = std::min(u[0], u[4]); bounds[1] = std::max(u[2], u[6]);
    bounds[2] = std::min(u[1], u[3]); bounds[3] = std::max(u[5], u[7]);
    return ORBX_OK;
}

int orbx_frames_rgbd(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t color_frame_stride,
                     size_t color_row_stride, const void* depth, int depth_is_u16, float depth_factor,
                     size_t depth_frame_stride, size_t depth_row_stride, int width, int height, int nframes,
                     int on_device, const orbx_camera* cam, int first_slot)
{
    ORBX_NVTX("orbx_frames_rgbd");
    if (!h || !color || !depth) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    int rc = check_cam(cam); if (rc) return rc;
    if (channels != 1 && channels != 3 && channels != 4) { orbx_set_error("colour images have 1, 3 or 4 channels, got %d", channels); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || nframes <= 0) return ORBX_OK;
    if (first_slot < 0 || first_slot + nframes > h->resultSlots) { orbx_set_error("slots [%d,%d) exceed result store (%d)", first_slot, first_slot + nframes, h->resultSlots); return ORBX_E_CAPACITY; }
    const size_t dElem = depth_is_u16 ? 2 : 4;
    if (color_row_stride < (size_t)width * channels || depth_row_stride < (size_t)width * dElem) { orbx_set_error("row stride smaller than a row"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    rc = orbx_ensure_match_buffers(h); if (rc) return rc;
    const size_t R = (size_t)h->resultSlots;
    if (!h->dKpsUn) ORBX_CUDA(cudaMalloc((void**)&h->dKpsUn, R * h->kpCap * sizeof(orbx_keypoint)));
    // ComputeImageBounds once per calibration and image size (mbInitialComputations, Frame.cc:170-186)
    float bounds[4];
    {
        const float key[11] = {cam->fx, cam->fy, cam->cx, cam->cy, cam->k1, cam->k2, cam->p1, cam->p2, cam->k3, (float)width, (float)height};
        if (!h->boundsValid || memcmp(key, h->boundsKey, sizeof(key)) != 0) {
            rc = orbx_image_bounds(h, width, height, cam, bounds); if (rc) return rc;
            memcpy(h->boundsKey, key, sizeof(key));
            for (int i = 0; i < 4; ++i) h->bounds[i] = bounds[i];
            h->boundsValid = true;
        }
        for (int i = 0; i < 4; ++i) bounds[i] = h->bounds[i];
    }
    cudaStream_t st = h->stream;
    const CamD cd = cam_double(cam);
    const int C = h->chunkFrames;
    const size_t gPitch = ((size_t)width + 15) & ~(size_t)15, gFrame = gPitch * height;
    // staging (kept in the handle): gray chunk, and colour / depth chunks when they arrive from the host
    const bool needGray = channels != 1 || !on_device;
    const size_t cRow = (size_t)width * channels, cFrame = cRow * height, zRow = (size_t)width * dElem, zFrame = zRow * height;
    const size_t grayBytes = needGray ? ((gFrame * C + 255) & ~(size_t)255) : 0;
    const size_t colorBytes = (!on_device && channels != 1) ? ((cFrame * C + 255) & ~(size_t)255) : 0;
    const size_t depthBytes = !on_device ? ((zFrame * C + 255) & ~(size_t)255) : 0;
    if (grayBytes + colorBytes + depthBytes > h->frontStageBytes) {
        ORBX_CUDA(cudaStreamSynchronize(st));
        if (h->dFrontStage) cudaFree(h->dFrontStage);
        h->dFrontStage = nullptr; h->frontStageBytes = 0;
        ORBX_CUDA(cudaMalloc((void**)&h->dFrontStage, grayBytes + colorBytes + depthBytes));
        h->frontStageBytes = grayBytes + colorBytes + depthBytes;
    }
    struct { u8* p; } dGray{h->dFrontStage}, dColor{h->dFrontStage + grayBytes}, dDepth{h->dFrontStage + grayBytes + colorBytes};
    const bool denseColor = color_row_stride == cRow && color_frame_stride == cFrame && channels != 1;
    const bool denseDepth = depth_row_stride == zRow && depth_frame_stride == zFrame;
    for (int f0 = 0; f0 < nframes; f0 += C) {
        const int nf = std::min(C, nframes - f0);
        const u8* csrc = color + (size_t)f0 * color_frame_stride;
        const u8* zsrc = (const u8*)depth + (size_t)f0 * depth_frame_stride;
        const u8* gsrc; size_t gfs, grs;
        const u8* zdev = zsrc; size_t zfs = depth_frame_stride, zrs = depth_row_stride;
        if (!on_device) {
            if (denseColor) ORBX_CUDA(cudaMemcpyAsync(dColor.p, csrc, cFrame * nf, cudaMemcpyHostToDevice, st));
            if (denseDepth) ORBX_CUDA(cudaMemcpyAsync(dDepth.p, zsrc, zFrame * nf, cudaMemcpyHostToDevice, st));
            for (int f = 0; f < nf && !(denseColor && denseDepth); ++f) {
                u8* cdst = channels == 1 ? (u8*)dGray.p + f * gFrame : (u8*)dColor.p + f * cFrame;
                if (!denseColor)
                    ORBX_CUDA(cudaMemcpy2DAsync(cdst, channels == 1 ? gPitch : cRow, csrc + f * color_frame_stride, color_row_stride, cRow, height, cudaMemcpyHostToDevice, st));
                if (!denseDepth)
                    ORBX_CUDA(cudaMemcpy2DAsync((u8*)dDepth.p + f * zFrame, zRow, zsrc + f * depth_frame_stride, depth_row_stride, zRow, height, cudaMemcpyHostToDevice, st));
            }
            zdev = (const u8*)dDepth.p; zfs = zFrame; zrs = zRow;
        }
        if (channels == 1) {
            if (on_device) { gsrc = csrc; gfs = color_frame_stride; grs = color_row_stride; }
            else { gsrc = (const u8*)dGray.p; gfs = gFrame; grs = gPitch; }
        } else {
            const u8* cdev = on_device ? csrc : (const u8*)dColor.p;
            rc = launch_cvt(cdev, on_device ? color_frame_stride : cFrame, on_device ? color_row_stride : cRow, channels, rgb_order,
                            (u8*)dGray.p, gFrame, gPitch, width, height, nf, st);
            if (rc) return rc;
            gsrc = (const u8*)dGray.p; gfs = gFrame; grs = gPitch;
        }
        rc = orbx_extract_batch(h, gsrc, gfs, grs, width, height, nf, 1, first_slot + f0);
        if (rc) return rc;
        frame_front_kernel<<<dim3((h->kpCap + 127) / 128, nf), 128, 0, st>>>(cd, cam->bf, h->dKps, h->dCount, h->kpCap, first_slot + f0, zdev,
                                                                              depth_is_u16, depth_factor, depth_factor != 1.0f, (long long)zfs,
                                                                              (long long)zrs, h->dKpsUn, h->dURight, h->dDepth);
        ORBX_CUDA(cudaGetLastError());
    }
    // the store now describes undistorted frames: matchers and the grid use mvKeysUn and these bounds
    h->storeUndistorted = true;
    rc = orbx_build_grids(h, first_slot, nframes, width, height);
    if (rc) return rc;
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

// One RGB-D frame, host images in, the whole Frame front out: what Tracking::GrabImageRGBD + Frame::Frame(RGB-D) produce for one
// image pair (src/Tracking.cc:300-339, src/Frame.cc:136-189).  The latency twin of orbx_frames_rgbd, built like orbx_extract and
// without a single copy command on its path (each small cudaMemcpyAsync costs ~8 us of latency):
//   colour -> pinned, device-MAPPED staging in row slices; cvtColor reads a slice through the mapping while the host copies the next
//   and writes gray straight into the extractor's level-0 buffer -> the frame's captured graph, whose last node writes mvKeys and
//   the descriptors into the mapped result pack; meanwhile the host copies the depth map into the mapped staging; the frame-front
//   kernel (UndistortKeyPoints, ComputeStereoFromRGBD) reads the ~1000 depth pixels it needs through the mapping and writes
//   mvKeysUn / mvuRight / mvDepth into the pack; ONE event synchronisation; the grid for later matcher calls is built behind it.
// Page-locked colour / depth images are read in place.
int orbx_frame_rgbd(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t color_row_stride, const void* depth,
                    int depth_is_u16, float depth_factor, size_t depth_row_stride, int width, int height, const orbx_camera* cam,
                    orbx_keypoint* kps, orbx_keypoint* keys_un, uint8_t* desc, float* u_right, float* depth_out, int capacity, int* n,
                    float bounds_out[4])
{
    ORBX_NVTX("orbx_frame_rgbd");
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || !color) return ORBX_OK;                 // empty image: outputs untouched, as the extractor does
    if (!depth) { orbx_set_error("null depth image"); return ORBX_E_INVALID; }
    int rc = check_cam(cam); if (rc) return rc;
    if (channels != 1 && channels != 3 && channels != 4) { orbx_set_error("colour images have 1, 3 or 4 channels, got %d", channels); return ORBX_E_INVALID; }
    if (width < 64 || height < 64) { orbx_set_error("image smaller than 64x64 is outside the extractor's domain"); return ORBX_E_INVALID; }
    const size_t dElem = depth_is_u16 ? 2 : 4;
    const size_t cRow = (size_t)width * channels, zRow = (size_t)width * dElem;
    if (color_row_stride < cRow || depth_row_stride < zRow) { orbx_set_error("row stride smaller than a row"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    rc = orbx_setup_shape(h, width, height); if (rc) return rc;
    rc = orbx_ensure_match_buffers(h); if (rc) return rc;
    const size_t R = (size_t)h->resultSlots, cap = (size_t)h->kpCap;
    if (!h->dKpsUn) ORBX_CUDA(cudaMalloc((void**)&h->dKpsUn, R * cap * sizeof(orbx_keypoint)));
    {   // ComputeImageBounds once per calibration and image size (mbInitialComputations, Frame.cc:170-186)
        const float key[11] = {cam->fx, cam->fy, cam->cx, cam->cy, cam->k1, cam->k2, cam->p1, cam->p2, cam->k3, (float)width, (float)height};
        if (!h->boundsValid || memcmp(key, h->boundsKey, sizeof(key)) != 0) {
            float b[4];
            rc = orbx_image_bounds(h, width, height, cam, b); if (rc) return rc;
            memcpy(h->boundsKey, key, sizeof(key));
            for (int i = 0; i < 4; ++i) h->bounds[i] = b[i];
            h->boundsValid = true;
        }
    }
    cudaStream_t st = h->stream;
    // ORBX_LAT_TRACE=1: host clock at every phase of the call, averaged over 200 calls, to stderr
    static const bool trace = getenv("ORBX_LAT_TRACE") != nullptr;
    static double tacc[8] = {0}; static int tcalls = 0;
    auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    const double tt0 = trace ? now() : 0.0;
    double tt[8] = {0};
    static cudaEvent_t tev[5] = {nullptr}; static double eacc[4] = {0};
    if (trace && !tev[0]) for (cudaEvent_t& e : tev) cudaEventCreate(&e);
    if (trace) cudaEventRecord(tev[0], st);
    // pinned, device-mapped staging for pageable inputs: [colour or gray][depth].  The kernels read it in place.
    const size_t grayRow = (size_t)h->l0Pitch;
    const size_t offDepth = ((channels == 1 ? grayRow : cRow) * height + 255) & ~(size_t)255;
    rc = orbx_ensure_host_stage(h, offDepth + ((zRow * height + 255) & ~(size_t)255)); if (rc) return rc;
    rc = orbx_ensure_pack(h); if (rc) return rc;
    u8* stage = (u8*)h->hStage;
    // a page-locked caller image: usable in place when the driver hands out a device alias for it
    auto mapped = [](const void* p) -> const u8* {
        cudaPointerAttributes pa;
        const bool host = cudaPointerGetAttributes(&pa, p) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer;
        cudaGetLastError();
        return host ? (const u8*)pa.devicePointer : nullptr;
    };
    // ---- colour in, gray into the extractor's level-0 buffer ------------------------------------------------------------
    if (channels == 1) {
        if (mapped(color)) ORBX_CUDA(cudaMemcpy2DAsync(h->dLevel0, grayRow, color, color_row_stride, width, height, cudaMemcpyHostToDevice, st));
        else {
            for (int y = 0; y < height; ++y) memcpy(stage + (size_t)y * grayRow, color + (size_t)y * color_row_stride, width);
            ORBX_CUDA(cudaMemcpyAsync(h->dLevel0, stage, grayRow * height, cudaMemcpyHostToDevice, st));
        }
    } else {
        // colour crosses the bus by DMA (a kernel reading 0.9 MB through the mapping in 4-byte pieces measured ~3x slower than
        // the copy engine), cvtColor then runs out of device memory
        const size_t colorDev = (cRow * height + 255) & ~(size_t)255;
        if (colorDev > h->frontStageBytes) {
            ORBX_CUDA(cudaStreamSynchronize(st));
            if (h->dFrontStage) cudaFree(h->dFrontStage);
            h->dFrontStage = nullptr; h->frontStageBytes = 0;
            ORBX_CUDA(cudaMalloc((void**)&h->dFrontStage, colorDev));
            h->frontStageBytes = colorDev;
        }
        u8* dColor = h->dFrontStage;
        if (mapped(color)) {
            ORBX_CUDA(cudaMemcpy2DAsync(dColor, cRow, color, color_row_stride, cRow, height, cudaMemcpyHostToDevice, st));
            rc = launch_cvt(dColor, cRow * height, cRow, channels, rgb_order, h->dLevel0, (size_t)h->l0FrameBytes, grayRow, width, height, 1, st);
            if (rc) return rc;
        } else {
            // pageable image: the host copies it into the mapped staging in row slices and cvtColor converts a slice THROUGH the
            // mapping while the next one is being copied (measured: 301 us per call against 328-360 us with a DMA per slice; a
            // 640x480 RGB image is 0.9 MB: ~70 us of memcpy that would otherwise all sit in front of the GPU)
            static const int slices = getenv("ORBX_RGBD_SLICES") ? std::max(1, atoi(getenv("ORBX_RGBD_SLICES"))) : 4;
            const int rowsPer = ((height + slices - 1) / slices + 3) & ~3;
            for (int y0 = 0; y0 < height; y0 += rowsPer) {
                const int rows = std::min(rowsPer, height - y0);
                if (color_row_stride == cRow) memcpy(stage + (size_t)y0 * cRow, color + (size_t)y0 * cRow, cRow * rows);
                else for (int y = y0; y < y0 + rows; ++y) memcpy(stage + (size_t)y * cRow, color + (size_t)y * color_row_stride, cRow);
                rc = launch_cvt(stage + (size_t)y0 * cRow, cRow * height, cRow, channels, rgb_order, h->dLevel0 + (size_t)y0 * grayRow,
                                (size_t)h->l0FrameBytes, grayRow, width, rows, 1, st);
                if (rc) return rc;
            }
        }
    }
    if (trace) { tt[0] = now(); cudaEventRecord(tev[1], st); }
    h->storeUndistorted = false;
    rc = orbx_launch_single_frame(h);                    // the extractor's graph: results in slot 0, mvKeys + descriptors in the pack
    if (rc) return rc;
    if (trace) { tt[1] = now(); cudaEventRecord(tev[2], st); }
    // ---- depth: only the pixels under the keypoints are read, through the mapping (the host copy of a pageable depth map
    // overlaps the extractor) --------------------------------------------------------------------------------------------
    const u8* dz = mapped(depth);
    size_t dzRow = depth_row_stride;
    if (!dz) {
        if (depth_row_stride == zRow) memcpy(stage + offDepth, depth, zRow * height);
        else for (int y = 0; y < height; ++y) memcpy(stage + offDepth + (size_t)y * zRow, (const u8*)depth + (size_t)y * depth_row_stride, zRow);
        dz = stage + offDepth; dzRow = zRow;
    }
    if (trace) tt[2] = now();
    const OrbxPackLayout PL = orbx_pack_layout(h->kpCap);
    frame_front_kernel<<<dim3((unsigned)((cap + 127) / 128), 1), 128, 0, st>>>(cam_double(cam), cam->bf, h->dKps, h->dCount, h->kpCap, 0, dz, depth_is_u16,
                                                                              depth_factor, depth_factor != 1.0f, (long long)(dzRow * height), (long long)dzRow,
                                                                              h->dKpsUn, h->dURight, h->dDepth, (orbx_keypoint*)(h->hPack + PL.kun),
                                                                              (float*)(h->hPack + PL.ur), (float*)(h->hPack + PL.z));
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaEventRecord(h->evPack, st));           // everything the caller gets is in the pack here
    if (trace) cudaEventRecord(tev[3], st);
    h->storeUndistorted = true;                          // the store now describes an undistorted frame (grid, matchers)
    rc = orbx_build_grids(h, 0, 1, width, height);       // for the matchers that run on slot 0 next: finishes behind the caller's back
    if (rc) return rc;
    if (trace) tt[3] = now();
    ORBX_CUDA(cudaEventSynchronize(h->evPack));
    if (trace) tt[4] = now();
    const int32_t* hdr = (const int32_t*)h->hPack;
    if (hdr[1]) {
        ORBX_CUDA(cudaMemsetAsync(h->dStatus, 0, sizeof(int), st));
        orbx_set_error("device-side capacity overflow (flags 0x%x): quadtree node/keypoint table too small", hdr[1]);
        return ORBX_E_CAPACITY;
    }
    const int cnt = hdr[0];
    if ((kps || keys_un || desc || u_right || depth_out) && cnt > capacity) { orbx_set_error("capacity %d < %d keypoints", capacity, cnt); return ORBX_E_CAPACITY; }
    if (kps && cnt > 0) memcpy(kps, h->hPack + PL.kps, (size_t)cnt * sizeof(orbx_keypoint));
    if (keys_un && cnt > 0) memcpy(keys_un, h->hPack + PL.kun, (size_t)cnt * sizeof(orbx_keypoint));
    if (desc && cnt > 0) memcpy(desc, h->hPack + PL.desc, (size_t)cnt * 32);
    if (u_right && cnt > 0) memcpy(u_right, h->hPack + PL.ur, (size_t)cnt * 4);
    if (depth_out && cnt > 0) memcpy(depth_out, h->hPack + PL.z, (size_t)cnt * 4);
    if (bounds_out) for (int i = 0; i < 4; ++i) bounds_out[i] = h->bounds[i];
    if (n) *n = cnt;
    if (trace) {
        tt[5] = now();
        cudaEventRecord(tev[4], st); cudaEventSynchronize(tev[4]);
        for (int i = 0; i < 4; ++i) { float ms = 0; cudaEventElapsedTime(&ms, tev[i], tev[i + 1]); eacc[i] += ms * 1e3; }
        for (int i = 0; i < 6; ++i) tacc[i] += tt[i] - tt0;
        if (++tcalls == 200) {
            fprintf(stderr, "orbx_frame_rgbd trace (us from entry, mean of 200): colour enqueued %.1f, graph launched %.1f, depth staged %.1f, "
                            "front+grid launched %.1f, synchronised %.1f, outputs copied %.1f\n",
                    tacc[0] / 200, tacc[1] / 200, tacc[2] / 200, tacc[3] / 200, tacc[4] / 200, tacc[5] / 200);
            fprintf(stderr, "orbx_frame_rgbd trace, device (us, CUDA events): colour in + cvtColor %.1f, extractor graph + pack %.1f, frame front %.1f, grid %.1f\n",
                    eacc[0] / 200, eacc[1] / 200, eacc[2] / 200, eacc[3] / 200);
            tcalls = 0; for (double& a : tacc) a = 0; for (double& a : eacc) a = 0;
        }
    }
    return ORBX_OK;
}

int orbx_download_frame_front(orbx_handle* h, int first_slot, int nframes, orbx_keypoint* keys_un, float* u_right,
                              float* depth, int capacity, float bounds[4])
{
    if (!h) { orbx_set_error("null handle"); return ORBX_E_INVALID; }
    if (nframes <= 0) return ORBX_OK;
    if (!h->dKpsUn || !h->storeUndistorted) { orbx_set_error("no RGB-D frames in the store"); return ORBX_E_STATE; }
    if (first_slot < 0 || first_slot + nframes > h->resultSlots) { orbx_set_error("slot range out of bounds"); return ORBX_E_INVALID; }
    if ((keys_un || u_right || depth) && capacity < h->kpCap) { orbx_set_error("output capacity %d < %d", capacity, h->kpCap); return ORBX_E_CAPACITY; }
    ORBX_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const size_t cap = (size_t)h->kpCap;
    if (keys_un)
        ORBX_CUDA(cudaMemcpy2DAsync(keys_un, (size_t)capacity * sizeof(orbx_keypoint), h->dKpsUn + (size_t)first_slot * cap, cap * sizeof(orbx_keypoint),
                                    cap * sizeof(orbx_keypoint), nframes, cudaMemcpyDeviceToHost, st));
    if (u_right)
        ORBX_CUDA(cudaMemcpy2DAsync(u_right, (size_t)capacity * 4, h->dURight + (size_t)first_slot * cap, cap * 4, cap * 4, nframes, cudaMemcpyDeviceToHost, st));
    if (depth)
        ORBX_CUDA(cudaMemcpy2DAsync(depth, (size_t)capacity * 4, h->dDepth + (size_t)first_slot * cap, cap * 4, cap * 4, nframes, cudaMemcpyDeviceToHost, st));
    if (bounds) for (int i = 0; i < 4; ++i) bounds[i] = h->bounds[i];
    ORBX_CUDA(cudaStreamSynchronize(st));
    return ORBX_OK;
}

} // extern "C"
