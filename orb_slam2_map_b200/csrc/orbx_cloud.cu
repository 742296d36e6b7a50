// orbx_cloud.cu — this fork's dense mapping step as an sm_100a kernel behind the C ABI:
//
//   K22 backproject_kernel   PointCloudMapping::generatePointCloud / convertToPointCloud (src/PointCloudMap.cc:78-139)
//                            fused with pcl::transformPointCloud<PointT, double>(tmp, cloud, T.inverse().matrix())
//
// One CTA per keyframe.  The reference's push_back order is the row-major scan order of the sampled pixels, so the
// output position of a point is an exclusive count of the valid samples before it: pass 1 = one warp per sampled row
// counts by ballot, a block scan turns the row counts into row offsets, pass 2 re-reads the row (L1/L2 hits) and each
// lane writes its 16-byte record at offset + popc(ballot below the lane).  HBM-bound: a third of the depth rows are
// read (every third sector of them is needed in full: a 32-byte sector holds 8 floats, the stride is 3), plus
// 16 bytes written per point.

#include "orbx_internal.h"

//   K23 pack_records_kernel  the per-feature record of Map::Save's keyframe block (src/Map.cc:160-180): 16 lanes per
//                            feature, one 4-byte word each, so the 64-byte records leave as coalesced stores

#include <algorithm>
#include <vector>

#define BP_THREADS 1024
#define BP_MAX_ROWS 4096
#define FULLMASK 0xffffffffu

struct BackprojectArgs {
    const float* depth; size_t dRow, dFrame;
    const uint8_t* bgr; size_t cRow, cFrame;
    int width, height, step, capacity;
    float fx, fy, cx, cy;
    const double* poseInv;              // nframes x 16 or nullptr
};

__device__ __forceinline__ bool bp_valid(float d) { return !((double)d < 0.01 || d > 10.f); }

__global__ void __launch_bounds__(BP_THREADS)
backproject_kernel(BackprojectArgs A, orbx_cloud_point* __restrict__ points, int* __restrict__ counts)
{
    __shared__ int rowOff[BP_MAX_ROWS + 1];
    __shared__ int warpSum[BP_THREADS / 32];
    __shared__ double T[12];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float* depth = A.depth + (size_t)f * A.dFrame;
    const uint8_t* bgr = A.bgr ? A.bgr + (size_t)f * A.cFrame : nullptr;
    const int nrows = (A.height + A.step - 1) / A.step, ncols = (A.width + A.step - 1) / A.step;
    if (A.poseInv && tid < 12) T[tid] = A.poseInv[(size_t)f * 16 + tid];
    // pass 1: valid samples per sampled row
    for (int r = wid; r < nrows; r += BP_THREADS / 32) {
        const float* row = depth + (size_t)(r * A.step) * A.dRow;
        int c = 0;
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane;
            const bool v = j < ncols && bp_valid(row[j * A.step]);
            c += __popc(__ballot_sync(FULLMASK, v));
        }
        if (lane == 0) rowOff[r + 1] = c;
    }
    if (tid == 0) rowOff[0] = 0;
    __syncthreads();
    // inclusive scan of rowOff[1..nrows] (chunks of BP_THREADS, carry in a register)
    int carry = 0;
    for (int base = 0; base < nrows; base += BP_THREADS) {
        const int i = base + tid;
        int v = i < nrows ? rowOff[i + 1] : 0;
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULLMASK, v, d); if (lane >= d) v += o; }
        if (lane == 31) warpSum[wid] = v;
        __syncthreads();
        if (wid == 0) {
            int w = warpSum[lane];
            for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULLMASK, w, d); if (lane >= d) w += o; }
            warpSum[lane] = w;
        }
        __syncthreads();
        const int before = carry + (wid ? warpSum[wid - 1] : 0);
        if (i < nrows) rowOff[i + 1] = v + before;
        carry += warpSum[BP_THREADS / 32 - 1];
        __syncthreads();
    }
    const int total = rowOff[nrows];
    if (tid == 0) counts[f] = total;
    if (total > A.capacity) return;                     // the host reports ORBX_E_CAPACITY
    orbx_cloud_point* out = points + (size_t)f * A.capacity;
    // pass 2: ordered write
    for (int r = wid; r < nrows; r += BP_THREADS / 32) {
        const int m = r * A.step;
        const float* row = depth + (size_t)m * A.dRow;
        const uint8_t* crow = bgr ? bgr + (size_t)m * A.cRow : nullptr;
        int off = rowOff[r];
        for (int j0 = 0; j0 < ncols; j0 += 32) {
            const int j = j0 + lane, n = j * A.step;
            const float d = j < ncols ? row[n] : 0.f;
            const bool v = j < ncols && bp_valid(d);
            const unsigned mask = __ballot_sync(FULLMASK, v);
            if (v) {
                const float x = __fdiv_rn(__fmul_rn(__fsub_rn((float)n, A.cx), d), A.fx);
                const float y = __fdiv_rn(__fmul_rn(__fsub_rn((float)m, A.cy), d), A.fy);
                float ox = x, oy = y, oz = d;
                if (A.poseInv) {
                    const double X = x, Y = y, Z = d;
                    ox = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[0], X), __dmul_rn(T[1], Y)), __dmul_rn(T[2], Z)), T[3]);
                    oy = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[4], X), __dmul_rn(T[5], Y)), __dmul_rn(T[6], Z)), T[7]);
                    oz = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[8], X), __dmul_rn(T[9], Y)), __dmul_rn(T[10], Z)), T[11]);
                }
                unsigned col = 0xff000000u;             // a = 255, b = g = r = 0
                if (crow) col |= (unsigned)crow[n * 3] | ((unsigned)crow[n * 3 + 1] << 8) | ((unsigned)crow[n * 3 + 2] << 16);
                float4 rec;
                rec.x = ox; rec.y = oy; rec.z = oz; rec.w = __uint_as_float(col);
                *reinterpret_cast<float4*>(out + off + __popc(mask & ((1u << lane) - 1))) = rec;
            }
            off += __popc(mask);
        }
    }
}

struct CBuf {
    void* p = nullptr;
    ~CBuf() { orbx_scratch_free(p); }
    int alloc(size_t bytes) { return orbx_scratch_alloc(&p, bytes); }
    template <typename T> T* as() { return (T*)p; }
};

extern "C" int orbm_generate_point_clouds(const float* depth, size_t depth_row_stride, size_t depth_frame_stride, const uint8_t* bgr,
                                          size_t bgr_row_stride, size_t bgr_frame_stride, int width, int height, int nframes,
                                          int step, const orbx_camera* cam, const double* pose_inv, int device,
                                          orbx_cloud_point* points, int capacity, int32_t* counts)
{
    static_assert(sizeof(orbx_cloud_point) == 16, "orbx_cloud_point is a 16-byte record");
    if (nframes <= 0) return ORBX_OK;
    if (!depth || !cam || !points || !counts) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (width <= 0 || height <= 0 || step <= 0 || capacity <= 0 || depth_row_stride < (size_t)width ||
        (bgr && bgr_row_stride < (size_t)width * 3)) { orbx_set_error("bad geometry"); return ORBX_E_INVALID; }
    if ((height + step - 1) / step > BP_MAX_ROWS) { orbx_set_error("more than %d sampled rows", BP_MAX_ROWS); return ORBX_E_CAPACITY; }
    if (!(cam->fx != 0.f) || !(cam->fy != 0.f)) { orbx_set_error("bad camera"); return ORBX_E_INVALID; }
    int ndev = 0;
    ORBX_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { orbx_set_error("CUDA device %d not available (%d devices); no CPU fallback", device, ndev); return ORBX_E_CUDA; }
    ORBX_CUDA(cudaSetDevice(device));
    // dense device copies: depth rows of `width` floats, colour rows of 3*width bytes
    CBuf dDepth, dBgr, dPose, dPts, dCnt;
    int rc;
    const size_t dFrame = (size_t)width * height, cFrame = (size_t)width * 3 * height;
    if ((rc = dDepth.alloc(dFrame * 4 * nframes)) || (rc = dPts.alloc((size_t)nframes * capacity * sizeof(orbx_cloud_point))) ||
        (rc = dCnt.alloc((size_t)nframes * 4))) return rc;
    for (int f = 0; f < nframes; ++f)
        ORBX_CUDA(cudaMemcpy2D(dDepth.as<float>() + f * dFrame, (size_t)width * 4, depth + f * depth_frame_stride, depth_row_stride * 4,
                               (size_t)width * 4, height, cudaMemcpyHostToDevice));
    if (bgr) {
        rc = dBgr.alloc(cFrame * nframes); if (rc) return rc;
        for (int f = 0; f < nframes; ++f)
            ORBX_CUDA(cudaMemcpy2D(dBgr.as<uint8_t>() + f * cFrame, (size_t)width * 3, bgr + f * bgr_frame_stride, bgr_row_stride,
                                   (size_t)width * 3, height, cudaMemcpyHostToDevice));
    }
    if (pose_inv) {
        rc = dPose.alloc((size_t)nframes * 16 * 8); if (rc) return rc;
        ORBX_CUDA(cudaMemcpy(dPose.p, pose_inv, (size_t)nframes * 16 * 8, cudaMemcpyHostToDevice));
    }
    BackprojectArgs A;
    A.depth = dDepth.as<float>(); A.dRow = (size_t)width; A.dFrame = dFrame;
    A.bgr = bgr ? dBgr.as<uint8_t>() : nullptr; A.cRow = (size_t)width * 3; A.cFrame = cFrame;
    A.width = width; A.height = height; A.step = step; A.capacity = capacity;
    A.fx = cam->fx; A.fy = cam->fy; A.cx = cam->cx; A.cy = cam->cy;
    A.poseInv = pose_inv ? dPose.as<double>() : nullptr;
    backproject_kernel<<<nframes, BP_THREADS>>>(A, dPts.as<orbx_cloud_point>(), dCnt.as<int>());
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpy(counts, dCnt.p, (size_t)nframes * 4, cudaMemcpyDeviceToHost));
    for (int f = 0; f < nframes; ++f) {
        if (counts[f] > capacity) { orbx_set_error("frame %d yields %d points, capacity %d", f, counts[f], capacity); return ORBX_E_CAPACITY; }
        if (counts[f] > 0)
            ORBX_CUDA(cudaMemcpy(points + (size_t)f * capacity, dPts.as<orbx_cloud_point>() + (size_t)f * capacity,
                                 (size_t)counts[f] * sizeof(orbx_cloud_point), cudaMemcpyDeviceToHost));
    }
    return ORBX_OK;
}

__global__ void pack_records_kernel(const orbx_keypoint* __restrict__ kps, const uint8_t* __restrict__ desc,
                                    const unsigned long long* __restrict__ mpIndex, int n, uint32_t* __restrict__ out)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x, i = t >> 4, w = t & 15;
    if (i >= n) return;
    uint32_t v;
    if (w < 5) v = reinterpret_cast<const uint32_t*>(kps + i)[w];                  // pt.x, pt.y, size, angle, response
    else if (w == 5) v = (uint32_t)kps[i].octave;
    else if (w < 14) v = reinterpret_cast<const uint32_t*>(desc + (size_t)i * 32)[w - 6];
    else { const unsigned long long m = mpIndex ? mpIndex[i] : ~0ull; v = w == 14 ? (uint32_t)m : (uint32_t)(m >> 32); }
    out[t] = v;
}

extern "C" int orbx_pack_map_records(orbx_handle* h, int slot, const uint64_t* mp_index, uint8_t* out, int capacity, int32_t* n)
{
    if (!h || !out || !n) { orbx_set_error("null argument"); return ORBX_E_INVALID; }
    if (slot < 0 || slot >= h->resultSlots) { orbx_set_error("slot out of bounds"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(h->device));
    int cnt = 0;
    ORBX_CUDA(cudaMemcpyAsync(&cnt, h->dCount + slot, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    int rc = orbx_sync(h); if (rc) return rc;
    cnt = std::min(cnt, h->kpCap);
    *n = cnt;
    if (cnt > capacity) { orbx_set_error("capacity %d < %d records", capacity, cnt); return ORBX_E_CAPACITY; }
    if (cnt == 0) return ORBX_OK;
    CBuf dOut, dIdx;
    if ((rc = dOut.alloc((size_t)cnt * 64))) return rc;
    if (mp_index) {
        if ((rc = dIdx.alloc((size_t)cnt * 8))) return rc;
        ORBX_CUDA(cudaMemcpyAsync(dIdx.p, mp_index, (size_t)cnt * 8, cudaMemcpyHostToDevice, h->stream));
    }
    pack_records_kernel<<<(cnt * 16 + 255) / 256, 256, 0, h->stream>>>(h->dKps + (size_t)slot * h->kpCap, h->dDesc + (size_t)slot * h->kpCap * 32,
                                                                      mp_index ? dIdx.as<unsigned long long>() : nullptr, cnt, dOut.as<uint32_t>());
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(out, dOut.p, (size_t)cnt * 64, cudaMemcpyDeviceToHost, h->stream));
    return orbx_sync(h);
}
