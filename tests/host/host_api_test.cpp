// Exercises the host-side C++ classes (orb_slam2_map_b200/host/orbx.hpp) the way Frame.cc /
// Tracking.cc use the reference classes.  usage: host_api_test <in.raw> <out.bin>
//   in.raw : int32 w, h, nframes(=2) followed by nframes*w*h bytes
//   out.bin: for each frame: int32 n, n*28 bytes keypoints, n*32 bytes descriptors; then the pyramid
//            level 3 of frame 1 with border (int32 rows, cols, bytes); then int32 nmatches,
//            int32 n1, n1*int32 vnMatches12, n1*2 floats vbPrevMatched; then the keyframe-rate section:
//            int32 nq, nq*int32 Fuse best index (frame 1 as keyframe, its own keypoints of frame 0 as queries);
//            int32 npoints, npoints*int32 distinctive-descriptor index (groups of 5 descriptors of frame 0);
//            int32 ncloud, ncloud*16 bytes point cloud of a synthetic depth map (= pixel value / 32 of frame 0)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "orbx.hpp"

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 3;
    int32_t hdr[3];
    if (fread(hdr, 4, 3, f) != 3) return 3;
    const int w = hdr[0], h = hdr[1], nf = hdr[2];
    std::vector<uint8_t> img((size_t)w * h * nf);
    if (fread(img.data(), 1, img.size(), f) != img.size()) return 3;
    fclose(f);
    FILE* o = fopen(argv[2], "wb");
    try {
        orbx::ORBextractor extractor(1000, 1.2f, 8, 20, 7, w, h);          // Tracking.cc:201 / TUM1.yaml
        std::vector<std::vector<orbx::KeyPoint>> keys(nf);
        std::vector<std::vector<uint8_t>> desc(nf);
        for (int i = 0; i < nf; ++i) {
            orbx::Image im{img.data() + (size_t)i * w * h, h, w, (size_t)w};
            extractor(im, orbx::Image(), keys[i], desc[i]);                  // Frame::ExtractORB (Frame.cc:250)
            int32_t n = (int32_t)keys[i].size();
            fwrite(&n, 4, 1, o);
            fwrite(keys[i].data(), 28, n, o);
            fwrite(desc[i].data(), 32, n, o);
        }
        // empty image: silent return, outputs untouched (ORBextractor.cc:1046)
        std::vector<orbx::KeyPoint> k0 = keys[0];
        std::vector<uint8_t> d0 = desc[0];
        extractor(orbx::Image(), orbx::Image(), k0, d0);
        if (k0.size() != keys[0].size()) return 4;
        // mvImagePyramid of the LAST image (read by Frame::ComputeStereoMatches in the reference)
        extractor(orbx::Image{img.data() + (size_t)(nf - 1) * w * h, h, w, (size_t)w}, orbx::Image(), k0, d0);
        const std::vector<orbx::OwnedImage>& pyr = extractor.FetchPyramid();
        int32_t dims[2] = {pyr[3].rows + 38, pyr[3].cols + 38};
        fwrite(dims, 4, 2, o);
        fwrite(pyr[3].buf.data(), 1, pyr[3].buf.size(), o);
        if (extractor.GetLevels() != 8 || extractor.GetScaleFactors().size() != 8) return 5;
        // Tracking::MonocularInitialization (Tracking.cc:862-877)
        orbx::FrameView F1, F2;
        F1.mvKeysUn = &keys[0]; F1.mDescriptors = desc[0].data(); F1.mnMaxX = (float)w; F1.mnMaxY = (float)h;
        F2.mvKeysUn = &keys[1]; F2.mDescriptors = desc[1].data(); F2.mnMaxX = (float)w; F2.mnMaxY = (float)h;
        std::vector<orbx::Point2f> vbPrevMatched(keys[0].size());
        for (size_t i = 0; i < keys[0].size(); ++i) vbPrevMatched[i] = orbx::Point2f{keys[0][i].x, keys[0][i].y};
        std::vector<int> vnMatches12;
        orbx::ORBmatcher matcher(0.9f, true);
        int32_t nmatches = matcher.SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, 100);
        int32_t n1 = (int32_t)vnMatches12.size();
        fwrite(&nmatches, 4, 1, o);
        fwrite(&n1, 4, 1, o);
        fwrite(vnMatches12.data(), 4, n1, o);
        fwrite(vbPrevMatched.data(), 8, n1, o);
        if (orbx::ORBmatcher::DescriptorDistance(desc[0].data(), desc[0].data()) != 0) return 6;
        // LocalMapping::SearchInNeighbors-style Fuse: keypoints of frame 0 as already projected MapPoints, searched in
        // frame 1 used as a keyframe (th = 3)
        std::vector<float> scale = extractor.GetScaleFactors(), invSigma2 = extractor.GetInverseScaleSigmaSquares(),
                           sigma2 = extractor.GetScaleSigmaSquares();
        orbx::KeyFrameView KF;
        KF.frame = F2; KF.frame.mvScaleFactors = &scale; KF.mvLevelSigma2 = &sigma2; KF.mvInvLevelSigma2 = &invSigma2;
        orbx::ProjectionQueries Q;
        for (size_t i = 0; i < keys[0].size(); ++i) {
            orbm_proj_query q{};
            q.u = keys[0][i].x; q.v = keys[0][i].y; q.radius = 3.0f * scale[keys[0][i].octave];
            q.min_level = -1; q.max_level = -1; q.u_right = -1.f; q.max_right_err = -1.f; q.angle = keys[0][i].angle;
            q.level_lo = keys[0][i].octave - 1; q.level_hi = keys[0][i].octave;
            Q.q.push_back(q);
        }
        Q.descriptors = desc[0];
        const std::vector<int32_t> fused = matcher.Fuse(ORBM_FUSE_MAPPOINTS, KF, Q);
        int32_t nq = (int32_t)fused.size();
        fwrite(&nq, 4, 1, o);
        fwrite(fused.data(), 4, nq, o);
        // MapPoint::ComputeDistinctiveDescriptors on groups of 5 consecutive descriptors
        std::vector<int32_t> offsets;
        for (int32_t i = 0; i + 5 <= (int32_t)keys[0].size() && offsets.size() < 101; i += 5) offsets.push_back(i);
        const std::vector<int32_t> dd = orbx::ComputeDistinctiveDescriptors(desc[0].data(), offsets);
        int32_t np = (int32_t)dd.size();
        fwrite(&np, 4, 1, o);
        fwrite(dd.data(), 4, np, o);
        // PointCloudMapping::generatePointCloud on depth = pixel / 32 (so part of the map is below 0.01 or 0)
        std::vector<float> depth((size_t)w * h);
        for (size_t i = 0; i < depth.size(); ++i) depth[i] = (float)img[i] / 32.0f;
        const double Tinv[16] = {0.36, 0.48, -0.8, 0.1, -0.8, 0.6, 0, -0.2, 0.48, 0.64, 0.6, 0.3, 0, 0, 0, 1};
        const std::vector<orbx_cloud_point> cloud = orbx::GeneratePointCloud(depth.data(), (size_t)w, nullptr, 0, w, h, 517.3f, 516.5f,
                                                                            318.6f, 255.3f, Tinv);
        int32_t nc = (int32_t)cloud.size();
        fwrite(&nc, 4, 1, o);
        fwrite(cloud.data(), 16, nc, o);
        // the rest of PointCloudMapping's chain: voxel.filter at resolution 0.04, then sor.filter (mean_k 50, multiplier 1.0)
        const std::vector<orbx_cloud_point> vox = orbx::VoxelGridFilter(cloud, 0.04f);
        const std::vector<orbx_cloud_point> kept = orbx::StatisticalOutlierFilter(vox, 50, 1.0);
        int32_t nv = (int32_t)vox.size(), nk = (int32_t)kept.size();
        fwrite(&nv, 4, 1, o);
        fwrite(vox.data(), 16, nv, o);
        fwrite(&nk, 4, 1, o);
        fwrite(kept.data(), 16, nk, o);
    } catch (const std::exception& e) {
        fprintf(stderr, "host_api_test: %s\n", e.what());
        fclose(o);
        return 1;
    }
    fclose(o);
    return 0;
}
