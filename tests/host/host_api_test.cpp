// Exercises the host-side C++ classes (orb_slam2_map_b200/host/orbx.hpp) the way Frame.cc /
// Tracking.cc use the reference classes.  usage: host_api_test <in.raw> <out.bin>
//   in.raw : int32 w, h, nframes(=2) followed by nframes*w*h bytes
//   out.bin: for each frame: int32 n, n*28 bytes keypoints, n*32 bytes descriptors; then the pyramid
//            level 3 of frame 1 with border (int32 rows, cols, bytes); then int32 nmatches,
//            int32 n1, n1*int32 vnMatches12, n1*2 floats vbPrevMatched
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
    } catch (const std::exception& e) {
        fprintf(stderr, "host_api_test: %s\n", e.what());
        fclose(o);
        return 1;
    }
    fclose(o);
    return 0;
}
