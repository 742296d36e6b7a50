// ref_driver.cpp — C entry points around the UNMODIFIED reference ORBextractor
// (/root/reference/src/ORBextractor.cc, compiled where it lies by oracle/Makefile over cvshim.h).
//
// TEST INFRASTRUCTURE ONLY.  Two uses: (1) tests cross-check the oracle's restatement of
// ORBextractor::operator() against the reference's own control flow; (2) bench.py's
// `--impl reference` / cpu_baseline legs time it ("kind": "reference").
//
// Determinism: DistributeOctTree sorts nodes by (size, heap address) (src/ORBextractor.cc:684),
// so its result depends on the allocator.  Inside ref_* calls `operator new` is a thread-local
// bump allocator: addresses grow with creation order and are never reused, which makes the
// reference's own tie-break equal to the oracle's documented rule (newest node of equal size is
// split first).  This is one legitimate execution of the reference, not a modification of it.

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>
#include <sys/mman.h>

#include "cvshim.h"
#include "ORBextractor.h"

namespace {

std::atomic<int> g_arena_overflow(0);

// One lazily-committed 1 GiB virtual range per thread: addresses only ever grow inside a call.
struct Arena {
    char* base = nullptr;
    size_t bytes = size_t(1) << 30, off = 0;
    bool active = false;
    ~Arena() { if (base) munmap(base, bytes); }
    void* alloc(size_t n)
    {
        n = (n + 15) & ~size_t(15);
        if (!base) {
            void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            if (p == MAP_FAILED) return nullptr;
            base = (char*)p;
        }
        if (off + n > bytes) { g_arena_overflow.store(1); return nullptr; }   // falls back to malloc: order no longer guaranteed
        void* p = base + off;
        off += n;
        return p;
    }
    bool owns(const void* p) const { return base && (const char*)p >= base && (const char*)p < base + bytes; }
    void reset() { off = 0; }
};

thread_local Arena t_arena;

// all arenas ever created stay alive until their thread exits; delete() of an arena pointer from
// another thread cannot happen because every ref_* call frees what it allocated on its own thread
struct Scope {
    Scope() { t_arena.reset(); t_arena.active = true; }
    ~Scope() { t_arena.active = false; }
};

} // namespace

void* operator new(size_t n)
{
    if (t_arena.active) { if (void* p = t_arena.alloc(n)) return p; }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !t_arena.owns(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

namespace {

int extract_one(ORB_SLAM2::ORBextractor& ex, const uint8_t* img, int w, int h, size_t stride,
                orbx_keypoint* kps, uint8_t* desc, int cap)
{
    Scope scope;
    int n = 0;
    {
        cv::Mat image(h, w, CV_8UC1, (void*)img, stride);
        std::vector<cv::KeyPoint> keys;
        cv::Mat descriptors;
        ex(image, cv::Mat(), keys, descriptors);
        n = (int)keys.size();
        if (n > cap) return -1;
        for (int i = 0; i < n; ++i) {
            const cv::KeyPoint& k = keys[i];
            if (kps) kps[i] = orbx_keypoint{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
            if (desc) std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
        }
    }
    return n;
}

} // namespace

extern "C" {

struct ref_extractor { ORB_SLAM2::ORBextractor ex; ref_extractor(const orbx_params& p) : ex(p.nfeatures, p.scale_factor, p.nlevels, p.ini_th_fast, p.min_th_fast) {} };

// 1 if any call ran out of arena (allocation order then no longer equals address order)
int ref_arena_overflowed(void) { return g_arena_overflow.load(); }

ref_extractor* ref_extractor_create(const orbx_params* p) { return new ref_extractor(*p); }
void ref_extractor_destroy(ref_extractor* e) { delete e; }

// ORBextractor::operator(); returns keypoint count or -1 if cap is too small
int ref_extract(ref_extractor* e, const uint8_t* img, int w, int h, size_t stride, orbx_keypoint* kps,
                uint8_t* desc, int cap)
{
    return extract_one(e->ex, img, w, h, stride, kps, desc, cap);
}

// mvImagePyramid[level] including its 19-px border (include/ORBextractor.h:85)
int ref_get_level(ref_extractor* e, int level, int with_border, uint8_t* dst, size_t dstride, int* w, int* h)
{
    if (level < 0 || level >= (int)e->ex.mvImagePyramid.size()) return -1;
    const cv::Mat& m = e->ex.mvImagePyramid[level];
    if (m.empty()) return -1;
    const int b = with_border ? 19 : 0;
    if (w) *w = m.cols + 2 * b;
    if (h) *h = m.rows + 2 * b;
    if (dst)
        for (int y = -b; y < m.rows + b; ++y)
            std::memcpy(dst + (size_t)(y + b) * dstride, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - b, m.cols + 2 * b);
    return 0;
}

void ref_get_tables(ref_extractor* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2)
{
    const std::vector<float> a = e->ex.GetScaleFactors(), b = e->ex.GetInverseScaleFactors();
    const std::vector<float> c = e->ex.GetScaleSigmaSquares(), d = e->ex.GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < a.size(); ++i) { scale[i] = a[i]; inv_scale[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i]; }
}

// CPU baseline, BASELINE.json config 2: the reference extractor on every frame (one extractor
// object per thread, as the reference uses one per eye) + SearchForInitialization on consecutive
// pairs (ORBmatcher.cc cannot be compiled without PCL/DBoW2/g2o headers, so the matching half is
// the oracle's restatement).  Same signature as oracle_extract_match_many.
int64_t ref_extract_match_many(const orbx_params* p, const uint8_t* imgs, size_t frame_stride, int w, int h,
                               size_t stride, int nframes, int nthreads, int window_size, float nnratio,
                               int check_orientation, int64_t* total_matches, uint64_t* ncompares)
{
    if (nthreads < 1) nthreads = 1;
    const int cap = p->nfeatures + 64 * p->nlevels + 64;
    std::vector<std::vector<orbx_keypoint>> K(nframes);
    std::vector<std::vector<uint8_t>> D(nframes);
    std::atomic<int> next(0);
    auto extract_work = [&]() {
        ref_extractor* e = ref_extractor_create(p);
        std::vector<orbx_keypoint> kps(cap);
        std::vector<uint8_t> desc((size_t)cap * 32);
        for (;;) {
            const int f = next.fetch_add(1);
            if (f >= nframes) break;
            int n = extract_one(e->ex, imgs + (size_t)f * frame_stride, w, h, stride, kps.data(), desc.data(), cap);
            if (n < 0) n = 0;
            K[f].assign(kps.begin(), kps.begin() + n);
            D[f].assign(desc.begin(), desc.begin() + (size_t)n * 32);
        }
        ref_extractor_destroy(e);
    };
    std::atomic<int> nextPair(0);
    std::atomic<int64_t> matches(0);
    std::atomic<uint64_t> cmps(0);
    auto match_work = [&]() {
        std::vector<float> prev;
        std::vector<int32_t> m12;
        for (;;) {
            const int i = nextPair.fetch_add(1);
            if (i >= nframes - 1) break;
            orbm_frame f1{(int32_t)K[i].size(), K[i].data(), D[i].data(), 0.f, (float)w, 0.f, (float)h, nullptr, nullptr, 0};
            orbm_frame f2{(int32_t)K[i + 1].size(), K[i + 1].data(), D[i + 1].data(), 0.f, (float)w, 0.f, (float)h, nullptr, nullptr, 0};
            prev.resize(2 * K[i].size() + 2);
            m12.resize(K[i].size() + 1);
            for (size_t k = 0; k < K[i].size(); ++k) { prev[2 * k] = K[i][k].x; prev[2 * k + 1] = K[i][k].y; }
            uint64_t c = 0;
            matches += oracle_search_for_initialization(&f1, &f2, prev.data(), m12.data(), window_size, nnratio, check_orientation, &c);
            cmps += c;
        }
    };
    auto run = [&](auto fn) {
        if (nthreads == 1) { fn(); return; }
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(fn);
        for (auto& t : th) t.join();
    };
    run(extract_work);
    run(match_work);
    int64_t total = 0;
    for (int f = 0; f < nframes; ++f) total += (int64_t)K[f].size();
    if (total_matches) *total_matches = matches;
    if (ncompares) *ncompares = cmps;
    return total;
}

} // extern "C"
