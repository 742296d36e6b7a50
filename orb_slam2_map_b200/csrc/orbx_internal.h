// orbx_internal.h — shared declarations of liborbx.so (not part of the public C ABI).
#ifndef ORBX_INTERNAL_H
#define ORBX_INTERNAL_H

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>      // header-only NVTX v3: ranges cost a pointer test when no tool is attached
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/orbx.h"

#define ORBX_MAX_LEVELS 16
#define ORBX_L0_BUFS 2        // chunk-sized input buffers of the host-image pipeline (3 measured no faster: the kernels, not the copies, pace it)
#define ORBX_GRID_COLS 64     // FRAME_GRID_COLS, reference include/Frame.h:38
#define ORBX_GRID_ROWS 48     // FRAME_GRID_ROWS, reference include/Frame.h:37
#define ORBX_GRID_CELLS (ORBX_GRID_COLS * ORBX_GRID_ROWS)
#define ORBX_TH_HIGH 100      // reference src/ORBmatcher.cc:37
#define ORBX_TH_LOW 50        // reference src/ORBmatcher.cc:38
#define ORBX_HISTO_LENGTH 30  // reference src/ORBmatcher.cc:39

void orbx_set_error(const char* fmt, ...);

// NVTX range around a stage / entry point (SURVEY.md §5 tooling): ORBX_NVTX("orbx:fast");
struct OrbxNvtxScope {
    explicit OrbxNvtxScope(const char* name) { nvtxRangePushA(name); }
    ~OrbxNvtxScope() { nvtxRangePop(); }
};
#define ORBX_NVTX_CAT2(a, b) a##b
#define ORBX_NVTX_CAT(a, b) ORBX_NVTX_CAT2(a, b)
#define ORBX_NVTX(name) OrbxNvtxScope ORBX_NVTX_CAT(orbxNvtxScope, __LINE__)(name)

#define ORBX_CUDA(call)                                                                         \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            orbx_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                 \
                           cudaGetErrorString(e__));                                            \
            return ORBX_E_CUDA;                                                                 \
        }                                                                                       \
    } while (0)

// ---- per-level geometry (host-computed, passed to kernels by value) -------------------------
struct LevelGeom {
    int w, h, pitch;        // level size, row pitch in bytes (multiple of 16)
    int off;                // byte offset of the level inside one frame's pyramid block
    int cellBeg, nCells;    // range of this level's FAST cells in the cell table
    int candOff, candCap;   // candidate region (entries) inside one frame's candidate block
    int cellCap;            // ... of which every FAST cell owns a fixed segment of cellCap entries
    int kpOff, kpCap;       // culled-keypoint region inside one frame's keypoint block
    int N;                  // mnFeaturesPerLevel[level]
    int nIni;               // quadtree roots
    float hX;               // root width (float, as the reference computes it)
    int winW, winH;         // maxBorderX-minBorderX, maxBorderY-minBorderY
    int coefX, coefY;       // offsets into the resize coefficient table (levels >= 1)
    int blurTH;             // output rows per blur strip (7g-6)
    int wCell, hCell;       // FAST cell stride (reference src/ORBextractor.cc:786-787)
    float scale;            // mvScaleFactor[level]
    float kpSize;           // (float)(int)(31*scale)
};

struct Geom {
    int nlevels, ncells, candTotal, kpTotal;
    int cwMax, chMax;            // largest FAST cell sub-image
    long long pyrFrameBytes;     // bytes of one frame's pyramid block (all levels, pitched)
    LevelGeom L[ORBX_MAX_LEVELS];
};

// one FAST cell = one cv::FAST call of the reference (src/ORBextractor.cc:789-829)
struct Cell {
    unsigned short x0, y0;      // iniX, iniY (level coordinates)
    unsigned short cw, ch;      // sub-image size (maxX-iniX, maxY-iniY)
    unsigned short offx, offy;  // j*wCell, i*hCell added to cell-local coordinates
    unsigned short level, pad;
};

// where level 0 of each frame of the current chunk lives (caller's device images or the
// workspace copy)
struct Level0 {
    const uint8_t* base;
    const uint8_t* base1;     // stereo: right-eye images (chunk frame 2p = left p, 2p+1 = right p)
    long long frameStride;
    int pitch;
    int stereo;
};
static __host__ __device__ __forceinline__ const uint8_t* l0_frame(const Level0& l0, int frame)
{
    if (l0.stereo) return ((frame & 1) ? l0.base1 : l0.base) + (long long)(frame >> 1) * l0.frameStride;
    return l0.base + (long long)frame * l0.frameStride;
}

struct orbx_handle {
    orbx_params params;
    int device;
    cudaStream_t stream;
    bool ownStream;
    int maxW, maxH, chunkFrames, resultSlots;
    // tables as the reference ctor computes them
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> perLevel;
    int umax[16];
    // geometry of the current image shape
    int curW, curH;
    Geom geom;
    // device tables
    Cell* dCells;
    int2* dCoef;        // resize coefficients
    int* dBlurTiles;
    int nBlurTiles;
    int* dKpMap;        // keypoint-slot -> level map for the describe kernel
    // workspace (chunkFrames deep)
    uint8_t* dLevel0;   // pitched copies of the input images (when needed): ORBX_L0_BUFS chunk-sized buffers
    uint8_t* dStageLin; // same size: landing zone for LINEAR host->device copies (then re-pitched on the device)
    cudaStream_t copyStream;        // host->device copies of chunk i+1 overlap the kernels of chunk i
    cudaStream_t d2hStream;         // device->host copies of chunk i overlap the kernels of chunk i+1
    cudaEvent_t evCopied[ORBX_L0_BUFS], evFree[ORBX_L0_BUFS], evDone;
    bool evFreeValid[ORBX_L0_BUFS];
    // page-locked staging of the batch pipeline for callers whose buffers are PAGEABLE (a driver-staged copy blocks the host and
    // serialises the pipeline): inputs are copied here by a few host threads and uploaded from here, results land here and are
    // handed to the caller one chunk later (orbx_api.cu: batch_to_host_impl)
    uint8_t* hInPool[ORBX_L0_BUFS]; size_t hInPoolBytes;
    uint8_t* hOutPool[2]; size_t hOutPoolBytes;
    cudaEvent_t evInPool[ORBX_L0_BUFS], evOutPool[2];
    bool evInPoolValid[ORBX_L0_BUFS];
    int l0Pitch;
    long long l0FrameBytes;
    uint8_t* dRaw;      // pyramid levels >= 1
    uint8_t* dBlur;     // blurred levels (all)
    uint32_t* dCand;    // FAST candidates, per-cell segments
    uint32_t* dKeysA;   // quadtree key buffers
    uint32_t* dKeysB;
    int2* dCellSeg;
    uint32_t* dLevelKps;    // culled keypoints per level (packed)
    int* dLevelKpCount;
    int* dStatus;           // device-side error flags
    Level0 lastL0;
    int lastChunkFrames;
    // result store (resultSlots deep)
    int kpCap;              // per-frame keypoint capacity (= geom.kpTotal at creation shape)
    orbx_keypoint* dKps;
    uint8_t* dDesc;
    int* dCount;
    int* dSlotLevelCount;
    // RGB-D frame front (orbx_frame.cu): mvKeysUn records; when storeUndistorted the matchers and the grid use them
    orbx_keypoint* dKpsUn;
    bool storeUndistorted;
    float bounds[4];        // mnMinX, mnMaxX, mnMinY, mnMaxY of the undistorted frames
    uint8_t* dFrontStage;   // gray / colour / depth staging of orbx_frames_rgbd (grown on demand)
    size_t frontStageBytes;
    float boundsKey[11];    // camera + image size the cached bounds belong to
    bool boundsValid;
    // frame grid (per slot)
    unsigned short* dGridStart;   // [slot][GRID_CELLS+1]
    unsigned short* dGridIdx;     // [slot][kpCap]
    // matcher outputs (per slot = pair index)
    int* dMatches12;
    int* dNMatches;
    unsigned long long* dNCompares;
    float* dURight;
    float* dDepth;
    uint32_t* dMatchScratch;
    long long matchScratchPerPair;
    // pinned, device-mapped staging of the single-frame calls: inputs are read and results written by the kernels
    // THROUGH the mapping (no copy commands: a small cudaMemcpyAsync costs ~8 us of latency each)
    void* hStage;
    size_t hStageBytes;
    uint8_t* hPack;         // [count, status, pad to 256 B][mvKeys cap x 28][descriptors cap x 32][mvKeysUn cap x 28][mvuRight cap x 4][mvDepth cap x 4]
    size_t hPackBytes;
    cudaEvent_t evPack;     // recorded when the results of a single-frame call are in hPack
    // single-frame call (orbx_extract): the kernels of one frame captured once per image shape
    cudaGraphExec_t graph1;
    bool graph1Valid;
    long long graph1Launches[ORBX_NSTAGES];
    // forked form of one frame (orbx_run_frame_forked): per-level side streams and events, blur tile range per level
    std::vector<cudaStream_t> sideStreams;
    std::vector<cudaEvent_t> sideEvents, levelEvents;
    int blurTileBeg[ORBX_MAX_LEVELS + 1];
    // per-stage device timing (CUDA events on the handle's stream) and launch counting
    bool profiling;
    std::vector<cudaEvent_t> evPool;
    std::vector<int> evStage;       // stage id of the interval that ENDS at event i (-1 = interval start)
    size_t evUsed;
    double stageMs[ORBX_NSTAGES];
    long long stageLaunches[ORBX_NSTAGES];
};

// record an event closing the interval of `stage` (or opening one when stage < 0)
void orbx_prof_mark(orbx_handle* h, int stage);
#define ORBX_COUNT(h, stage, n) ((h)->stageLaunches[stage] += (n))

// orbx_extract.cu
// Device scratch of the stateless entry points (orbm_*, the frame-front helpers): a per-thread cache of blocks in
// power-of-two size classes, so a call repeated every frame does no cudaMalloc / cudaFree after its first time.
// A block is reused only by the thread that released it, on the same (legacy default) stream: stream order makes the
// reuse safe even if the previous call's kernels were still in flight.
// The stream of the stateless orbm_* calls: one non-blocking stream per calling thread and device, so that calls from
// Tracking / LocalMapping / LoopClosing threads neither serialise on the legacy default stream nor wait for the extractor handles'
// streams.  A thread's scratch blocks are reused by that thread on this stream only (stream order keeps the reuse safe).
cudaStream_t orbx_call_stream();
int orbx_scratch_alloc(void** p, size_t bytes);
void orbx_scratch_free(void* p);
int orbx_scratch_host(void** p, size_t bytes);     // this thread's pinned staging buffer (grown on demand, kept)
int orbx_setup_shape(orbx_handle* h, int w, int hgt);
int orbx_grow_result_store(orbx_handle* h, int newCap);   // orbx_api.cu: re-allocates everything sized by kpCap (stored slots are discarded)
void orbx_drop_graph(orbx_handle* h);      // the workspace moved: the captured single-frame graph is stale
int orbx_run_chunk(orbx_handle* h, Level0 l0, int nframes, int firstSlot);
int orbx_run_frame_forked(orbx_handle* h, Level0 l0, int firstSlot);   // one frame, per-level fork/join (latency form)
int orbx_launch_single_frame(orbx_handle* h);    // orbx_api.cu: graph replay of one frame whose level 0 is in dLevel0 -> slot 0
int orbx_ensure_host_stage(orbx_handle* h, size_t bytes);   // orbx_api.cu: the handle's pinned staging buffer, grown on demand
int orbx_ensure_pack(orbx_handle* h);                      // orbx_api.cu: the result pack of the single-frame calls (sized by kpCap)
struct OrbxPackLayout { size_t kps, desc, kun, ur, z, total; };
OrbxPackLayout orbx_pack_layout(int kpCap);

// orbx_api.cu: frames [nf][height][row_stride] -> pitched workspace frames, on `st`
int orbx_copy_frames(orbx_handle* h, uint8_t* dst, long long dstFrameBytes, uint8_t* stage, size_t stageBytes, const uint8_t* src, size_t frame_stride,
                     size_t row_stride, int width, int height, int nf, bool srcOnDevice, cudaStream_t st, long long* frameStrideOut);

// orbx_match.cu
int orbx_build_grids(orbx_handle* h, int firstSlot, int nframes, int width, int height);
int orbx_ensure_match_buffers(orbx_handle* h);
// grids of result slots [gridFirst, +gridCount) then SearchForInitialization for pairs (s, s+1), s in [pairFirst, +npairs); async on h->stream
int orbx_match_chunk(orbx_handle* h, int gridFirst, int gridCount, int pairFirst, int npairs, int window, float nnratio, int checkOri);

#endif
