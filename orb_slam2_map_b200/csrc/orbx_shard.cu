// orbx_shard.cu — BASELINE.json config 5 across GPUs behind the C ABI (SURVEY.md §8(b) "Matcher signature", §8(e)):
// the database is sharded by rows (one shard per rank), queries are replicated, every rank scans its shard with K10
// (knn2_kernel), the per-query (best, second) records are exchanged with ncclAllGather (12 bytes per query and rank) and
// merged with K11 (knn2_merge_kernel) under the (distance, global index) order.  Queries go in chunks: the gather + merge
// of chunk i runs on a second stream while the scan of chunk i+1 occupies the SMs.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): liborbx.so keeps linking cudart only, a process that already
// loaded NCCL (torch) gets that copy, and a consumer that never calls the sharded entry points needs no NCCL at all.

#include "orbx_internal.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <mutex>

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*CommCount)(const ncclComm_t, int*);
    ncclResult_t (*CommUserRank)(const ncclComm_t, int*);
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
    bool ok = false;
};

NcclApi g_nccl;
std::once_flag g_ncclOnce;

int nccl_api(NcclApi** out)
{
    std::call_once(g_ncclOnce, [] {
        void* h = nullptr;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) { h = dlopen(name, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) return;
        bool all = true;
        auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p) all = false; return p; };
        g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))sym("ncclGetUniqueId");
        g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))sym("ncclCommInitRank");
        g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
        g_nccl.CommCount = (decltype(g_nccl.CommCount))sym("ncclCommCount");
        g_nccl.CommUserRank = (decltype(g_nccl.CommUserRank))sym("ncclCommUserRank");
        g_nccl.AllGather = (decltype(g_nccl.AllGather))sym("ncclAllGather");
        g_nccl.GroupStart = (decltype(g_nccl.GroupStart))sym("ncclGroupStart");
        g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))sym("ncclGroupEnd");
        g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
        g_nccl.ok = all;
    });
    if (!g_nccl.ok) { orbx_set_error("NCCL not available (dlopen libnccl.so.2 failed): the sharded search needs it"); return ORBX_E_STATE; }
    *out = &g_nccl;
    return ORBX_OK;
}

#define ORBX_NCCL(api, call)                                                                              \
    do {                                                                                                  \
        ncclResult_t r__ = (call);                                                                        \
        if (r__ != ncclSuccess) {                                                                         \
            orbx_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, (api)->GetErrorString(r__)); \
            return ORBX_E_CUDA;                                                                           \
        }                                                                                                 \
    } while (0)

} // namespace

extern "C" {

int orbm_nccl_unique_id(uint8_t id[ORBM_NCCL_ID_BYTES])
{
    NcclApi* n; int rc = nccl_api(&n); if (rc) return rc;
    static_assert(sizeof(ncclUniqueId) == ORBM_NCCL_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId u;
    ORBX_NCCL(n, n->GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return ORBX_OK;
}

int orbm_nccl_comm_create(const uint8_t id[ORBM_NCCL_ID_BYTES], int rank, int nranks, void** comm)
{
    if (!id || !comm || nranks < 1 || rank < 0 || rank >= nranks) { orbx_set_error("bad argument"); return ORBX_E_INVALID; }
    NcclApi* n; int rc = nccl_api(&n); if (rc) return rc;
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclComm_t c = nullptr;
    ORBX_NCCL(n, n->CommInitRank(&c, nranks, u, rank));         // on the calling thread's current device
    *comm = c;
    return ORBX_OK;
}

int orbm_nccl_comm_destroy(void* comm)
{
    if (!comm) return ORBX_OK;
    NcclApi* n; int rc = nccl_api(&n); if (rc) return rc;
    ORBX_NCCL(n, n->CommDestroy((ncclComm_t)comm));
    return ORBX_OK;
}

int orbm_knn2_sharded(void* comm, const uint8_t* queries, int64_t nq, const uint8_t* db_shard, int64_t ndb_local,
                      int32_t db_index_offset, int64_t query_chunk, uint16_t* out_dist, int32_t* out_idx, void* stream)
{
    ORBX_NVTX("orbm_knn2_sharded");
    if (nq <= 0) return ORBX_OK;
    if (!comm) return orbm_knn2_device(queries, nq, db_shard, ndb_local, db_index_offset, out_dist, out_idx, stream);
    NcclApi* n; int rc = nccl_api(&n); if (rc) return rc;
    ncclComm_t c = (ncclComm_t)comm;
    int nranks = 0, rank = 0;
    ORBX_NCCL(n, n->CommCount(c, &nranks));
    ORBX_NCCL(n, n->CommUserRank(c, &rank));
    // a one-rank communicator takes the short cut unless the caller asked for chunks (tests walk the exchange path on one GPU)
    if (nranks == 1 && query_chunk <= 0) return orbm_knn2_device(queries, nq, db_shard, ndb_local, db_index_offset, out_dist, out_idx, stream);
    cudaStream_t st = (cudaStream_t)stream;
    // chunks: a multiple of the kernel's 1024-query tile.  Default: two chunks — enough to hide all but the last exchange —
    // but never fewer query tiles per chunk than CTAs the GPU holds (2 per SM): below that the scan splits the database to
    // fill the machine and pays the pipeline prologue twice (measured: 4 chunks of 256 tiles ran 13 % slower than 1024 tiles)
    int64_t chunk = query_chunk;
    if (chunk <= 0) {
        int dev = 0, sms = 148;
        ORBX_CUDA(cudaGetDevice(&dev));
        ORBX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        chunk = std::max<int64_t>((nq + 1) / 2, (int64_t)2 * sms * 1024);
    }
    chunk = std::min<int64_t>((chunk + 1023) / 1024 * 1024, (nq + 1023) / 1024 * 1024);
    const int nbuf = 2;
    // per buffer: local (dist, idx) of one chunk + the gathered [nranks] copies
    const size_t distBytes = (size_t)chunk * 2 * sizeof(uint16_t), idxBytes = (size_t)chunk * 2 * sizeof(int32_t);
    const size_t perBuf = (distBytes + idxBytes) * (size_t)(1 + nranks);
    uint8_t* scratch = nullptr;
    ORBX_CUDA(cudaMallocAsync((void**)&scratch, perBuf * nbuf, st));
    cudaStream_t cs = nullptr;
    ORBX_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaEvent_t evScan[nbuf], evFree[nbuf], evEnd;
    for (int b = 0; b < nbuf; ++b) {
        ORBX_CUDA(cudaEventCreateWithFlags(&evScan[b], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&evFree[b], cudaEventDisableTiming));
    }
    ORBX_CUDA(cudaEventCreateWithFlags(&evEnd, cudaEventDisableTiming));
    // the scratch allocation is ordered on st: the communication stream must not touch it earlier
    ORBX_CUDA(cudaEventRecord(evEnd, st));
    ORBX_CUDA(cudaStreamWaitEvent(cs, evEnd, 0));
    int ci = 0;
    for (int64_t q0 = 0; q0 < nq; q0 += chunk, ++ci) {
        const int64_t cq = std::min(chunk, nq - q0);
        const int b = ci % nbuf;
        uint8_t* base = scratch + (size_t)b * perBuf;
        int32_t* li = (int32_t*)base;                               // local idx, then local dist, then gathered idx, gathered dist
        uint16_t* ld = (uint16_t*)(base + idxBytes);
        int32_t* gi = (int32_t*)(base + idxBytes + distBytes);
        uint16_t* gd = (uint16_t*)(base + idxBytes + distBytes + (size_t)nranks * idxBytes);
        if (ci >= nbuf) ORBX_CUDA(cudaStreamWaitEvent(st, evFree[b], 0));           // the merge of chunk ci-2 has read this buffer
        rc = orbm_knn2_device(queries + q0 * 32, cq, db_shard, ndb_local, db_index_offset, ld, li, st);
        if (rc) return rc;
        ORBX_CUDA(cudaEventRecord(evScan[b], st));
        ORBX_CUDA(cudaStreamWaitEvent(cs, evScan[b], 0));
        // rank-major gather: gathered[r][q][2]; every rank contributes cq records (the tail chunk is shorter on all ranks alike)
        ORBX_NCCL(n, n->GroupStart());
        ORBX_NCCL(n, n->AllGather(li, gi, (size_t)cq * 2 * sizeof(int32_t), ncclUint8, c, cs));
        ORBX_NCCL(n, n->AllGather(ld, gd, (size_t)cq * 2 * sizeof(uint16_t), ncclUint8, c, cs));
        ORBX_NCCL(n, n->GroupEnd());
        rc = orbm_knn2_merge_device(gd, gi, nranks, cq, out_dist + q0 * 2, out_idx + q0 * 2, cs);
        if (rc) return rc;
        ORBX_CUDA(cudaEventRecord(evFree[b], cs));
    }
    ORBX_CUDA(cudaEventRecord(evEnd, cs));
    ORBX_CUDA(cudaStreamWaitEvent(st, evEnd, 0));                   // the caller's stream sees the merged result
    ORBX_CUDA(cudaFreeAsync(scratch, st));
    for (int b = 0; b < nbuf; ++b) { cudaEventDestroy(evScan[b]); cudaEventDestroy(evFree[b]); }
    cudaEventDestroy(evEnd);
    cudaStreamDestroy(cs);                                          // released when its work has drained
    return ORBX_OK;
}

} // extern "C"
