// Peer exchange of the packed genotype rows between the ranks of a multi-GPU run ("each GPU H2D-copies
// 1/G of a chunk, an all-gather over NVSwitch rebuilds the full chunk on every GPU").
//
// One process per GPU.  Every rank owns a gather buffer of the plan's geometry (cudaMalloc, exported with CUDA IPC).  The
// file rows are cut into segments (the same on every rank, whatever SNP chunking each rank's plan uses - that depends
// on its band); per segment a rank uploads its 1/G slice over its own PCIe link ("post"), then pulls the other slices
// straight out of the peers' buffers with the copy engines over NVLink ("pull").  No SM is used, so nothing
// competes with - or can dead-lock against - the persistent SYRK grid and its wave barrier (an NCCL all-gather kernel per
// chunk did exactly that: both grids were half resident and waited for each other's SMs).  Cross-process ordering is
// by inter-process CUDA events, one per (rank, segment); the plan waits on a local "chunk complete" event per chunk
// (gcta_b200_plan_set_chunk_events; recorded behind the last segment the chunk reads), so the exchange of the later
// rows runs behind the compute of the earlier chunks.
//
// The reference has no counterpart: it reads the whole .bed on one host thread (data.cpp:140-194).
#include "common.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

namespace {
constexpr int HANDLE = GCTA_B200_IPC_HANDLE_BYTES;
constexpr int MAX_PULL_STREAMS = 4;
constexpr int64_t SEGMENTS = 32;

// segments depend on the file's row count only
void seg_geometry(int64_t m_file, int64_t* seg_rows, int64_t* nseg)
{
    *seg_rows = std::max<int64_t>(1, (m_file + SEGMENTS - 1) / SEGMENTS);
    *nseg = (m_file + *seg_rows - 1) / *seg_rows;
}
static_assert(sizeof(cudaIpcMemHandle_t) == HANDLE && sizeof(cudaIpcEventHandle_t) == HANDLE, "CUDA IPC handle size");
}  // namespace

struct gcta_b200_exchange {
    gcta_b200_plan* plan = nullptr;
    int rank = 0, world = 1;
    int32_t device = 0;
    int64_t nchunks = 0, nseg = 0, seg_rows = 0, pitch = 0, stride = 0, m_file = 0;
    uint8_t* bed = nullptr;                  // this rank's gather buffer (the plan reads it)
    std::vector<uint8_t*> peer_bed;          // [world]: the peers' buffers mapped into this process (null for self)
    std::vector<cudaEvent_t> posted;         // [nseg] inter-process: my slice of segment s is in my buffer
    std::vector<cudaEvent_t> peer_posted;    // [world * nseg]: the peers' events opened here (null for self)
    std::vector<cudaEvent_t> ready;          // [nchunks] local: every row chunk c reads is in my buffer
    std::vector<cudaEvent_t> joined;         // [nstreams] scratch events to join the pull streams
    std::vector<int64_t> c0, c1;             // file rows of each of the plan's SNP chunks
    cudaStream_t s_up = nullptr;
    std::vector<cudaStream_t> s_pull;
    bool connected = false;

    // rows of segment s that rank q uploads
    void slice(int64_t s, int q, int64_t* r0, int64_t* r1) const { gcta_b200_exchange_slice(m_file, world, q, s, r0, r1); }
};

// Pure geometry (no device needed): the file rows are cut into at most 32 equal segments, each segment into `world`
// equal slices; returns the rows of segment `seg` that `rank` uploads and the number of segments.
extern "C" int64_t gcta_b200_exchange_slice(int64_t m_file, int32_t world, int32_t rank, int64_t seg, int64_t* row_begin, int64_t* row_end)
{
    if (m_file <= 0 || world < 1 || rank < 0 || rank >= world) return 0;
    int64_t seg_rows, nseg;
    seg_geometry(m_file, &seg_rows, &nseg);
    if (row_begin && row_end) {
        if (seg < 0 || seg >= nseg) { *row_begin = *row_end = m_file; return nseg; }
        const int64_t s0 = seg * seg_rows, s1 = std::min(m_file, s0 + seg_rows);
        const int64_t per = (s1 - s0 + world - 1) / world;
        *row_begin = std::min(s1, s0 + rank * per);
        *row_end = std::min(s1, s0 + (rank + 1) * per);
    }
    return nseg;
}

extern "C" int64_t gcta_b200_exchange_blob_bytes(gcta_b200_plan* plan)
{
    int64_t m_file = 0, seg_rows, nseg;
    if (!plan || gcta_b200_plan_geometry(plan, nullptr, &m_file, nullptr, nullptr)) return 0;
    seg_geometry(m_file, &seg_rows, &nseg);
    return (int64_t)HANDLE * (1 + nseg);
}

// Unmap the peers' buffers and events (after the last step; every rank does this, then a host barrier, then destroy,
// so that no buffer is freed while a peer still maps it).
extern "C" void gcta_b200_exchange_disconnect(gcta_b200_exchange* x)
{
    if (!x) return;
    cudaSetDevice(x->device);
    if (x->s_up) cudaStreamSynchronize(x->s_up);
    for (cudaStream_t s : x->s_pull) cudaStreamSynchronize(s);
    for (uint8_t*& q : x->peer_bed)
        if (q) { cudaIpcCloseMemHandle(q); q = nullptr; }
    for (cudaEvent_t& e : x->peer_posted)
        if (e) { cudaEventDestroy(e); e = nullptr; }
    x->connected = false;
    cudaGetLastError();
}

extern "C" void gcta_b200_exchange_destroy(gcta_b200_exchange* x)
{
    if (!x) return;
    gcta_b200_exchange_disconnect(x);
    if (x->plan) {   // the plan goes back to its own genotype buffer and stops waiting on this exchange's events
        gcta_b200_plan_set_chunk_events(x->plan, nullptr, 0, nullptr, 0);
        gcta_b200_plan_use_own_bed(x->plan);
    }
    for (cudaEvent_t e : x->posted) cudaEventDestroy(e);
    for (cudaEvent_t e : x->ready) cudaEventDestroy(e);
    for (cudaEvent_t e : x->joined) cudaEventDestroy(e);
    if (x->s_up) cudaStreamDestroy(x->s_up);
    for (cudaStream_t s : x->s_pull) cudaStreamDestroy(s);
    if (x->bed) cudaFree(x->bed);
    cudaGetLastError();
    delete x;
}

#define EX_CUDA(call)                                                                                                  \
    do {                                                                                                               \
        cudaError_t e_ = (call);                                                                                       \
        if (e_ != cudaSuccess) {                                                                                       \
            gcta_b200_exchange_destroy(x);                                                                             \
            return gb::set_err(err, errcap, e_ == cudaErrorMemoryAllocation ? GCTA_B200_ERR_NOMEM : GCTA_B200_ERR_CUDA, \
                               std::string("Error: CUDA call failed: ") + #call + ": " + cudaGetErrorString(e_));      \
        }                                                                                                              \
    } while (0)

extern "C" int gcta_b200_exchange_create(gcta_b200_plan* plan, int32_t rank, int32_t world, uint8_t* my_blob, gcta_b200_exchange** out,
                                         char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(plan && my_blob && out, "exchange_create: null argument");
    GB_ARG(world >= 1 && rank >= 0 && rank < world, "exchange_create: bad rank / world");
    gcta_b200_exchange* x = new gcta_b200_exchange();
    x->plan = plan;
    x->rank = rank;
    x->world = world;
    int rc = gcta_b200_plan_geometry(plan, &x->device, &x->m_file, &x->stride, &x->pitch);
    if (rc) { delete x; return set_err(err, errcap, rc, "Error: exchange_create: bad plan"); }
    seg_geometry(x->m_file, &x->seg_rows, &x->nseg);
    x->nchunks = gcta_b200_plan_num_chunks(plan);
    x->c0.resize(x->nchunks);
    x->c1.resize(x->nchunks);
    for (int64_t c = 0; c < x->nchunks; c++) gcta_b200_plan_chunk_rows(plan, c, &x->c0[c], &x->c1[c]);
    EX_CUDA(cudaSetDevice(x->device));
    EX_CUDA(cudaMalloc((void**)&x->bed, (size_t)x->pitch * x->m_file));
    EX_CUDA(cudaMemset(x->bed, 0, (size_t)x->pitch * x->m_file));   // pad bytes of a re-pitched row stay zero
    EX_CUDA(cudaStreamCreateWithFlags(&x->s_up, cudaStreamNonBlocking));
    const int nstreams = std::max(1, std::min(MAX_PULL_STREAMS, world - 1));
    x->s_pull.assign(nstreams, nullptr);
    x->joined.assign(nstreams, nullptr);
    for (int i = 0; i < nstreams; i++) {
        EX_CUDA(cudaStreamCreateWithFlags(&x->s_pull[i], cudaStreamNonBlocking));
        EX_CUDA(cudaEventCreateWithFlags(&x->joined[i], cudaEventDisableTiming));
    }
    x->posted.assign(x->nseg, nullptr);
    x->ready.assign(x->nchunks, nullptr);
    cudaIpcMemHandle_t mh;
    EX_CUDA(cudaIpcGetMemHandle(&mh, x->bed));
    memcpy(my_blob, &mh, HANDLE);
    for (int64_t c = 0; c < x->nchunks; c++) EX_CUDA(cudaEventCreateWithFlags(&x->ready[c], cudaEventDisableTiming));
    for (int64_t s = 0; s < x->nseg; s++) {
        EX_CUDA(cudaEventCreateWithFlags(&x->posted[s], cudaEventDisableTiming | cudaEventInterprocess));
        cudaIpcEventHandle_t eh;
        EX_CUDA(cudaIpcGetEventHandle(&eh, x->posted[s]));
        memcpy(my_blob + (size_t)HANDLE * (1 + s), &eh, HANDLE);
    }
    int64_t pitch = 0;
    rc = gcta_b200_plan_use_device_bed(plan, x->bed, &pitch, err, errcap);
    if (rc) { gcta_b200_exchange_destroy(x); return rc; }
    *out = x;
    return 0;
}

extern "C" int gcta_b200_exchange_connect(gcta_b200_exchange* x, const uint8_t* all_blobs, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(x && all_blobs, "exchange_connect: null argument");
    GB_ARG(!x->connected, "exchange_connect: already connected");
    GB_CUDA(cudaSetDevice(x->device));
    const size_t blob = (size_t)HANDLE * (1 + x->nseg);
    x->peer_bed.assign(x->world, nullptr);
    x->peer_posted.assign((size_t)x->world * x->nseg, nullptr);
    for (int q = 0; q < x->world; q++) {
        if (q == x->rank) continue;
        const uint8_t* b = all_blobs + blob * q;
        cudaIpcMemHandle_t mh;
        memcpy(&mh, b, HANDLE);
        GB_CUDA(cudaIpcOpenMemHandle((void**)&x->peer_bed[q], mh, cudaIpcMemLazyEnablePeerAccess));
        for (int64_t s = 0; s < x->nseg; s++) {
            cudaIpcEventHandle_t eh;
            memcpy(&eh, b + (size_t)HANDLE * (1 + s), HANDLE);
            GB_CUDA(cudaIpcOpenEventHandle(&x->peer_posted[(size_t)q * x->nseg + s], eh));
        }
    }
    x->connected = true;
    return 0;
}

// Phase 1 of a step: enqueue the host->device copy of this rank's slice of every segment and record its inter-process
// event.  Returns without waiting.  The caller must have made sure (host barrier) that no peer still reads this
// rank's buffer from the previous step.
extern "C" int gcta_b200_exchange_post(gcta_b200_exchange* x, const uint8_t* bed_host, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(x && bed_host, "exchange_post: null argument");
    GB_CUDA(cudaSetDevice(x->device));
    for (int64_t s = 0; s < x->nseg; s++) {
        int64_t r0, r1;
        x->slice(s, x->rank, &r0, &r1);
        int rc = gcta_b200_plan_upload_rows_async(x->plan, bed_host, r0, r1, x->s_up, err, errcap);
        if (rc) return rc;
        GB_CUDA(cudaEventRecord(x->posted[s], x->s_up));
    }
    return 0;
}

// Phase 2, after a host barrier that follows every rank's post (a wait captures the event's latest record, so the
// peers' records of THIS step must exist): per segment, wait for each peer's slice and copy it out of the peer's buffer
// (copy engines, NVLink); record a "chunk complete" event behind the last segment each of the plan's chunks reads and
// hand those to the plan.  Returns without waiting.
extern "C" int gcta_b200_exchange_pull(gcta_b200_exchange* x, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(x, "exchange_pull: null argument");
    GB_ARG(x->connected || x->world == 1, "exchange_pull: exchange_connect has not been called");
    GB_CUDA(cudaSetDevice(x->device));
    const int ns = (int)x->s_pull.size();
    int64_t c = 0;   // next chunk without a "complete" event (chunks ascend in file rows)
    for (int64_t s = 0; s < x->nseg; s++) {
        for (int k = 1; k < x->world; k++) {
            const int q = (x->rank + k) % x->world;   // every rank starts with a different peer
            cudaStream_t st = x->s_pull[(k - 1) % ns];
            int64_t r0, r1;
            x->slice(s, q, &r0, &r1);
            if (r1 <= r0) continue;
            GB_CUDA(cudaStreamWaitEvent(st, x->peer_posted[(size_t)q * x->nseg + s], 0));
            GB_CUDA(cudaMemcpyAsync(x->bed + r0 * x->pitch, x->peer_bed[q] + r0 * x->pitch, (size_t)(r1 - r0) * x->pitch, cudaMemcpyDefault, st));
        }
        // join: pull stream 0 waits for my own upload of this segment and for the other pull streams
        GB_CUDA(cudaStreamWaitEvent(x->s_pull[0], x->posted[s], 0));
        for (int i = 1; i < ns; i++) {
            GB_CUDA(cudaEventRecord(x->joined[i], x->s_pull[i]));
            GB_CUDA(cudaStreamWaitEvent(x->s_pull[0], x->joined[i], 0));
        }
        const int64_t s1 = std::min(x->m_file, (s + 1) * x->seg_rows);
        for (; c < x->nchunks && x->c1[c] <= s1; c++) GB_CUDA(cudaEventRecord(x->ready[c], x->s_pull[0]));
    }
    GB_ARG(c == x->nchunks, "exchange_pull: a chunk reads rows beyond the file");
    return gcta_b200_plan_set_chunk_events(x->plan, (void* const*)x->ready.data(), x->nchunks, err, errcap);
}

// End of a step: wait until this rank's uploads and pulls are complete (they are, once the plan's run has returned;
// this makes it explicit).  A host barrier after it tells every rank that its buffer may be overwritten.
extern "C" int gcta_b200_exchange_finish(gcta_b200_exchange* x, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(x, "exchange_finish: null argument");
    GB_CUDA(cudaSetDevice(x->device));
    GB_CUDA(cudaStreamSynchronize(x->s_up));
    for (cudaStream_t s : x->s_pull) GB_CUDA(cudaStreamSynchronize(s));
    return 0;
}
