// plan.cu - orchestration of the --make-grm path on one GPU (one row band of the GRM).
//
// Mirrors the sequence inside gcta::make_grm (grm.cpp:219-301), one SNP chunk at a time into a reused Z buffer:
//   keep list             -> K1c compact_rows (only with --keep / --remove)
//   calcu_mu              -> K1s snp_stats
//   make_XMat + std_XMat  -> K1f decode_emit_diag (also the diagonal and, with missing genotypes, the indicators);
//                            K1 bed_decode + K1d diag_accum for the X chromosome, --grm-alg 1 and split-FP16 operands
//   triple loop numerator -> K2 syrk_tiles accumulating FP32 S tiles over the chunks, in TMEM chains of <= 32,768 SNPs
//   A_N counts            -> K2 in counting mode on the E4M3 indicators, chunk by chunk (K1m + K3 AND+popc: cross-check)
//   A = S / N, zero rules -> K4 finalize_tiles into the packed lower triangle
// Around it: uploads one chunk ahead of the kernels (resident buffer, or two ring buffers when the file exceeds its HBM
// allowance), pinned staging by helper threads for pageable host buffers, and the download of the band behind the last
// chunk's SYRK.
// Multi-GPU: one plan per GPU with (part_rank, part_world); each plan owns a contiguous band of rows
// [row_begin,row_end) of the lower triangle, balanced by cost, so outputs are disjoint contiguous ranges of the packed
// triangle and no collective is needed on the data path (exchange.cu / multi.cu bring the packed rows to every GPU).
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <memory>
#include <thread>
#include <vector>

#include "common.cuh"

namespace gb {

static int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// column blocks needed by row block bi: columns 0 .. 128*bi+127 (clipped to n)
static int64_t tiles_in_row_block(int64_t bi, int64_t n)
{
    const int64_t last_row = std::min<int64_t>(bi * TILE_M + TILE_M - 1, n - 1);
    return last_row / TILE_N + 1;
}

// row blocks per raster group: with the wave barrier of K2 a wave of 148 tiles is ~12 x 12 tiles sharing 12 + 12 operand
// panels (measured on cfg3 with the barrier: 6: 1420, 12: 1433, 24: 1406 TFLOP/s; without it 6 was best; profiles/README.md)
static int raster_rows()
{
    static int v = 0;
    if (!v) {
        const char* e = getenv("GCTA_B200_RASTER");
        v = e ? atoi(e) : 12;
        if (v < 1) v = 12;
    }
    return v;
}

}  // namespace gb

extern "C" const char* gcta_b200_version(void) { return "gcta_b200 0.1.0 (sm_100a)"; }

extern "C" int gcta_b200_device_count(void)
{
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return c;
}

extern "C" int gcta_b200_partition(int64_t n, int32_t part_rank, int32_t part_world, int64_t* row_begin, int64_t* row_end)
{
    using namespace gb;
    if (n <= 0 || !row_begin || !row_end) return GCTA_B200_ERR_ARG;
    if (part_world <= 1) { *row_begin = 0; *row_end = n; return 0; }
    if (part_rank < 0 || part_rank >= part_world) return GCTA_B200_ERR_ARG;
    const int64_t nrb = (n + TILE_M - 1) / TILE_M;
    // Cost of a band of row blocks [b0, b1), in units of one 256x256 tile of SYRK work:
    //   SMS * ceil(tiles(b0, b1) / SMS)   the tensor work: K2 runs one persistent CTA per SM and all CTAs start each wave
    //                                     of SMS tiles together, so a launch takes whole waves
    // + ALPHA * b1                        K1 standardises columns [0, 256*b1) for this band (HBM-bound, replicated per GPU)
    // + BETA * (b1 - b0)                  K1d sums the diagonal of the band's own individuals
    // ALPHA and BETA are measured ratios on B200, re-measured in round 2 with the fused decode (2-GPU run of the headline
    // workload: a 256x256 tile 0.087 ms = a wave of 148 tiles 12.8 ms; K1f 0.083 ms per block of 256 individuals at M = 1M,
    // the diagonal of a band's own rows ~0.004 ms per block; all scale with M, so the ratios do not depend on M; round 1's
    // two-kernel decode had 1.33 / 0.83).  Boundaries minimise the largest band, counted in whole waves.
    double ALPHA = 0.96, BETA = 0.05;
    if (const char* e = getenv("GCTA_B200_PART_ALPHA")) ALPHA = atof(e);   // for re-calibration runs
    if (const char* e = getenv("GCTA_B200_PART_BETA")) BETA = atof(e);
    const int64_t SMS = 148;
    std::vector<int64_t> pref(nrb + 1, 0);
    for (int64_t b = 0; b < nrb; b++) pref[b + 1] = pref[b] + tiles_in_row_block(b, n);
    auto cost = [&](int64_t b0, int64_t b1) {
        const int64_t t = pref[b1] - pref[b0];
        return (double)((t + SMS - 1) / SMS * SMS) + ALPHA * b1 + BETA * (b1 - b0);
    };
    // the same with tiles counted one by one: used to break ties, so that among the partitions that reach the smallest
    // wave-counted maximum the bands are also as even as possible in plain work
    auto cost_lin = [&](int64_t b0, int64_t b1) { return (double)(pref[b1] - pref[b0]) + ALPHA * b1 + BETA * (b1 - b0); };
    auto cut = [&](double cap, double cap_lin, std::vector<int64_t>& bounds) {   // greedy: longest band within the caps, G bands at most
        bounds.assign(part_world + 1, nrb);
        bounds[0] = 0;
        int64_t b = 0;
        for (int g = 0; g < part_world; g++) {
            int64_t e = b;
            while (e < nrb && cost(b, e + 1) <= cap && cost_lin(b, e + 1) <= cap_lin) e++;
            bounds[g + 1] = e;
            b = e;
        }
        return b >= nrb;
    };
    const double INF = 1e300;
    double lo = 0.0, hi = cost(0, nrb);
    std::vector<int64_t> bounds, best;
    cut(hi, INF, best);
    for (int it = 0; it < 60; it++) {
        const double mid = 0.5 * (lo + hi);
        if (cut(mid, INF, bounds)) { hi = mid; best = bounds; }
        else lo = mid;
    }
    const double cap_wave = hi * 1.02;   // within 2 % of the best wave-counted maximum (less than one wave for any real size)
    lo = 0.0;
    hi = cost_lin(0, nrb);
    for (int it = 0; it < 60; it++) {
        const double mid = 0.5 * (lo + hi);
        if (cut(cap_wave, mid, bounds)) { hi = mid; best = bounds; }
        else lo = mid;
    }
    *row_begin = std::min<int64_t>(best[part_rank] * TILE_M, n);
    *row_end = std::min<int64_t>(best[part_rank + 1] * TILE_M, n);
    return 0;
}

extern "C" int64_t gcta_b200_tile_list(int64_t n, int64_t row_begin, int64_t row_end, int32_t* tiles)
{
    using namespace gb;
    if (n <= 0 || row_begin < 0 || row_end > n) return -1;
    if (row_begin >= row_end) return 0;
    if (row_begin % TILE_M != 0) return -1;
    const int64_t b0 = row_begin / TILE_M, b1 = (row_end + TILE_M - 1) / TILE_M;
    int64_t cnt = 0;
    const int RASTER_ROWS = raster_rows();
    for (int64_t g0 = b0; g0 < b1; g0 += RASTER_ROWS) {
        const int64_t g1 = std::min<int64_t>(g0 + RASTER_ROWS, b1);
        const int64_t maxcb = tiles_in_row_block(g1 - 1, n);
        for (int64_t bj = 0; bj < maxcb; bj++)
            for (int64_t bi = g0; bi < g1; bi++) {
                if (bj >= tiles_in_row_block(bi, n)) continue;
                if (tiles) { tiles[2 * cnt] = (int32_t)bi; tiles[2 * cnt + 1] = (int32_t)bj; }
                cnt++;
            }
    }
    return cnt;
}

struct gcta_b200_plan {
    // problem
    int64_t n_file = 0, m_file = 0, n = 0, m = 0, stride = 0, pitch = 0, ldz = 0;
    int xchr = 0, alg = 0, operand = 0, device = 0;
    bool has_keep = false, has_include = false, has_mu = false;
    int64_t row_begin = 0, row_end = 0, ntiles = 0, chunk = 0, band_count = 0;
    // device
    cudaStream_t stream = nullptr, stream2 = nullptr, stream_copy = nullptr;
    std::vector<cudaEvent_t> upload_events;
    std::vector<cudaEvent_t> chunk_ready;   // caller-owned events: chunk c's packed rows are in place (plan_set_chunk_events)
    // the same as a call-back: invoked on the host right before chunk c's first kernel is queued on `stream`; it may
    // block, and it queues whatever waits make the chunk's rows complete (gcta_b200_plan_set_chunk_hook; multi.cu)
    int (*chunk_hook)(void* ctx, int64_t chunk, void* stream) = nullptr;
    void* chunk_hook_ctx = nullptr;
    std::vector<int32_t> h_include, h_tiles;
    std::vector<int64_t> cstart;   // SNP chunk c = included SNPs [cstart[c], cstart[c+1]); every chunk <= `chunk` SNPs
    int64_t nchunks() const { return (int64_t)cstart.size() - 1; }
    bool downloaded = false;   // the last run already streamed A (and filled N) into the caller's buffers ...
    const float* streamed_A = nullptr;   // ... namely these (plan_download into other buffers still copies)
    const int32_t* streamed_N = nullptr;
    const int32_t* prefilled_N = nullptr;   // this N band was filled with m on the host while the last run computed
    int64_t diag_f0 = 0, diag_f1 = 0;   // file individuals whose diagonal this band needs (multiple of 16 at the low end)
    uint8_t *d_bed = nullptr, *d_bed_own = nullptr;   // d_bed: the buffer the kernels read; d_bed_own: the plan's allocation
    // Packed file larger than its HBM allowance: no resident copy; the rows of one SNP chunk at a time pass through two
    // ring buffers fed from host memory (plan_run_streamed / plan_run_to_host / make_grm only)
    bool stream_bed = false;
    uint8_t* d_ring[2] = {nullptr, nullptr};
    int64_t ring_rows = 0;
    cudaEvent_t slot_free[2] = {nullptr, nullptr};   // the last kernel that reads ring slot j has finished
    // pinned bounce slots of the helper threads that move pageable host buffers (Stager: genotypes in, Drainer: results out)
    // (pinning host memory costs ~1.5 ms per MB on this box, so the slots are few and small and the helpers only serve
    // transfers of GBs: STAGE_MIN_BYTES / DRAIN_MIN_BYTES)
    static constexpr int HELPERS = 4, HSLOTS = 2;
    static constexpr int64_t HSLOT_BYTES = 8ll << 20;
    static constexpr int64_t STAGE_MIN_BYTES = 2ll << 30, DRAIN_MIN_BYTES = 1ll << 30;
    uint8_t* h_up[HELPERS][HSLOTS] = {};
    uint8_t* h_down[HELPERS][HSLOTS] = {};
    uint32_t *d_keepmask = nullptr, *d_sexmale = nullptr, *d_planes = nullptr;
    int32_t *d_keep = nullptr, *d_include = nullptr, *d_cnt = nullptr, *d_sum = nullptr, *d_nii = nullptr, *d_tiles = nullptr;
    int32_t *d_Ntiles = nullptr, *d_N_lower = nullptr, *d_pairs = nullptr;   // d_pairs: CTA-pair work list of the cta_group::2 SYRK
    int64_t npairs = 0;
    double *d_mu_in = nullptr, *d_mu = nullptr, *d_diag = nullptr, *d_dk = nullptr, *d_dw = nullptr;   // d_dk, d_dw: --grm-alg 1
    float* d_Dtiles = nullptr;
    uint8_t* d_OKbuf = nullptr;   // E4M3 non-missing indicators of one SNP chunk, written by K1f when the counts run chunk by chunk
    float *d_zlut = nullptr, *d_S = nullptr, *d_A_lower = nullptr;
    void* d_Zbuf[2] = {nullptr, nullptr};
    unsigned long long* d_total_missing = nullptr;
    std::vector<cudaEvent_t> events;
    bool ran = false, overlap = false;
    bool fused_decode = false;   // K1s + K1f (decode2.cu) instead of K1 + K1d
    // keep list: each chunk's rows are first compacted to the kept individuals (K1c), and the fused kernels then see a
    // chunk in which everyone is kept - pitch cpitch, all-ones keep mask, row kk of the chunk at d_cbed + kk * cpitch
    bool compact = false;
    uint8_t* d_cbed = nullptr;
    uint32_t* d_ckeepmask = nullptr;
    int64_t cpitch = 0;
    int64_t chain = 32768;       // SNPs accumulated in one TMEM chain before the accumulator is added into S
    int64_t total_missing = 0;
    std::vector<void*> allocs;
};

namespace gb {

template <typename T>
static int dev_alloc(gcta_b200_plan* p, T** ptr, size_t count, char* err, size_t errcap)
{
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_err(err, errcap, e == cudaErrorMemoryAllocation ? GCTA_B200_ERR_NOMEM : GCTA_B200_ERR_CUDA,
                       std::string("Error: device allocation of ") + std::to_string(count * sizeof(T)) + " bytes failed: " + cudaGetErrorString(e));
    }
    p->allocs.push_back(q);
    *ptr = (T*)q;
    return 0;
}

}  // namespace gb

extern "C" void gcta_b200_plan_destroy(gcta_b200_plan* p)
{
    if (!p) return;
    cudaSetDevice(p->device);
    if (p->stream) cudaStreamSynchronize(p->stream);
    if (p->stream2) cudaStreamSynchronize(p->stream2);
    for (cudaEvent_t e : p->events) cudaEventDestroy(e);
    for (void* q : p->allocs) cudaFree(q);
    if (p->stream) { gb::wave_slot_release(p->device, p->stream); cudaStreamDestroy(p->stream); }
    if (p->stream2) { gb::wave_slot_release(p->device, p->stream2); cudaStreamDestroy(p->stream2); }
    if (p->stream_copy) cudaStreamDestroy(p->stream_copy);
    for (cudaEvent_t e : p->upload_events) cudaEventDestroy(e);
    for (int j = 0; j < 2; j++)
        if (p->slot_free[j]) cudaEventDestroy(p->slot_free[j]);
    for (int t = 0; t < gcta_b200_plan::HELPERS; t++)
        for (int i = 0; i < gcta_b200_plan::HSLOTS; i++) {
            if (p->h_up[t][i]) cudaFreeHost(p->h_up[t][i]);
            if (p->h_down[t][i]) cudaFreeHost(p->h_down[t][i]);
        }
    delete p;
}

extern "C" int gcta_b200_plan_create(const gcta_b200_args* a, gcta_b200_plan** out_plan, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(a && out_plan, "plan_create: null argument");
    GB_ARG(a->n_file > 0 && a->m_file > 0, "plan_create: empty genotype file");
    GB_ARG(a->n_file < (1ll << 31) && a->m_file < (1ll << 31), "plan_create: more than 2^31 individuals or SNPs");
    const int64_t n = a->keep ? a->n : a->n_file, m = a->include ? a->m : a->m_file;
    if (m <= 0) return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: No SNP is retained for analysis.");           // data.cpp:149
    if (n <= 0) return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: No individual is retained for analysis.");    // data.cpp:150
    GB_ARG(n <= a->n_file && m <= a->m_file, "plan_create: more kept than present");
    GB_ARG(a->alg == 0 || a->alg == 1, "plan_create: alg must be 0 or 1");
    GB_ARG(a->operand == GCTA_B200_FP16 || a->operand == GCTA_B200_BF16 || a->operand == GCTA_B200_FP16X2, "plan_create: unknown operand type");
    GB_ARG(!a->xchr || a->sex, "plan_create: xchr needs the sex array");
    if (gcta_b200_device_count() <= 0)
        return set_err(err, errcap, GCTA_B200_ERR_NO_DEVICE, "Error: no CUDA device is available (gcta_b200 has no CPU fallback).");
    GB_ARG(a->device >= 0 && a->device < gcta_b200_device_count(), "plan_create: bad device ordinal");
    if (a->keep)
        for (int64_t i = 0; i < n; i++)
            GB_ARG(a->keep[i] >= 0 && a->keep[i] < a->n_file && (i == 0 || a->keep[i] > a->keep[i - 1]), "plan_create: keep must be ascending file indices");
    if (a->include)
        for (int64_t k = 0; k < m; k++)
            GB_ARG(a->include[k] >= 0 && a->include[k] < a->m_file, "plan_create: include index out of range");
    if (a->xchr)
        for (int64_t i = 0; i < n; i++) {
            const int s = a->sex[a->keep ? a->keep[i] : i];
            if (s != 1 && s != 2)  // check_sex, grm.cpp:29-34
                return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: Sex information of some individuals are missing. Use --update-sex option to update the sex information of the individuals.");
        }

    GB_CUDA(cudaSetDevice(a->device));
    gcta_b200_plan* p = new gcta_b200_plan();
    *out_plan = nullptr;
    p->n_file = a->n_file; p->m_file = a->m_file; p->n = n; p->m = m;
    p->stride = (a->n_file + 3) / 4;
    p->pitch = round_up(p->stride, 16);
    p->ldz = round_up(n, TILE_N);
    p->xchr = a->xchr; p->alg = a->alg; p->operand = a->operand; p->device = a->device;
    // Rounding errors of 16-bit operands average out like 1/sqrt(M); with few SNPs they do not (a single FP16 product
    // is only good to ~1e-3 relative and A_ij is then O(1)).  Short contractions therefore use the split-FP16 mode,
    // which costs nothing measurable at these sizes and keeps max|dA| <= 1e-3 for any M.
    // BF16 (8 significand bits) needs the averaging even more, so it is upgraded the same way.
    if ((p->operand == GCTA_B200_FP16 || p->operand == GCTA_B200_BF16) && m < 8192) p->operand = GCTA_B200_FP16X2;
    p->has_keep = a->keep != nullptr && !(n == a->n_file);
    p->has_include = a->include != nullptr;
    p->has_mu = a->mu != nullptr;
    // the two streaming passes of decode2.cu cover the common case; GCTA_B200_DECODE=legacy forces the row-per-CTA kernels
    {
        const bool fast_ok = !a->xchr && a->alg == 0 && p->operand != GCTA_B200_FP16X2 &&
                             !(getenv("GCTA_B200_DECODE") && strcmp(getenv("GCTA_B200_DECODE"), "legacy") == 0);
        p->compact = fast_ok && p->has_keep && !(getenv("GCTA_B200_COMPACT") && atoi(getenv("GCTA_B200_COMPACT")) == 0);
        p->fused_decode = fast_ok && (!p->has_keep || p->compact);
    }
#define PLAN_TRY(x) do { int rc_ = (x); if (rc_) { gcta_b200_plan_destroy(p); return rc_; } } while (0)
#define PLAN_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { gcta_b200_plan_destroy(p); \
        return set_err(err, errcap, GCTA_B200_ERR_CUDA, std::string("Error: CUDA call failed: ") + #x + ": " + cudaGetErrorString(e_)); } } while (0)
    if (gcta_b200_partition(n, a->part_rank, a->part_world, &p->row_begin, &p->row_end) != 0) {
        gcta_b200_plan_destroy(p);
        return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: plan_create: bad part_rank / part_world");
    }
    // a row band only multiplies individuals [0, row_end): standardise (and store) no more columns than that
    p->ldz = round_up(std::max<int64_t>(p->row_end, 1), TILE_N);
    p->diag_f0 = p->row_begin;   // identity keep: file individual == kept individual (overridden below for a keep list)
    p->diag_f1 = p->row_end;
    p->ntiles = gcta_b200_tile_list(n, p->row_begin, p->row_end, nullptr);
    p->band_count = p->row_end * (p->row_end + 1) / 2 - p->row_begin * (p->row_begin + 1) / 2;
    PLAN_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    PLAN_CUDA(cudaStreamCreateWithFlags(&p->stream2, cudaStreamNonBlocking));

    // fixed-size buffers.  The packed genotypes stay resident unless they exceed their allowance: half of the free HBM,
    // or GCTA_B200_BED_CAP_GIB (a decimal number of GiB; for tests) - then they are streamed chunk by chunk (see stream_bed)
    {
        size_t free_b = 0, total_b = 0;
        PLAN_CUDA(cudaMemGetInfo(&free_b, &total_b));
        double cap = 0.5 * (double)free_b;
        if (const char* e = getenv("GCTA_B200_BED_CAP_GIB")) cap = atof(e) * 1073741824.0;
        p->stream_bed = (double)p->pitch * (double)a->m_file > cap;
    }
    if (!p->stream_bed) {
        PLAN_TRY(dev_alloc(p, &p->d_bed, (size_t)p->pitch * a->m_file, err, errcap));
        p->d_bed_own = p->d_bed;
        if (p->pitch != p->stride) PLAN_CUDA(cudaMemset(p->d_bed, 0, (size_t)p->pitch * a->m_file));
    }
    PLAN_TRY(dev_alloc(p, &p->d_keepmask, (size_t)p->pitch / 4, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_mu, (size_t)m, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_cnt, (size_t)m, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_sum, (size_t)m, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_zlut, (size_t)m * 8, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_diag, (size_t)a->n_file, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_nii, (size_t)a->n_file, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_total_missing, 1, err, errcap));
    if (a->alg == 1) {
        PLAN_TRY(dev_alloc(p, &p->d_dk, (size_t)m, err, errcap));
        PLAN_TRY(dev_alloc(p, &p->d_dw, (size_t)a->n_file, err, errcap));
    }
    PLAN_TRY(dev_alloc(p, &p->d_tiles, (size_t)p->ntiles * 2, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_S, (size_t)p->ntiles * TILE_M * TILE_N, err, errcap));
    PLAN_TRY(dev_alloc(p, &p->d_A_lower, (size_t)p->band_count, err, errcap));

    // host-built index structures
    std::vector<uint32_t> km((size_t)p->pitch / 4, 0u), mm;
    if (a->xchr) mm.assign((size_t)p->pitch / 4, 0u);
    for (int64_t i = 0; i < n; i++) {
        const int64_t f = p->has_keep ? a->keep[i] : i;
        km[f >> 4] |= 1u << ((f & 15) * 2);
        if (a->xchr && a->sex[f] == 1) mm[f >> 4] |= 1u << ((f & 15) * 2);
    }
    PLAN_CUDA(cudaMemcpy(p->d_keepmask, km.data(), km.size() * 4, cudaMemcpyHostToDevice));
    if (a->xchr) {
        PLAN_TRY(dev_alloc(p, &p->d_sexmale, mm.size(), err, errcap));
        PLAN_CUDA(cudaMemcpy(p->d_sexmale, mm.data(), mm.size() * 4, cudaMemcpyHostToDevice));
    }
    if (p->has_keep) {
        PLAN_TRY(dev_alloc(p, &p->d_keep, (size_t)n, err, errcap));
        PLAN_CUDA(cudaMemcpy(p->d_keep, a->keep, (size_t)n * 4, cudaMemcpyHostToDevice));
        if (p->row_end > p->row_begin) { p->diag_f0 = a->keep[p->row_begin] / 16 * 16; p->diag_f1 = (int64_t)a->keep[p->row_end - 1] + 1; }
        if (p->compact) {   // the diagonal is indexed by kept individual
            p->diag_f0 = p->row_begin;
            p->diag_f1 = p->row_end;
            p->cpitch = round_up((n + 3) / 4, 16);
            std::vector<uint32_t> ckm((size_t)p->cpitch / 4, 0u);
            for (int64_t i = 0; i < n; i++) ckm[i >> 4] |= 1u << ((i & 15) * 2);
            PLAN_TRY(dev_alloc(p, &p->d_ckeepmask, ckm.size(), err, errcap));
            PLAN_CUDA(cudaMemcpy(p->d_ckeepmask, ckm.data(), ckm.size() * 4, cudaMemcpyHostToDevice));
        }
    }
    if (p->has_include) {
        PLAN_TRY(dev_alloc(p, &p->d_include, (size_t)m, err, errcap));
        PLAN_CUDA(cudaMemcpy(p->d_include, a->include, (size_t)m * 4, cudaMemcpyHostToDevice));
        p->h_include.assign(a->include, a->include + m);
    }
    if (p->has_mu) {
        PLAN_TRY(dev_alloc(p, &p->d_mu_in, (size_t)m, err, errcap));
        PLAN_CUDA(cudaMemcpy(p->d_mu_in, a->mu, (size_t)m * 8, cudaMemcpyHostToDevice));
    }
    {
        std::vector<int32_t> tl((size_t)p->ntiles * 2 + 2);
        gcta_b200_tile_list(n, p->row_begin, p->row_end, tl.data());
        PLAN_CUDA(cudaMemcpy(p->d_tiles, tl.data(), (size_t)p->ntiles * 8, cudaMemcpyHostToDevice));
        p->h_tiles.assign(tl.begin(), tl.begin() + p->ntiles * 2);
        // GCTA_B200_2CTA=1: run K2 as CTA pairs (tcgen05.mma.cta_group::2, syrk2.cu) over the same tile storage
        if (getenv("GCTA_B200_2CTA") && atoi(getenv("GCTA_B200_2CTA")) > 0 && p->ntiles > 0) {
            p->npairs = gcta_b200_pair_list(tl.data(), p->ntiles, raster_rows(), nullptr);
            std::vector<int32_t> pl((size_t)p->npairs * 4 + 4);
            gcta_b200_pair_list(tl.data(), p->ntiles, raster_rows(), pl.data());
            PLAN_TRY(dev_alloc(p, &p->d_pairs, (size_t)p->npairs * 4, err, errcap));
            PLAN_CUDA(cudaMemcpy(p->d_pairs, pl.data(), (size_t)p->npairs * 16, cudaMemcpyHostToDevice));
        }
    }
    // SNP chunk: Z buffer of at most ~8 GiB (or what is left of HBM), multiple of 64 SNPs
    {
        size_t free_b = 0, total_b = 0;
        PLAN_CUDA(cudaMemGetInfo(&free_b, &total_b));
        // per Z buffer (two when overlapping).  Larger chunks mean fewer SYRK launches, hence fewer epilogues and
        // less read-modify-write of S; 16 GiB leaves room for everything else at every BASELINE configuration.
        int64_t zcap = 16ll << 30;
        if (const char* e = getenv("GCTA_B200_ZCAP_GIB")) zcap = std::max<int64_t>(1, atoll(e)) << 30;
        int64_t budget = std::min<int64_t>((int64_t)(free_b / 4), zcap);
        const int planes = p->operand == GCTA_B200_FP16X2 ? 2 : 1;   // split mode keeps a residual plane behind the first
        int64_t kc = a->chunk_snps > 0 ? a->chunk_snps : budget / (p->ldz * 2 * planes);
        kc = std::max<int64_t>(64, kc / 64 * 64);
        kc = std::min<int64_t>(kc, round_up(m, 64));
        kc = std::min<int64_t>(kc, 1 << 23);   // counts of one chunk must stay exact in the FP32 accumulator
        p->chunk = kc;
        // Length of one accumulation chain in TMEM.  The tensor cores add each MMA's partial sum into the FP32 accumulator
        // with truncation, so a long chain of same-sign products (the off-diagonal of duplicates or close relatives, which
        // grows like the diagonal) drifts low by ~2^-25 per MMA of the chain; the add of a finished chain into S is a
        // round-to-nearest FP32 add.  32,768 SNPs = 2,048 MMAs keeps that bias below 5e-5 relative; the SYRK of a Z chunk
        // is issued as one launch per chain-length slice of it.  GCTA_B200_CHAIN_SNPS overrides (0 = whole chunk).
        if (const char* e = getenv("GCTA_B200_CHAIN_SNPS")) p->chain = atoll(e);
        if (p->chain <= 0) p->chain = 1ll << 40;
        p->chain = std::max<int64_t>(128, p->chain / 128 * 128);
        // Chunk table.  When the plan picks the chunk length itself and there are many SNPs, the chunks ramp up: 1/64 of
        // the SNPs first, then twice the previous chunk until the Z buffer's length is reached.  A run that receives its
        // packed rows while it computes (plan_run_streamed / plan_run_to_host, per-chunk events of an exchange) can only
        // start chunk c once its rows are in HBM; with the ramp it waits for 1/64 of the file at the start and the
        // transfer of each later chunk has the (at least half as long) compute of the one before to hide behind -
        // a 16 GiB chunk of a narrow band would otherwise be up to half of the file.  Costs a few more launches.
        int64_t head = round_up((m + 63) / 64, 64);
        if (const char* e = getenv("GCTA_B200_HEAD_CHUNK")) head = round_up(std::max<int64_t>(0, atoll(e)), 64);
        if (a->chunk_snps > 0 || m < 65536 || head >= kc) head = 0;
        // Lengths in ascending order, so that the remainder is not the last chunk: a run that delivers to host buffers
        // downloads its band behind the last chunk's SYRK (run_last_chunk_pipelined), and with several GPUs the band's
        // copy (0.6 - 2.5 GB per rank) does not fit behind a short remainder (measured at 8 GPUs: the last chunk was 8 %
        // of the SNPs = 18 ms of compute against 25 ms of copy).
        std::vector<int64_t> lens;
        for (int64_t len = head > 0 ? head : kc, done = 0; done < m; len = std::min<int64_t>(kc, 2 * len)) {
            lens.push_back(std::min<int64_t>(len, m - done));
            done += lens.back();
        }
        std::stable_sort(lens.begin(), lens.end());
        p->cstart.assign(1, 0);
        for (int64_t l : lens) p->cstart.push_back(p->cstart.back() + l);
        PLAN_TRY(dev_alloc(p, (uint16_t**)&p->d_Zbuf[0], (size_t)kc * p->ldz * planes, err, errcap));
        if (p->compact) PLAN_TRY(dev_alloc(p, &p->d_cbed, (size_t)kc * p->cpitch, err, errcap));
        if (p->stream_bed) {
            // two ring buffers, each large enough for the file rows any one chunk reads
            for (int64_t c = 0; c < p->nchunks(); c++) {
                int64_t r0 = 0, r1 = 0;
                gcta_b200_plan_chunk_rows(p, c, &r0, &r1);
                p->ring_rows = std::max(p->ring_rows, r1 - r0);
            }
            for (int j = 0; j < 2; j++) {
                PLAN_TRY(dev_alloc(p, &p->d_ring[j], (size_t)p->ring_rows * p->pitch, err, errcap));
                PLAN_CUDA(cudaMemset(p->d_ring[j], 0, (size_t)p->ring_rows * p->pitch));   // the pad bytes of every row stay zero
                PLAN_CUDA(cudaEventCreateWithFlags(&p->slot_free[j], cudaEventDisableTiming));
            }
        }
        p->overlap = getenv("GCTA_B200_OVERLAP") && atoi(getenv("GCTA_B200_OVERLAP")) > 0;
        if (p->overlap && kc < round_up(m, 64)) PLAN_TRY(dev_alloc(p, (uint16_t**)&p->d_Zbuf[1], (size_t)kc * p->ldz * planes, err, errcap));
    }
    *out_plan = p;
    return 0;
}

extern "C" int gcta_b200_plan_upload(gcta_b200_plan* p, const uint8_t* bed_host, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && bed_host, "plan_upload: null argument");
    GB_ARG(!p->stream_bed, "plan_upload: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    GB_CUDA(cudaSetDevice(p->device));
    // re-pitch rows to a multiple of 16 bytes so K1 can use 128-bit loads; pad bytes are zero
    if (p->pitch != p->stride) GB_CUDA(cudaMemsetAsync(p->d_bed, 0, (size_t)p->pitch * p->m_file, p->stream));
    GB_CUDA(cudaMemcpy2DAsync(p->d_bed, (size_t)p->pitch, bed_host, (size_t)p->stride, (size_t)p->stride, (size_t)p->m_file,
                              cudaMemcpyHostToDevice, p->stream));
    GB_CUDA(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int gcta_b200_plan_use_device_bed(gcta_b200_plan* p, void* d_bed, int64_t* pitch, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p, "plan_use_device_bed: null plan");
    GB_ARG(!p->stream_bed || !d_bed, "plan_use_device_bed: this plan streams its genotypes from the host");
    if (pitch) *pitch = p->pitch;
    if (d_bed) {
        GB_ARG(((uintptr_t)d_bed & 15) == 0, "plan_use_device_bed: buffer must be 16-byte aligned");
        p->d_bed = (uint8_t*)d_bed;   // the plan's own buffer stays allocated and is freed with the plan
    }
    return 0;
}

extern "C" int gcta_b200_plan_use_own_bed(gcta_b200_plan* p)
{
    if (!p) return GCTA_B200_ERR_ARG;
    p->d_bed = p->d_bed_own;
    return 0;
}

extern "C" int gcta_b200_plan_geometry(gcta_b200_plan* p, int32_t* device, int64_t* m_file, int64_t* stride, int64_t* pitch)
{
    if (!p) return GCTA_B200_ERR_ARG;
    if (device) *device = p->device;
    if (m_file) *m_file = p->m_file;
    if (stride) *stride = p->stride;
    if (pitch) *pitch = p->pitch;
    return 0;
}

// ---- chunk-wise delivery of the packed rows by the caller (multi-GPU: upload 1/G + all-gather per chunk) ----
extern "C" int64_t gcta_b200_plan_num_chunks(gcta_b200_plan* p) { return p ? p->nchunks() : 0; }

extern "C" int gcta_b200_plan_chunk_rows(gcta_b200_plan* p, int64_t c, int64_t* row_begin, int64_t* row_end)
{
    if (!p || c < 0 || c >= gcta_b200_plan_num_chunks(p) || !row_begin || !row_end) return GCTA_B200_ERR_ARG;
    const int64_t k0 = p->cstart[c], k1 = p->cstart[c + 1];
    *row_begin = k0;
    *row_end = k1;
    if (p->has_include) {
        *row_begin = *std::min_element(p->h_include.begin() + k0, p->h_include.begin() + k1);
        *row_end = *std::max_element(p->h_include.begin() + k0, p->h_include.begin() + k1) + 1;
    }
    return 0;
}

extern "C" int gcta_b200_plan_upload_rows_async(gcta_b200_plan* p, const uint8_t* bed_host, int64_t row_begin, int64_t row_end, void* stream,
                                                char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && bed_host, "plan_upload_rows_async: null argument");
    GB_ARG(!p->stream_bed, "plan_upload_rows_async: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    GB_ARG(row_begin >= 0 && row_begin <= row_end && row_end <= p->m_file, "plan_upload_rows_async: bad row range");
    GB_CUDA(cudaSetDevice(p->device));
    if (row_end > row_begin)
        GB_CUDA(cudaMemcpy2DAsync(p->d_bed + row_begin * p->pitch, (size_t)p->pitch, bed_host + row_begin * p->stride, (size_t)p->stride,
                                  (size_t)p->stride, (size_t)(row_end - row_begin), cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}

extern "C" int gcta_b200_plan_set_chunk_events(gcta_b200_plan* p, void* const* events, int64_t count, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && (events || count == 0), "plan_set_chunk_events: null argument");
    GB_ARG(count == 0 || count == gcta_b200_plan_num_chunks(p), "plan_set_chunk_events: one event per SNP chunk (or none)");
    for (int64_t i = 0; i < count; i++) GB_ARG(events[i], "plan_set_chunk_events: null event");
    p->chunk_ready.clear();
    for (int64_t i = 0; i < count; i++) p->chunk_ready.push_back((cudaEvent_t)events[i]);
    return 0;
}

extern "C" int gcta_b200_plan_set_chunk_hook(gcta_b200_plan* p, int (*hook)(void*, int64_t, void*), void* ctx)
{
    if (!p) return GCTA_B200_ERR_ARG;
    p->chunk_hook = hook;
    p->chunk_hook_ctx = ctx;
    return 0;
}

extern "C" void* gcta_b200_plan_own_bed(gcta_b200_plan* p) { return p ? p->d_bed_own : nullptr; }

extern "C" int gcta_b200_plan_upload_rows(gcta_b200_plan* p, const uint8_t* bed_host, int64_t row_begin, int64_t row_end, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && bed_host, "plan_upload_rows: null argument");
    GB_ARG(!p->stream_bed, "plan_upload_rows: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    GB_ARG(row_begin >= 0 && row_begin <= row_end && row_end <= p->m_file, "plan_upload_rows: bad row range");
    GB_CUDA(cudaSetDevice(p->device));
    if (row_end > row_begin)
        GB_CUDA(cudaMemcpy2DAsync(p->d_bed + row_begin * p->pitch, (size_t)p->pitch, bed_host + row_begin * p->stride, (size_t)p->stride,
                                  (size_t)p->stride, (size_t)(row_end - row_begin), cudaMemcpyHostToDevice, p->stream));
    GB_CUDA(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int gcta_b200_plan_download_bed(gcta_b200_plan* p, uint8_t* bed_host, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && bed_host, "plan_download_bed: null argument");
    GB_ARG(!p->stream_bed, "plan_download_bed: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    GB_CUDA(cudaSetDevice(p->device));
    GB_CUDA(cudaMemcpy2DAsync(bed_host, (size_t)p->stride, p->d_bed, (size_t)p->pitch, (size_t)p->stride, (size_t)p->m_file,
                              cudaMemcpyDeviceToHost, p->stream));
    GB_CUDA(cudaStreamSynchronize(p->stream));
    return 0;
}

namespace gb {
__device__ __forceinline__ uint64_t mix64(uint64_t x)
{
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}
// same generator as gcta_b200/synth.py:bed_rows - one thread per packed byte
__global__ void synth_bed_kernel(uint8_t* bed, int64_t pitch, int64_t stride, int64_t n, int64_t m, uint64_t seed, const uint4* thr)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= pitch * m) return;
    const int64_t k = idx / pitch, b = idx % pitch;
    uint32_t byte = 0;
    if (b < stride) {
        const uint4 t = thr[k];
        for (int e = 0; e < 4; e++) {
            const int64_t i = b * 4 + e;
            if (i >= n) break;
            const uint64_t h = mix64(seed + (uint64_t)k * 0x9E3779B97F4A7C15ull + (uint64_t)i * 0xC2B2AE3D27D4EB4Full);
            const uint32_t g = (uint32_t)h, r = (uint32_t)(h >> 32);
            uint32_t code = 3;
            if (g < t.y) code = 2;
            if (g < t.x) code = 0;
            if (r < t.z) code = 1;
            byte |= code << (2 * e);
        }
    }
    bed[idx] = (uint8_t)byte;
}
}  // namespace gb

extern "C" int gcta_b200_plan_synth(gcta_b200_plan* p, uint64_t seed, const uint32_t* thr_host, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && thr_host, "plan_synth: null argument");
    GB_ARG(!p->stream_bed, "plan_synth: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    GB_CUDA(cudaSetDevice(p->device));
    uint4* d_thr = nullptr;
    GB_CUDA(cudaMalloc(&d_thr, (size_t)p->m_file * 16));
    cudaError_t e = cudaMemcpy(d_thr, thr_host, (size_t)p->m_file * 16, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const int64_t total = p->pitch * p->m_file;
        synth_bed_kernel<<<(unsigned)((total + 255) / 256), 256, 0, p->stream>>>(p->d_bed, p->pitch, p->stride, p->n_file, p->m_file, seed, d_thr);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    }
    cudaFree(d_thr);
    GB_CUDA(e);
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Pageable host genotypes.  cudaMemcpy2DAsync from pageable memory lets the driver stage the rows itself, on the
// calling thread, at a few GB/s (measured: 12.5 GB in ~3 s, twice the compute of the headline workload).  Instead, a few
// host threads copy the rows into pinned slots of their own - re-pitched on the way, so that the device copy is one
// contiguous block - and queue the host->device copies on their own streams, running ahead of the compute by the depth
// of their slots.  Per thread and SNP chunk an event marks "my slabs of this chunk are queued"; the run waits for all of
// them before it standardises the chunk (on the host until the record exists, on the device until it fires).
// ------------------------------------------------------------------------------------------------
namespace gb {

struct Stager {
    static constexpr int K = gcta_b200_plan::HELPERS, SLOTS = gcta_b200_plan::HSLOTS;
    static constexpr int64_t SLOT_BYTES = gcta_b200_plan::HSLOT_BYTES;
    gcta_b200_plan* p = nullptr;
    const uint8_t* host = nullptr;
    std::vector<int64_t> r0, r1;                        // file rows of each chunk
    uint8_t* pinned[K][SLOTS] = {};                     // the plan's h_up slots (allocated on first use, freed with the plan)
    cudaStream_t streams[K] = {};
    cudaEvent_t slot_done[K][SLOTS] = {};
    std::vector<cudaEvent_t> chunk_done[K];             // [t][c]
    std::atomic<int64_t> posted[K];                     // chunks whose event thread t has recorded
    std::atomic<int> failed{0}, stop{0};
    std::vector<std::thread> threads;

    ~Stager() { finish(); }
    void finish()
    {
        stop.store(1);
        for (auto& t : threads) t.join();
        threads.clear();
        for (int t = 0; t < K; t++) {
            if (streams[t]) cudaStreamSynchronize(streams[t]);
            for (int i = 0; i < SLOTS; i++) {
                if (slot_done[t][i]) cudaEventDestroy(slot_done[t][i]);
                slot_done[t][i] = nullptr;
            }
            for (cudaEvent_t e : chunk_done[t]) cudaEventDestroy(e);
            chunk_done[t].clear();
            if (streams[t]) cudaStreamDestroy(streams[t]);
            streams[t] = nullptr;
        }
        cudaGetLastError();
    }
    bool start(gcta_b200_plan* plan, const uint8_t* bed_host, uint8_t* d_bed, int64_t pitch, int64_t stride, const std::vector<int64_t>& rb,
               const std::vector<int64_t>& re)
    {
        p = plan; host = bed_host; r0 = rb; r1 = re;
        const int64_t nch = (int64_t)r0.size();
        for (int t = 0; t < K; t++) {
            posted[t].store(0);
            if (cudaStreamCreateWithFlags(&streams[t], cudaStreamNonBlocking) != cudaSuccess) return false;
            for (int i = 0; i < SLOTS; i++) {
                if (!plan->h_up[t][i]) {
                    if (cudaHostAlloc((void**)&plan->h_up[t][i], (size_t)SLOT_BYTES, cudaHostAllocDefault) != cudaSuccess) return false;
                    memset(plan->h_up[t][i], 0, (size_t)SLOT_BYTES);   // the pad bytes of every re-pitched row stay zero
                }
                pinned[t][i] = plan->h_up[t][i];
                if (cudaEventCreateWithFlags(&slot_done[t][i], cudaEventDisableTiming) != cudaSuccess) return false;
            }
            chunk_done[t].assign((size_t)nch, nullptr);
            for (int64_t c = 0; c < nch; c++)
                if (cudaEventCreateWithFlags(&chunk_done[t][(size_t)c], cudaEventDisableTiming) != cudaSuccess) return false;
        }
        const int64_t slab_rows = std::max<int64_t>(1, SLOT_BYTES / pitch);
        const int device = plan_device();
        for (int t = 0; t < K; t++)
            threads.emplace_back([=]() {
                bool ok = cudaSetDevice(device) == cudaSuccess;
                int slot = 0;
                bool used[SLOTS] = {};
                for (int64_t c = 0; ok && c < nch && !stop.load(); c++) {
                    int64_t j = 0;
                    for (int64_t a = r0[(size_t)c]; ok && a < r1[(size_t)c]; a += slab_rows, j++) {
                        if (j % K != t) continue;
                        const int64_t b = std::min(r1[(size_t)c], a + slab_rows);
                        if (used[slot]) ok = cudaEventSynchronize(slot_done[t][slot]) == cudaSuccess;   // its last copy has left the slot
                        uint8_t* dst = pinned[t][slot];
                        if (pitch == stride) memcpy(dst, host + a * stride, (size_t)((b - a) * stride));
                        else
                            for (int64_t r = a; r < b; r++) memcpy(dst + (r - a) * pitch, host + r * stride, (size_t)stride);
                        ok = ok && cudaMemcpyAsync(d_bed + a * pitch, dst, (size_t)((b - a) * pitch), cudaMemcpyHostToDevice, streams[t]) == cudaSuccess &&
                             cudaEventRecord(slot_done[t][slot], streams[t]) == cudaSuccess;
                        used[slot] = true;
                        slot = (slot + 1) % SLOTS;
                    }
                    ok = ok && cudaEventRecord(chunk_done[t][(size_t)c], streams[t]) == cudaSuccess;
                    if (ok) posted[t].store(c + 1, std::memory_order_release);
                }
                if (!ok) failed.store(1);
            });
        return true;
    }
    int plan_device() const;
    // queue on `s` the waits that make chunk c's rows complete
    bool wait_chunk(int64_t c, cudaStream_t s)
    {
        for (int t = 0; t < K; t++) {
            while (posted[t].load(std::memory_order_acquire) <= c) {
                if (failed.load()) return false;
                std::this_thread::yield();
            }
            if (cudaStreamWaitEvent(s, chunk_done[t][(size_t)c], 0) != cudaSuccess) return false;
        }
        return true;
    }
};
int Stager::plan_device() const { return p->device; }

}  // namespace gb

// ------------------------------------------------------------------------------------------------
// Pageable host results.  A device->host copy into pageable memory is staged by the driver on the calling thread at a
// few GB/s - slower still when the pages have never been touched (a fresh `new float[...]`), which is the usual case
// for the 5 GB of A at the headline size (measured: 1.2 s, against 1.75 s for the whole compute).  The Drainer moves a
// list of (device range -> host range, optional "ready" event) pieces with the same helper-thread scheme as the Stager:
// each thread copies its slabs into its own pinned slots on its own stream and memcpy's them out, so bus transfers,
// page faults and memcpy run in parallel.
// ------------------------------------------------------------------------------------------------
namespace gb {

struct DrainPiece { void* dst; const void* src; int64_t bytes; cudaEvent_t ready; };

// smallest transfers the helper threads take over; GCTA_B200_STAGE_MIN_MB overrides both (tests use it to reach the code
// with small inputs)
static int64_t stage_min_bytes()
{
    if (const char* e = getenv("GCTA_B200_STAGE_MIN_MB")) return atoll(e) << 20;
    return gcta_b200_plan::STAGE_MIN_BYTES;
}
static int64_t drain_min_bytes()
{
    if (const char* e = getenv("GCTA_B200_STAGE_MIN_MB")) return atoll(e) << 20;
    return gcta_b200_plan::DRAIN_MIN_BYTES;
}

static bool host_is_pageable(const void* ptr)
{
    cudaPointerAttributes attr;
    const bool pg = cudaPointerGetAttributes(&attr, ptr) != cudaSuccess || attr.type == cudaMemoryTypeUnregistered;
    cudaGetLastError();
    return pg;
}

// blocks until every piece is in host memory; false on a CUDA error
static bool drain_to_pageable(gcta_b200_plan* p, const std::vector<DrainPiece>& pieces)
{
    constexpr int K = gcta_b200_plan::HELPERS, SLOTS = gcta_b200_plan::HSLOTS;
    constexpr int64_t SLOT = gcta_b200_plan::HSLOT_BYTES;
    for (int t = 0; t < K; t++)
        for (int i = 0; i < SLOTS; i++)
            if (!p->h_down[t][i] && cudaHostAlloc((void**)&p->h_down[t][i], (size_t)SLOT, cudaHostAllocDefault) != cudaSuccess) {
                cudaGetLastError();
                return false;
            }
    struct Slab { char* dst; const char* src; int64_t bytes; cudaEvent_t ready; };
    std::vector<Slab> slabs;
    for (const DrainPiece& pc : pieces)
        for (int64_t o = 0; o < pc.bytes; o += SLOT)
            slabs.push_back({(char*)pc.dst + o, (const char*)pc.src + o, std::min(SLOT, pc.bytes - o), pc.ready});
    std::atomic<int> failed{0};
    const int device = p->device;
    std::vector<std::thread> th;
    for (int t = 0; t < K; t++)
        th.emplace_back([&, t]() {
            cudaStream_t s = nullptr;
            cudaEvent_t ev[SLOTS] = {};
            bool ok = cudaSetDevice(device) == cudaSuccess && cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess;
            for (int i = 0; ok && i < SLOTS; i++) ok = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) == cudaSuccess;
            // slab j of this thread sits in slot j % SLOTS: queue the copy of slab j+1, then move slab j out
            std::vector<size_t> mine;
            for (size_t j = (size_t)t; j < slabs.size(); j += K) mine.push_back(j);
            auto queue = [&](size_t q) {
                const Slab& sl = slabs[mine[q]];
                if (sl.ready) ok = ok && cudaStreamWaitEvent(s, sl.ready, 0) == cudaSuccess;
                ok = ok && cudaMemcpyAsync(p->h_down[t][q % SLOTS], sl.src, (size_t)sl.bytes, cudaMemcpyDeviceToHost, s) == cudaSuccess &&
                     cudaEventRecord(ev[q % SLOTS], s) == cudaSuccess;
            };
            if (ok && !mine.empty()) queue(0);
            for (size_t q = 0; ok && q < mine.size(); q++) {
                if (q + 1 < mine.size()) queue(q + 1);
                ok = ok && cudaEventSynchronize(ev[q % SLOTS]) == cudaSuccess;
                if (ok) memcpy(slabs[mine[q]].dst, p->h_down[t][q % SLOTS], (size_t)slabs[mine[q]].bytes);
            }
            if (s) cudaStreamSynchronize(s);
            for (int i = 0; i < SLOTS; i++)
                if (ev[i]) cudaEventDestroy(ev[i]);
            if (s) cudaStreamDestroy(s);
            if (!ok) failed.store(1);
        });
    for (auto& t : th) t.join();
    cudaGetLastError();
    return failed.load() == 0;
}

}  // namespace gb

// Last SNP chunk of a run whose results go straight to host buffers and in which nothing is missing: the tile list
// is cut at raster-group boundaries into a few segments; each segment's SYRK is followed by its own finalize and an
// asynchronous device->host copy of its rows on the copy stream, so the download of A hides behind the remaining
// tensor work, and N (== m everywhere) is filled on the host meanwhile.
static int run_last_chunk_pipelined(gcta_b200_plan* p, const void* Z, int64_t kp, gcta_b200_out* out, cudaStream_t st, int* launches,
                                    char* err, size_t errcap)
{
    using namespace gb;
    if (!p->stream_copy) GB_CUDA(cudaStreamCreateWithFlags(&p->stream_copy, cudaStreamNonBlocking));
    const int64_t band_base = p->row_begin * (p->row_begin + 1) / 2;
    // segment boundaries: tile indices where a new raster group of row blocks starts
    const int R = raster_rows();
    const int64_t b0 = p->row_begin / TILE_M;
    std::vector<int64_t> gstart;   // first tile index of each group
    int64_t last_g = -1;
    for (int64_t t = 0; t < p->ntiles; t++) {
        const int64_t g = (p->h_tiles[2 * t] - b0) / R;
        if (g != last_g) { gstart.push_back(t); last_g = g; }
    }
    gstart.push_back(p->ntiles);
    const int ngroups = (int)gstart.size() - 1;
    const int nseg = std::min(8, ngroups);
    std::vector<cudaEvent_t> evs;
    struct Seg { int64_t base, cnt; };
    std::vector<Seg> segs;
    int rc = 0;
    int gi = 0;
    for (int s = 0; s < nseg && !rc && gi < ngroups; s++) {
        // groups [gi, ge): about 1/nseg of the tiles each (the first groups of a triangle are small, so early segments
        // may swallow several groups and use them up before the last segment)
        const int64_t want = p->ntiles * (s + 1) / nseg;
        int ge = gi + 1;
        while (ge < ngroups && (s == nseg - 1 || gstart[ge] < want)) ge++;
        if (s == nseg - 1) ge = ngroups;
        const int64_t t0 = gstart[gi], t1 = gstart[ge];
        const int64_t r0 = (b0 + (int64_t)gi * R) * TILE_M, r1 = std::min<int64_t>((b0 + (int64_t)ge * R) * TILE_M, p->row_end);
        gi = ge;
        if (t1 <= t0) continue;
        rc = gcta_b200_syrk_tiles_chained(Z, Z, p->ldz, kp, std::min<int64_t>(p->chain, round_up(kp, 128)), p->operand, p->d_tiles + 2 * t0, t1 - t0,
                                          p->d_S + t0 * (TILE_M * TILE_N), st, err, errcap);   // TMEM chains as in plan_run_impl
        if (rc) break;
        const int64_t seg_base = r0 * (r0 + 1) / 2, seg_cnt = r1 * (r1 + 1) / 2 - seg_base;
        rc = gcta_b200_finalize_tiles(p->d_S + t0 * (TILE_M * TILE_N), nullptr, (int32_t)p->m, p->d_tiles + 2 * t0, t1 - t0, p->n, p->compact ? nullptr : p->d_keep, p->d_diag,
                                      p->d_nii, r0, r1, p->d_A_lower + (seg_base - band_base), nullptr, st, err, errcap);
        if (rc) break;
        *launches += 2;
        cudaEvent_t ev;
        cudaError_t e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        if (e == cudaSuccess) { evs.push_back(ev); e = cudaEventRecord(ev, st); }
        if (e != cudaSuccess) rc = set_err(err, errcap, GCTA_B200_ERR_CUDA, std::string("Error: CUDA call failed in the pipelined download: ") + cudaGetErrorString(e));
        segs.push_back({seg_base, seg_cnt});
    }
    // every launch of the chunk is queued: only now start the host threads that fill N (== m everywhere), so that
    // nothing competes with the launching thread for cores while the GPU waits for its next kernel
    std::vector<std::thread> fill;
    if (out->N_lower && p->prefilled_N != out->N_lower) {
        const int nt = (int)std::min<int64_t>(16, std::max<int64_t>(1, p->band_count >> 22));
        int32_t* dst = out->N_lower + band_base;
        const int32_t mval = (int32_t)p->m;
        const int64_t cnt = p->band_count;
        for (int t = 0; t < nt; t++)
            fill.emplace_back([=]() { std::fill(dst + cnt * t / nt, dst + cnt * (t + 1) / nt, mval); });
    }
    // The copies are issued only now, after every kernel of the chunk is in the queue, so the GPU keeps computing the later
    // segments whatever the host does: into pinned memory they are asynchronous copies on the copy stream; pageable memory
    // is served by the Drainer's helper threads (or, for small bands, by the driver's own staging, which blocks this thread
    // piece by piece).
    const bool drain = !rc && p->band_count * 4 >= drain_min_bytes() && host_is_pageable(out->A_lower) &&
                       !(getenv("GCTA_B200_STAGE") && atoi(getenv("GCTA_B200_STAGE")) == 0);
    if (drain) {
        std::vector<DrainPiece> pieces;
        for (size_t i = 0; i < segs.size() && i < evs.size(); i++)
            pieces.push_back({out->A_lower + segs[i].base, p->d_A_lower + (segs[i].base - band_base), segs[i].cnt * 4, evs[i]});
        if (!drain_to_pageable(p, pieces)) rc = set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: CUDA call failed in the pipelined download (pinned staging).");
    }
    for (size_t i = 0; !drain && i < segs.size() && i < evs.size() && !rc; i++) {
        cudaError_t e = cudaStreamWaitEvent(p->stream_copy, evs[i], 0);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(out->A_lower + segs[i].base, p->d_A_lower + (segs[i].base - band_base), (size_t)segs[i].cnt * 4, cudaMemcpyDeviceToHost, p->stream_copy);
        if (e != cudaSuccess) rc = set_err(err, errcap, GCTA_B200_ERR_CUDA, std::string("Error: CUDA call failed in the pipelined download: ") + cudaGetErrorString(e));
    }
    cudaError_t e2 = cudaStreamSynchronize(p->stream_copy);
    for (auto& t : fill) t.join();
    for (cudaEvent_t ev : evs) cudaEventDestroy(ev);
    if (!rc && e2 != cudaSuccess) rc = set_err(err, errcap, GCTA_B200_ERR_CUDA, std::string("Error: CUDA call failed: ") + cudaGetErrorString(e2));
    return rc;
}

// plan_run and plan_run_streamed.  With bed_host != NULL the packed rows of chunk c+1 travel host -> HBM on the
// copy stream while chunk c is standardised and contracted, so the upload hides behind the tensor work.
static int plan_run_impl(gcta_b200_plan* p, const uint8_t* bed_host, gcta_b200_out* out, bool to_host, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && out, "plan_run: null argument");
    p->downloaded = false;
    bool pipelined = false;
    GB_CUDA(cudaSetDevice(p->device));
    std::vector<cudaEvent_t>& up = p->upload_events;
    const int64_t nch_up = p->nchunks();
    if (bed_host) {
        if (!p->stream_copy) GB_CUDA(cudaStreamCreateWithFlags(&p->stream_copy, cudaStreamNonBlocking));
        while ((int64_t)up.size() < nch_up) {
            cudaEvent_t ev;
            GB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            up.push_back(ev);
        }
    }
    GB_ARG(!p->stream_bed || bed_host, "plan_run: the packed genotypes exceed their HBM allowance and are streamed from the host; use plan_run_streamed / plan_run_to_host");
    // file rows chunk c reads
    auto chunk_rows = [&](int64_t c, int64_t& r0, int64_t& r1) {
        const int64_t k0 = p->cstart[c], k1 = p->cstart[c + 1];
        r0 = k0; r1 = k1;
        if (p->has_include) {
            r0 = *std::min_element(p->h_include.begin() + k0, p->h_include.begin() + k1);
            r1 = *std::max_element(p->h_include.begin() + k0, p->h_include.begin() + k1) + 1;
        }
    };
    // host -> HBM copy of the file rows chunk c needs; issued one chunk ahead of the compute so that it overlaps
    // even when the host buffer is pageable (then the call blocks the host, but the GPU is already busy).
    // Resident genotypes: into place in the whole-file buffer.  Streamed genotypes: into ring slot seq & 1, once the
    // kernels that read the slot's previous contents have finished (`seq` counts the chunks staged during this run).
    int64_t seq_next = 0;
    bool slot_used[2] = {false, false};
    auto issue_upload = [&](int64_t c) -> cudaError_t {
        if (!bed_host || c >= nch_up) return cudaSuccess;
        int64_t r0, r1;
        chunk_rows(c, r0, r1);
        uint8_t* dst = p->d_bed + r0 * p->pitch;
        if (p->stream_bed) {
            const int slot = (int)(seq_next++ & 1);
            dst = p->d_ring[slot];
            if (slot_used[slot]) {
                cudaError_t e = cudaStreamWaitEvent(p->stream_copy, p->slot_free[slot], 0);
                if (e != cudaSuccess) return e;
            }
        }
        cudaError_t e = cudaMemcpy2DAsync(dst, (size_t)p->pitch, bed_host + r0 * p->stride, (size_t)p->stride,
                                          (size_t)p->stride, (size_t)(r1 - r0), cudaMemcpyHostToDevice, p->stream_copy);
        if (e != cudaSuccess) return e;
        return cudaEventRecord(up[c], p->stream_copy);
    };
    // what the kernels of the seq-th staged chunk (chunk c) take as "the packed file": row r lives at bed + r * pitch
    auto bed_of = [&](int64_t c, int64_t seq) -> const uint8_t* {
        if (!p->stream_bed) return p->d_bed;
        int64_t r0, r1;
        chunk_rows(c, r0, r1);
        return p->d_ring[seq & 1] - r0 * p->pitch;
    };
    // the last kernel reading the seq-th staged chunk has been queued on `s`
    auto release_slot = [&](int64_t seq, cudaStream_t s) -> cudaError_t {
        if (!p->stream_bed) return cudaSuccess;
        slot_used[seq & 1] = true;
        return cudaEventRecord(p->slot_free[seq & 1], s);
    };
    // pageable host memory and a resident genotype buffer: pinned staging by helper threads (see Stager)
    std::unique_ptr<Stager> stager;
    if (bed_host && !p->stream_bed && !(getenv("GCTA_B200_STAGE") && atoi(getenv("GCTA_B200_STAGE")) == 0)) {
        cudaPointerAttributes attr;
        const bool pageable = cudaPointerGetAttributes(&attr, bed_host) != cudaSuccess || attr.type == cudaMemoryTypeUnregistered;
        cudaGetLastError();
        if (pageable && p->pitch * p->m_file >= stage_min_bytes() && p->pitch <= gcta_b200_plan::HSLOT_BYTES) {   // a slot holds whole rows
            std::vector<int64_t> rb((size_t)nch_up), re((size_t)nch_up);
            for (int64_t c = 0; c < nch_up; c++) chunk_rows(c, rb[(size_t)c], re[(size_t)c]);
            stager.reset(new Stager());
            if (!stager->start(p, bed_host, p->d_bed, p->pitch, p->stride, rb, re)) stager.reset();   // fall back to the driver's staging
        }
    }
    if (!stager) GB_CUDA(issue_upload(0));
    // Two streams: `sd` standardises chunk c+1 (K1, K1d - HBM / ALU work) while `st` runs the SYRK of
    // chunk c on the tensor cores; two Z buffers alternate.  Ordering is by events only.
    // Measured on cfg3 (profiles/README.md): co-running the decode with the SYRK steals issue slots and power
    // from the tensor loop (K2 +30 %), so by default both run on one stream; GCTA_B200_OVERLAP=1 enables it.
    cudaStream_t st = p->stream, sd = p->overlap ? p->stream2 : p->stream;
    const int64_t nchunks = p->nchunks();
    // events: [0] start, [1] buffers cleared, then 6 per chunk (decode start/end, syrk start/end, counts start/end), then 2
    const size_t need = (size_t)(6 * nchunks + 4);
    while (p->events.size() < need) {
        cudaEvent_t ev;
        GB_CUDA(cudaEventCreate(&ev));
        p->events.push_back(ev);
    }
    auto EV = [&](int64_t c, int which) { return p->events[(size_t)(2 + 6 * c + which)]; };
    int launches = 0;
    // K2 dispatch: out (+)= Za^T Zb over the band's tiles, as single CTAs (syrk.cu) or as CTA pairs (syrk2.cu)
    auto syrk = [&](const void* Za, const void* Zb, int64_t kp, int op, void* outp, bool int_acc, cudaStream_t s) -> int {
        if (p->d_pairs) {   // CTA-pair kernel: one launch per TMEM chain (counts need no chains: exact)
            const int64_t step = int_acc ? kp : p->chain, eb = op == GCTA_B200_E4M3 ? 1 : 2;
            for (int64_t ks = 0; ks < kp; ks += step) {
                int rc = gcta_b200_syrk_pairs((const uint8_t*)Za + ks * p->ldz * eb, (const uint8_t*)Zb + ks * p->ldz * eb, p->ldz, std::min<int64_t>(step, kp - ks), op,
                                              p->d_pairs, p->npairs, outp, int_acc, s, err, errcap);
                if (rc) return rc;
            }
            return 0;
        }
        if (int_acc) return gcta_b200_count_tiles(Za, p->ldz, kp, op, p->d_tiles, p->ntiles, (int32_t*)outp, s, err, errcap);
        // TMEM chains of p->chain SNPs (see plan_create), all inside one launch
        int rc = gcta_b200_syrk_tiles_chained(Za, Zb, p->ldz, kp, std::min<int64_t>(p->chain, round_up(kp, 128)), op, p->d_tiles, p->ntiles,
                                              (float*)outp, s, err, errcap);
        if (rc) return rc;
        return 0;
    };
    GB_CUDA(cudaEventRecord(p->events[0], st));
    GB_CUDA(cudaMemsetAsync(p->d_S, 0, (size_t)p->ntiles * TILE_M * TILE_N * sizeof(float), st));
    GB_CUDA(cudaMemsetAsync(p->d_diag, 0, (size_t)p->n_file * 8, st));
    GB_CUDA(cudaMemsetAsync(p->d_nii, 0, (size_t)p->n_file * 4, st));
    GB_CUDA(cudaMemsetAsync(p->d_total_missing, 0, 8, st));
    if (p->d_dw) GB_CUDA(cudaMemsetAsync(p->d_dw, 0, (size_t)p->n_file * 8, st));
    GB_CUDA(cudaEventRecord(p->events[1], st));
    GB_CUDA(cudaStreamWaitEvent(sd, p->events[1], 0));
    // Pair counts chunk by chunk (grm.cpp:270, 291): when the first chunk already shows missing genotypes, K1f writes the
    // E4M3 indicators of each chunk next to its Z in the same pass over the packed rows, and the counting SYRK of the chunk
    // follows its S SYRK at once - no second decode of the file (and, for streamed genotypes, no second trip over PCIe).
    // If missing genotypes only turn up in a later chunk, the separate indicator pass after the loop does the counting.
    const char* nmode = getenv("GCTA_B200_N_MODE");
    const bool inline_ok = p->fused_decode && !p->d_pairs && !p->overlap && p->ntiles > 0 && !(nmode && *nmode) &&
                           !(getenv("GCTA_B200_COUNT_INLINE") && atoi(getenv("GCTA_B200_COUNT_INLINE")) == 0);
    bool count_inline = false;
    void* okbuf = nullptr;
    // Results that go to host buffers: when nothing is missing N_ij == m for every pair (what grm.cpp:291 yields), filled
    // by host threads.  They start as soon as the first chunk is queued - not behind the last chunk, whose compute is too
    // short to hide 0.6 - 5 GB of host writes once several GPUs share the work (measured: +55 ms on a 233 ms step at
    // 8 GPUs) - and are joined before this call returns.  If genotypes turn out to be missing, the real counts are
    // downloaded over the fill afterwards.
    struct FillGuard {
        std::vector<std::thread> th;
        ~FillGuard() { for (auto& t : th) t.join(); }
    } nfill;
    p->prefilled_N = nullptr;
    for (int64_t c = 0; c < nchunks; c++) {
        const int64_t k0 = p->cstart[c], kc = p->cstart[c + 1] - k0;
        void* Z = (p->overlap && p->d_Zbuf[1]) ? p->d_Zbuf[c & 1] : p->d_Zbuf[0];
        if (c >= 2) GB_CUDA(cudaStreamWaitEvent(sd, EV(c - 2, 3), 0));   // the SYRK that last read this Z buffer
        if (stager) {
            if (!stager->wait_chunk(c, sd)) return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: plan_run: staging the packed genotypes through pinned memory failed.");
        } else if (bed_host) GB_CUDA(cudaStreamWaitEvent(sd, up[c], 0));        // this chunk's packed rows have landed
        if (!bed_host && c < (int64_t)p->chunk_ready.size() && p->chunk_ready[c])   // ... or were delivered by the caller
            GB_CUDA(cudaStreamWaitEvent(sd, p->chunk_ready[c], 0));                  // (e.g. an NVLink all-gather per chunk)
        if (!bed_host && p->chunk_hook && p->chunk_hook(p->chunk_hook_ctx, c, sd) != 0)
            return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: plan_run: the delivery of a chunk's packed rows failed.");
        GB_CUDA(cudaEventRecord(EV(c, 0), sd));
        int rc = 0;
        const uint8_t* bedc = bed_of(c, c);   // first pass: chunk c is the c-th staged chunk
        if (p->fused_decode) {
            // what K1s / K1f take as the packed file: the file itself, or this chunk compacted to the kept individuals
            const uint8_t* fb = bedc;
            int64_t fpitch = p->pitch, fn = p->n_file;
            const uint32_t* fkm = p->d_keepmask;
            const int32_t* finc = p->d_include;
            if (p->compact) {
                rc = gcta_b200_compact_rows(bedc, p->pitch, p->d_include, k0, kc, p->d_keep, p->n, p->d_cbed, p->cpitch, sd, err, errcap);
                if (rc) return rc;
                launches++;
                fb = p->d_cbed - k0 * p->cpitch;   // row k of the "file" = row k - k0 of the chunk
                fpitch = p->cpitch; fn = p->n; fkm = p->d_ckeepmask; finc = nullptr;
            }
            rc = gcta_b200_snp_stats(fb, fpitch, fn, fkm, p->n, finc, k0, kc, p->d_mu_in, p->alg, p->d_mu,
                                     p->d_cnt, p->d_sum, p->d_zlut, p->d_total_missing, sd, err, errcap);
            if (!rc && c == 0 && inline_ok) {
                // anything missing in the first chunk?  (one small wait at the very start of the run)
                unsigned long long tm0 = 0;
                GB_CUDA(cudaMemcpyAsync(&tm0, p->d_total_missing, 8, cudaMemcpyDeviceToHost, sd));
                GB_CUDA(cudaStreamSynchronize(sd));
                if (tm0 > 0) {
                    if (!p->d_OKbuf) rc = dev_alloc(p, &p->d_OKbuf, (size_t)round_up(p->chunk, 128) * p->ldz, err, errcap);
                    if (!rc && !p->d_Ntiles) rc = dev_alloc(p, &p->d_Ntiles, (size_t)p->ntiles * TILE_M * TILE_N, err, errcap);
                    if (!rc && !p->d_N_lower) rc = dev_alloc(p, &p->d_N_lower, (size_t)p->band_count, err, errcap);
                    if (rc) return rc;
                    GB_CUDA(cudaMemsetAsync(p->d_Ntiles, 0, (size_t)p->ntiles * TILE_M * TILE_N * 4, st));
                    count_inline = true;
                    okbuf = p->d_OKbuf;
                }
            }
            // Z and the band's diagonal (and the indicators) in one pass over the packed rows
            if (!rc) rc = gcta_b200_decode_emit_diag(fb, fpitch, fn, finc, k0, kc, p->d_zlut, p->operand, Z, p->ldz,
                                                     p->d_diag, p->d_nii, p->diag_f0, p->diag_f1, okbuf, sd, err, errcap);
            if (rc) return rc;
        } else {
        rc = gcta_b200_bed_decode_ex(bedc, p->pitch, p->n_file, p->d_keepmask, p->d_keep, p->n, p->d_sexmale, p->d_include, k0, kc,
                                         p->d_mu_in, p->xchr, p->alg, p->operand, Z, p->ldz, p->d_mu, p->d_cnt, p->d_sum, p->d_zlut,
                                         p->d_dk, p->d_total_missing, sd, err, errcap);
        if (rc) return rc;
        // the diagonal is only needed for this band's own individuals: hand K1d that slice of columns
        rc = p->diag_f1 > p->diag_f0 ? gcta_b200_diag_accum_ex(bedc + p->diag_f0 / 4, p->pitch, p->diag_f1 - p->diag_f0, p->d_sexmale ? p->d_sexmale + p->diag_f0 / 16 : nullptr,
                                                                p->d_include, k0, kc, p->d_zlut, p->d_diag + p->diag_f0, p->d_nii + p->diag_f0, p->d_dk,
                                                                p->d_dw ? p->d_dw + p->diag_f0 : nullptr, sd, err, errcap) : 0;
        if (rc) return rc;
        }
        GB_CUDA(cudaEventRecord(EV(c, 1), sd));
        GB_CUDA(release_slot(c, sd));
        GB_CUDA(cudaStreamWaitEvent(st, EV(c, 1), 0));
        GB_CUDA(cudaEventRecord(EV(c, 2), st));
        if (to_host && c == nchunks - 1 && out->A_lower && p->operand != GCTA_B200_FP16X2 && !p->d_pairs && p->alg == 0 && p->ntiles > 0) {
            // results go to host buffers: if nothing is missing, pipeline finalize + download behind this last SYRK
            unsigned long long tmq = 1;
            GB_CUDA(cudaMemcpyAsync(&tmq, p->d_total_missing, 8, cudaMemcpyDeviceToHost, sd));
            GB_CUDA(cudaStreamSynchronize(sd));
            pipelined = tmq == 0;
        }
        if (pipelined) {
            rc = run_last_chunk_pipelined(p, Z, round_up(kc, 64), out, st, &launches, err, errcap);
        } else if (p->operand == GCTA_B200_FP16X2) {
            // z = hi + lo: S += hi'hi + hi'lo + lo'hi (lo'lo ~ 2^-22 z^2 is dropped)
            const int64_t kp = round_up(kc, 64);
            const uint16_t* hi = (const uint16_t*)Z;
            const uint16_t* lo = hi + kp * p->ldz;
            rc = syrk(hi, hi, kp, GCTA_B200_FP16, p->d_S, false, st);
            if (!rc) rc = syrk(hi, lo, kp, GCTA_B200_FP16, p->d_S, false, st);
            if (!rc) rc = syrk(lo, hi, kp, GCTA_B200_FP16, p->d_S, false, st);
            launches += 2;
        } else
            rc = syrk(Z, Z, round_up(kc, 64), p->operand, p->d_S, false, st);
        if (rc) return rc;
        GB_CUDA(cudaEventRecord(EV(c, 3), st));
        launches += 3;
        GB_CUDA(cudaEventRecord(EV(c, 4), st));
        if (count_inline) {
            rc = gcta_b200_count_tiles(okbuf, p->ldz, round_up(kc, 128), GCTA_B200_E4M3, p->d_tiles, p->ntiles, p->d_Ntiles, st, err, errcap);
            if (rc) return rc;
            launches++;
        }
        GB_CUDA(cudaEventRecord(EV(c, 5), st));
        if (c == 0 && !pipelined && to_host && out->N_lower && p->band_count > 0) {
            const int nt = (int)std::min<int64_t>(8, std::max<int64_t>(1, p->band_count >> 22));
            int32_t* dst = out->N_lower + p->row_begin * (p->row_begin + 1) / 2;
            const int32_t mval = (int32_t)p->m;
            const int64_t cnt = p->band_count;
            for (int t = 0; t < nt; t++)
                nfill.th.emplace_back([=]() { std::fill(dst + cnt * t / nt, dst + cnt * (t + 1) / nt, mval); });
            p->prefilled_N = out->N_lower;
        }
        if (!stager) GB_CUDA(issue_upload(c + 1));
    }
    if (stager) stager->finish();
    unsigned long long tm = 0;
    GB_CUDA(cudaMemcpyAsync(&tm, p->d_total_missing, 8, cudaMemcpyDeviceToHost, st));
    GB_CUDA(cudaStreamSynchronize(st));
    p->total_missing = (int64_t)tm;
    const size_t e_pc = 2 + 6 * nchunks, e_fin = e_pc + 1;
    // N_ij is only needed when something is missing (otherwise N == m, which is what grm.cpp:291 yields).
    // Default: indicator matrix + the SYRK kernel in counting mode (exact, tensor rate).  GCTA_B200_N_MODE=popc
    // selects the AND+popc kernel over bit planes instead (same integers, ~8x slower; kept as the cross-check).
    const int op16 = p->operand == GCTA_B200_FP16X2 ? GCTA_B200_FP16 : p->operand;   // indicator matrices are single-plane
    const bool use_popc = nmode && strcmp(nmode, "popc") == 0;
    if (tm > 0 && !p->d_Ntiles) {
        int rc = dev_alloc(p, &p->d_Ntiles, (size_t)p->ntiles * TILE_M * TILE_N, err, errcap);
        if (rc) return rc;
        rc = dev_alloc(p, &p->d_N_lower, (size_t)p->band_count, err, errcap);
        if (rc) return rc;
    }
    GB_ARG(!(p->stream_bed && tm > 0 && use_popc), "plan_run: GCTA_B200_N_MODE=popc needs resident genotypes");
    // Streamed genotypes: a pass that reads the packed rows again (the indicator passes below) stages every chunk a second
    // time through the ring, one chunk ahead of the kernels like the first pass.
    auto restage_begin = [&]() -> cudaError_t { return p->stream_bed ? issue_upload(0) : cudaSuccess; };
    auto restage_wait = [&](int64_t c) -> cudaError_t { return p->stream_bed ? cudaStreamWaitEvent(st, up[c], 0) : cudaSuccess; };
    auto restage_next = [&](int64_t c, int64_t seq) -> cudaError_t {
        cudaError_t e = release_slot(seq, st);
        return e != cudaSuccess ? e : (p->stream_bed ? issue_upload(c + 1) : cudaSuccess);
    };
    if (tm > 0 && !use_popc && !count_inline) {
        GB_CUDA(cudaMemsetAsync(p->d_Ntiles, 0, (size_t)p->ntiles * TILE_M * TILE_N * 4, st));
        const int64_t seq0 = seq_next;
        GB_CUDA(restage_begin());
        for (int64_t c = 0; c < nchunks; c++) {   // the Z buffer is free again: reuse it for the indicator chunks
            const int64_t k0 = p->cstart[c], kc = p->cstart[c + 1] - k0;
            GB_CUDA(restage_wait(c));
            // 8-bit E4M3 indicators (kind::f8f6f4: twice the MMA rate, half the bytes) unless the CTA-pair kernel or
            // GCTA_B200_N_MODE=tensor16 asks for the 16-bit route; both give the same integers
            const int opN = (p->d_pairs || (nmode && strcmp(nmode, "tensor16") == 0)) ? op16 : GCTA_B200_E4M3;
            int rc = 0;
            if (p->compact) {   // the indicator kernel too runs on the compacted chunk
                rc = gcta_b200_compact_rows(bed_of(c, seq0 + c), p->pitch, p->d_include, k0, kc, p->d_keep, p->n, p->d_cbed, p->cpitch, st, err, errcap);
                if (!rc) rc = gcta_b200_nonmiss_matrix(p->d_cbed - k0 * p->cpitch, p->cpitch, p->n, nullptr, p->n, nullptr, k0, kc, opN,
                                                        p->d_Zbuf[0], p->ldz, st, err, errcap);
                launches++;
            } else
                rc = gcta_b200_nonmiss_matrix(bed_of(c, seq0 + c), p->pitch, p->n_file, p->d_keep, p->n, p->d_include, k0, kc, opN,
                                              p->d_Zbuf[0], p->ldz, st, err, errcap);
            if (rc) return rc;
            GB_CUDA(restage_next(c, seq0 + c));
            rc = syrk(p->d_Zbuf[0], p->d_Zbuf[0], round_up(kc, opN == GCTA_B200_E4M3 ? 128 : 64), opN, p->d_Ntiles, true, st);
            if (rc) return rc;
            launches += 2;
        }
    }
    if (tm > 0 && use_popc) {
        const int64_t nwords = (p->m + 31) / 32;
        if (!p->d_planes) {
            int rc = dev_alloc(p, &p->d_planes, (size_t)nwords * p->ldz, err, errcap);
            if (rc) return rc;
        }
        GB_CUDA(cudaMemsetAsync(p->d_Ntiles, 0, (size_t)p->ntiles * TILE_M * TILE_N * 4, st));
        // planes in slabs of at most 32768 words (grid.y limit)
        for (int64_t w0 = 0; w0 < nwords; w0 += 32768) {
            const int64_t k0 = w0 * 32, kc = std::min<int64_t>(32768 * 32, p->m - k0);
            int rc = gcta_b200_nonmiss_planes(p->d_bed, p->pitch, p->n_file, p->d_keep, p->n, p->d_include, k0, kc, p->d_planes, p->ldz, st, err, errcap);
            if (rc) return rc;
            launches++;
        }
        int rc = gcta_b200_paircount_tiles(p->d_planes, p->ldz, nwords, p->d_tiles, p->ntiles, p->d_Ntiles, st, err, errcap);
        if (rc) return rc;
        launches++;
    }
    // --grm-alg 1: denominators sum_k ok_ik ok_jk d_k.  Nothing missing: the constant sum_k d_k, added on the host in
    // SNP order like grm.cpp:283.  Otherwise one more SYRK pass over ok * sqrt(d_k) into float tiles.
    double d_const = 0.0;
    if (p->alg == 1) {
        std::vector<double> dk((size_t)p->m);
        GB_CUDA(cudaMemcpyAsync(dk.data(), p->d_dk, (size_t)p->m * 8, cudaMemcpyDeviceToHost, st));
        GB_CUDA(cudaStreamSynchronize(st));
        for (int64_t k = 0; k < p->m; k++) d_const += dk[(size_t)k];
        if (tm > 0) {
            if (!p->d_Dtiles) {
                int rc = dev_alloc(p, &p->d_Dtiles, (size_t)p->ntiles * TILE_M * TILE_N, err, errcap);
                if (rc) return rc;
            }
            GB_CUDA(cudaMemsetAsync(p->d_Dtiles, 0, (size_t)p->ntiles * TILE_M * TILE_N * 4, st));
            const int64_t seq0 = seq_next;
            GB_CUDA(restage_begin());
            for (int64_t c = 0; c < nchunks; c++) {
                const int64_t k0 = p->cstart[c], kc = p->cstart[c + 1] - k0;
                GB_CUDA(restage_wait(c));
                int rc = gcta_b200_nonmiss_matrix_ex(bed_of(c, seq0 + c), p->pitch, p->n_file, p->d_keep, p->n, p->d_include, k0, kc, op16, p->d_dk,
                                                     p->d_Zbuf[0], p->ldz, st, err, errcap);
                if (rc) return rc;
                GB_CUDA(restage_next(c, seq0 + c));
                rc = syrk(p->d_Zbuf[0], p->d_Zbuf[0], round_up(kc, 64), op16, p->d_Dtiles, false, st);
                if (rc) return rc;
                launches += 2;
            }
        }
    }
    GB_CUDA(cudaEventRecord(p->events[e_pc], st));
    if (pipelined) {   // A is on the host already, N filled
        p->downloaded = true;
        p->streamed_A = out->A_lower;
        p->streamed_N = out->N_lower;
    }
    else {
        int rc = gcta_b200_finalize_tiles_ex(p->d_S, tm > 0 ? p->d_Ntiles : nullptr, (int32_t)p->m, (p->alg == 1 && tm > 0) ? p->d_Dtiles : nullptr,
                                             d_const, p->alg == 1 ? p->d_dw : nullptr, p->d_tiles, p->ntiles, p->n, p->compact ? nullptr : p->d_keep,
                                             p->d_diag, p->d_nii, p->row_begin, p->row_end, p->d_A_lower, tm > 0 ? p->d_N_lower : nullptr, st, err, errcap);
        if (rc) return rc;
        launches++;
    }
    GB_CUDA(cudaEventRecord(p->events[e_fin], st));
    GB_CUDA(cudaStreamSynchronize(st));
    // decode and SYRK overlap, so the phase times are per-stream busy times and do not add up to ms_total
    out->ms_decode = out->ms_syrk = 0.f;
    for (int64_t c = 0; c < nchunks; c++) {
        float a = 0.f, b = 0.f;
        GB_CUDA(cudaEventElapsedTime(&a, EV(c, 0), EV(c, 1)));
        GB_CUDA(cudaEventElapsedTime(&b, EV(c, 2), EV(c, 3)));
        out->ms_decode += a;
        out->ms_syrk += b;
    }
    GB_CUDA(cudaEventElapsedTime(&out->ms_paircount, EV(nchunks - 1, 5), p->events[e_pc]));
    for (int64_t c = 0; c < nchunks; c++) {   // counts that ran chunk by chunk
        float a = 0.f;
        GB_CUDA(cudaEventElapsedTime(&a, EV(c, 4), EV(c, 5)));
        out->ms_paircount += a;
    }
    GB_CUDA(cudaEventElapsedTime(&out->ms_finalize, p->events[e_pc], p->events[e_fin]));
    GB_CUDA(cudaEventElapsedTime(&out->ms_total, p->events[0], p->events[e_fin]));
    out->launches = launches;
    out->total_missing = p->total_missing;
    out->row_begin = p->row_begin;
    out->row_end = p->row_end;
    p->ran = true;
    return 0;
}

extern "C" int gcta_b200_plan_run(gcta_b200_plan* p, gcta_b200_out* out, char* err, size_t errcap)
{
    return plan_run_impl(p, nullptr, out, false, err, errcap);
}

extern "C" int gcta_b200_plan_run_streamed(gcta_b200_plan* p, const uint8_t* bed_host, gcta_b200_out* out, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(bed_host, "plan_run_streamed: null bed");
    return plan_run_impl(p, bed_host, out, false, err, errcap);
}

extern "C" int gcta_b200_plan_run_to_host(gcta_b200_plan* p, const uint8_t* bed_host, gcta_b200_out* out, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(out && out->A_lower, "plan_run_to_host: null argument");   // bed_host == NULL: genotypes already resident
    int rc = plan_run_impl(p, bed_host, out, true, err, errcap);
    if (!rc) rc = gcta_b200_plan_download(p, out, err, errcap);   // whatever the pipelined path has not delivered yet
    return rc;
}

extern "C" int gcta_b200_plan_device_results(gcta_b200_plan* p, const float** dA, const int32_t** dN, int64_t* count)
{
    if (!p || !p->ran) return GCTA_B200_ERR_ARG;
    if (dA) *dA = p->d_A_lower;
    if (dN) *dN = p->total_missing > 0 ? p->d_N_lower : nullptr;
    if (count) *count = p->band_count;
    return 0;
}

extern "C" int gcta_b200_plan_download(gcta_b200_plan* p, gcta_b200_out* out, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(p && out && out->A_lower, "plan_download: null argument");
    GB_ARG(p->ran, "plan_download: plan_run has not completed");
    GB_CUDA(cudaSetDevice(p->device));
    const int64_t base = p->row_begin * (p->row_begin + 1) / 2;
    // plan_run_to_host already delivered A (and N) behind the last SYRK - into these very buffers?
    const bool have = p->downloaded && p->streamed_A == out->A_lower && (!out->N_lower || p->streamed_N == out->N_lower);
    // large pageable destinations go through the Drainer's helper threads (everything on p->stream has completed: plan_run
    // synchronises before it returns)
    const bool big = p->band_count * 4 >= drain_min_bytes() && !(getenv("GCTA_B200_STAGE") && atoi(getenv("GCTA_B200_STAGE")) == 0);
    std::vector<DrainPiece> pieces;
    if (!have) {
        if (big && host_is_pageable(out->A_lower)) pieces.push_back({out->A_lower + base, p->d_A_lower, p->band_count * 4, nullptr});
        else GB_CUDA(cudaMemcpyAsync(out->A_lower + base, p->d_A_lower, (size_t)p->band_count * 4, cudaMemcpyDeviceToHost, p->stream));
    }
    if (out->N_lower && p->total_missing > 0) {
        if (big && host_is_pageable(out->N_lower)) pieces.push_back({out->N_lower + base, p->d_N_lower, p->band_count * 4, nullptr});
        else GB_CUDA(cudaMemcpyAsync(out->N_lower + base, p->d_N_lower, (size_t)p->band_count * 4, cudaMemcpyDeviceToHost, p->stream));
    }
    if (!pieces.empty() && !drain_to_pageable(p, pieces))
        return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: CUDA call failed in plan_download (pinned staging).");
    if (out->mu_out) GB_CUDA(cudaMemcpyAsync(out->mu_out, p->d_mu, (size_t)p->m * 8, cudaMemcpyDeviceToHost, p->stream));
    if (out->N_lower && p->total_missing == 0 && !have && p->prefilled_N != out->N_lower) {
        // nothing missing: N_ij == m for every pair (what grm.cpp:291 yields); fill on the host while the copies run
        const int nt = (int)std::min<int64_t>(16, std::max<int64_t>(1, p->band_count >> 22));
        std::vector<std::thread> th;
        int32_t* dst = out->N_lower + base;
        const int32_t mval = (int32_t)p->m;
        for (int t = 0; t < nt; t++)
            th.emplace_back([=]() {
                const int64_t lo = p->band_count * t / nt, hi = p->band_count * (t + 1) / nt;
                std::fill(dst + lo, dst + hi, mval);
            });
        for (auto& t : th) t.join();
    }
    GB_CUDA(cudaStreamSynchronize(p->stream));
    out->row_begin = p->row_begin;
    out->row_end = p->row_end;
    out->total_missing = p->total_missing;
    return 0;
}

extern "C" int gcta_b200_make_grm(const gcta_b200_args* a, gcta_b200_out* out, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(a && out && a->bed && out->A_lower, "make_grm: null argument");
    gcta_b200_plan* p = nullptr;
    int rc = gcta_b200_plan_create(a, &p, err, errcap);
    if (rc) return rc;
    cudaEvent_t e0, e1, e2, e3;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2); cudaEventCreate(&e3);
    cudaEventRecord(e0, p->stream);
    cudaEventRecord(e1, p->stream);   // the upload is streamed chunk by chunk inside the run (ms_h2d ~ 0)
    // upload, compute and download pipelined: plan_run_to_host
    if (!rc) rc = gcta_b200_plan_run_to_host(p, a->bed, out, err, errcap);
    cudaEventRecord(e2, p->stream);
    cudaEventRecord(e3, p->stream);
    cudaEventSynchronize(e3);
    if (!rc) {
        cudaEventElapsedTime(&out->ms_h2d, e0, e1);
        cudaEventElapsedTime(&out->ms_d2h, e2, e3);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); cudaEventDestroy(e3);
    gcta_b200_plan_destroy(p);
    return rc;
}
