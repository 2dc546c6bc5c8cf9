/*
 * gcta_b200.h - C ABI of the B200-native (sm_100a) replacement for GCTA's `--make-grm` hot path.
 *
 * The reference (GCTA 0.93.9, cooljeanius/gcta) has no plugin / FFI interface; the seam this
 * library fills is the C++ member  void gcta::make_grm(bool grm_xchr_flag, bool output_bin,
 * int grm_mtd)  (gcta.h:83, defined grm.cpp:219-301, called from option.cpp:620-621).
 * A maintainer keeps the reference's readers, filters and writers and replaces the body of
 * make_grm by one call of gcta_b200_make_grm() (see INTEGRATION.md for the stub).
 *
 * Conventions
 *  - plain C types only; no exceptions cross the boundary.  Every function returns 0 on
 *    success or a non-zero gcta_b200_status and, where an `err` buffer is passed, writes a
 *    NUL-terminated message into it.  The reference reports errors as throw(std::string)
 *    caught in main (option.cpp:42-44); the host shim rethrows `err` as std::string.
 *  - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *    GCTA_B200_ERR_NO_DEVICE.
 *  - "included SNP k" / "kept individual i" follow the reference's _include / _keep index
 *    lists (gcta.h:216-238): ascending indices into the .bim / .fam order.
 *  - the packed lower triangle uses the .grm.gz line order  idx(i,j) = i(i+1)/2 + j, j <= i
 *    (grm.cpp:328-330).
 */
#ifndef GCTA_B200_H
#define GCTA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    GCTA_B200_OK = 0,
    GCTA_B200_ERR_ARG = 1,       /* bad argument (message says which) */
    GCTA_B200_ERR_NO_DEVICE = 2, /* no usable CUDA device / driver */
    GCTA_B200_ERR_CUDA = 3,      /* a CUDA call or kernel failed */
    GCTA_B200_ERR_NOMEM = 4      /* device or pinned-host allocation failed */
} gcta_b200_status;

/* operand type of the standardised genotypes fed to the tensor cores (FP32 accumulate). */
typedef enum {
    GCTA_B200_FP16 = 0, /* default: 11-bit significand, |z| <= 65504 (MAF >= 1e-9); with fewer than 8192 included
                           SNPs the plan upgrades it to GCTA_B200_FP16X2 (rounding no longer averages out) */
    GCTA_B200_BF16 = 1, /* 8-bit significand; upgraded to GCTA_B200_FP16X2 below 8192 included SNPs like FP16 */
    /* "tight" mode: z = hi + lo with hi = fp16(z), lo = fp16(z - hi); S = hi'hi + hi'lo + lo'hi on the tensor
     * cores (3x the SYRK work, twice the Z traffic).  Products are then exact to ~2^-22 relative, so A agrees
     * with the reference to FP32-accumulation accuracy (max|dA| ~1e-7).  The north-star's "TF32 mode for tighter
     * agreement": TF32 itself has FP16's 10-bit mantissa and would not be tighter than the default. */
    GCTA_B200_FP16X2 = 2,
    /* 8-bit E4M3 0/1 indicators: only for gcta_b200_nonmiss_matrix + gcta_b200_count_tiles (the N pass), where the
     * tensor cores run kind::f8f6f4 at twice the 16-bit rate on half the bytes; counts stay exact. */
    GCTA_B200_E4M3 = 3
} gcta_b200_operand;

/* ------------------------------------------------------------------------------------------
 * Whole-path entry point: replaces the body of gcta::make_grm (grm.cpp:219-301) up to, but
 * not including, output_grm_vec (grm.cpp:303-341).  Host pointers in, host pointers out.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    /* .bed payload after the 3 magic bytes (data.cpp:164 skips them unchecked): m_file rows of
     * ceil(n_file/4) bytes, SNP-major, 2 bits per individual LSB first (data.cpp:171-185). */
    const uint8_t* bed;
    int64_t n_file;         /* individuals in the .fam  (_indi_num) */
    int64_t m_file;         /* SNPs in the .bim         (_snp_num)  */
    const int32_t* keep;    /* _keep: ascending file indices of kept individuals, NULL = all */
    int64_t n;              /* number kept (n_file when keep == NULL) */
    const int32_t* include; /* _include: ascending file indices of included SNPs, NULL = all */
    int64_t m;              /* number included (m_file when include == NULL) */
    const double* mu;       /* per included SNP 2p OF ALLELE 1 (.bim column 5), i.e. the mean dosage the packed codes count.
                               The reference's _mu (data.cpp:794, 915-916) is 2p of its reference allele: a caller passes
                               _mu[k] where _allele1[k] == _ref_A[k] and 2 - _mu[k] where --update-ref-allele flipped it.
                               NULL = compute like calcu_mu (data.cpp:759-803) */
    const int32_t* sex;     /* per FILE individual, 1 = male 2 = female (_sex); only read when xchr */
    int32_t xchr;           /* grm_xchr_flag: sex-weighted mu, male rows * sqrt(0.5) (data.cpp:1045-1055) */
    int32_t alg;            /* grm_mtd: 0 default, 1 = --grm-alg 1 (grm.cpp:230-237) */
    int32_t operand;        /* gcta_b200_operand */
    int32_t device;         /* CUDA device ordinal */
    int32_t part_rank;      /* this call computes row band part_rank of part_world (0 / 1 = everything) */
    int32_t part_world;
    int64_t chunk_snps;     /* SNPs standardised per pass, 0 = choose from free HBM */
} gcta_b200_args;

typedef struct {
    /* caller-allocated; the call fills rows [row_begin, row_end) of the packed lower triangle,
     * i.e. elements [row_begin(row_begin+1)/2, row_end(row_end+1)/2), and leaves the rest alone. */
    float* A_lower;          /* A_ij, grm.cpp:273-275, 294 */
    int32_t* N_lower;        /* A_N_ij, grm.cpp:270, 291; may be NULL (--make-grm-bin drops N, grm.cpp:307-318) */
    double* mu_out;          /* [m] 2p per included SNP, may be NULL */
    int64_t row_begin, row_end; /* written by the call */
    int64_t total_missing;   /* number of missing genotypes among kept x included */
    /* device-side milliseconds (CUDA events on the compute stream) */
    float ms_h2d, ms_decode, ms_syrk, ms_paircount, ms_finalize, ms_d2h, ms_total;
    int32_t launches;        /* kernels of this library launched by the call */
} gcta_b200_out;

int gcta_b200_make_grm(const gcta_b200_args* args, gcta_b200_out* out, char* err, size_t errcap);

/* The same on n_gpus devices of THIS process (args->device .. args->device + n_gpus - 1; args->part_* are ignored): the
 * lower triangle is cut into n_gpus row bands, one plan and one host thread per device.  Every device needs all packed
 * rows; instead of n_gpus full uploads, each device uploads 1/n_gpus of every row segment over its own PCIe link and
 * pushes it to the others with peer copies (copy engines over NVLink), and each band starts a SNP chunk as soon as
 * the segments it reads are complete - the seam stays the reference's one call per process from the main thread
 * (option.cpp:620-621).  out->A_lower / N_lower are filled completely; the timings in *out are the slowest device's,
 * per_gpu (NULL or n_gpus entries) receives each device's. */
int gcta_b200_make_grm_multi(const gcta_b200_args* args, int32_t n_gpus, gcta_b200_out* out, gcta_b200_out* per_gpu,
                             char* err, size_t errcap);

/* Allele counts without building a GRM: the counting half of calcu_mu / mu_func (data.cpp:759-803)
 * that --freq, --maf and --max-maf need (option.cpp:604-615).  Host pointers; per included SNP the
 * non-missing count and the allele-1 dosage sum among kept individuals.  When args->sex and all four
 * split outputs are given, the same counts restricted to kept males (sex == 1) and kept females
 * (sex == 2) are returned too, so the host can apply the X-chromosome weights of data.cpp:765-777. */
int gcta_b200_allele_counts(const gcta_b200_args* args, int32_t* cnt, int32_t* sum, int32_t* cntM, int32_t* sumM,
                            int32_t* cntF, int32_t* sumF, char* err, size_t errcap);

/* ------------------------------------------------------------------------------------------
 * Plan API: the same path split into upload / compute / download so that a caller (bench.py,
 * a pipelined host) can keep the packed genotypes resident in HBM and time the phases.
 * gcta_b200_make_grm() is plan_create + upload + run + download + destroy.
 * All device memory, streams and events are owned by the plan.
 * ---------------------------------------------------------------------------------------- */
typedef struct gcta_b200_plan gcta_b200_plan;

int gcta_b200_plan_create(const gcta_b200_args* args, gcta_b200_plan** plan, char* err, size_t errcap);
/* copy args->bed (pageable or pinned host memory) into the plan's pitched HBM buffer */
int gcta_b200_plan_upload(gcta_b200_plan* plan, const uint8_t* bed_host, char* err, size_t errcap);
/* fill the HBM genotype buffer on the device from the counter-based generator of
 * gcta_b200/synth.py (thr = uint32[m_file][4] thresholds in HOST memory); bench/test use only */
int gcta_b200_plan_synth(gcta_b200_plan* plan, uint64_t seed, const uint32_t* thr_host, char* err, size_t errcap);
/* Multi-GPU hosts (one process per GPU): instead of every rank pulling the whole .bed over PCIe, each rank uploads
 * 1/G of the SNP rows (plan_upload_rows) and the ranks all-gather the packed rows over NVLink (NCCL, by the host -
 * bench.py uses torch.distributed).  plan_use_device_bed makes the plan read its genotypes from a caller-owned device
 * buffer of m_file rows x *pitch bytes (pad bytes zero), e.g. the all-gather's output; d_bed == NULL only queries the
 * pitch.  plan_upload_rows copies host rows [row_begin,row_end) into whichever buffer the plan currently uses. */
int gcta_b200_plan_use_device_bed(gcta_b200_plan* plan, void* d_bed, int64_t* pitch, char* err, size_t errcap);
int gcta_b200_plan_upload_rows(gcta_b200_plan* plan, const uint8_t* bed_host, int64_t row_begin, int64_t row_end, char* err, size_t errcap);
/* The same per SNP chunk, so that the exchange of chunk c+1 overlaps the compute of chunk c: plan_num_chunks /
 * plan_chunk_rows give the file rows each chunk reads; plan_upload_rows_async copies rows on the caller's stream without
 * synchronising; plan_set_chunk_events hands the plan one caller-owned cudaEvent_t per chunk (recorded after that
 * chunk's rows are complete in the genotype buffer, e.g. after its all-gather) - plan_run / plan_run_to_host with
 * bed_host == NULL then wait on event c before standardising chunk c.  count == 0 clears them. */
int64_t gcta_b200_plan_num_chunks(gcta_b200_plan* plan);
int gcta_b200_plan_chunk_rows(gcta_b200_plan* plan, int64_t chunk, int64_t* row_begin, int64_t* row_end);
int gcta_b200_plan_upload_rows_async(gcta_b200_plan* plan, const uint8_t* bed_host, int64_t row_begin, int64_t row_end,
                                     void* stream, char* err, size_t errcap);
int gcta_b200_plan_set_chunk_events(gcta_b200_plan* plan, void* const* events, int64_t count, char* err, size_t errcap);
/* plan_set_chunk_events as a call-back: `hook(ctx, chunk, stream)` is called on the host right before the first kernel of
 * SNP chunk `chunk` is queued on `stream` (a cudaStream_t) by plan_run / plan_run_to_host with bed_host == NULL.  It may
 * block, and it queues on `stream` whatever waits make that chunk's file rows complete; non-zero return aborts the run.
 * plan_own_bed: device address of the plan's own genotype buffer (m_file rows x pitch; NULL when streamed). */
int gcta_b200_plan_set_chunk_hook(gcta_b200_plan* plan, int (*hook)(void* ctx, int64_t chunk, void* stream), void* ctx);
void* gcta_b200_plan_own_bed(gcta_b200_plan* plan);
/* back to the plan's own genotype buffer after plan_use_device_bed; the plan's device, file rows, row stride, row pitch */
int gcta_b200_plan_use_own_bed(gcta_b200_plan* plan);
int gcta_b200_plan_geometry(gcta_b200_plan* plan, int32_t* device, int64_t* m_file, int64_t* stride, int64_t* pitch);

/* ---- peer exchange of the packed rows between the ranks of a multi-GPU run (csrc/exchange.cu) ----
 * One process per GPU.  Each rank uploads 1/world of every SNP chunk's rows over its own PCIe link and pulls the rest
 * out of its peers' HBM with the copy engines over NVLink (CUDA IPC memory + inter-process events; no SM, no NCCL
 * kernel next to the persistent SYRK grid).  The plan then waits per chunk, so the exchange of later chunks runs
 * behind the compute of earlier ones.  No reference counterpart (the reference reads the .bed on one thread,
 * data.cpp:140-194).  Protocol, every rank:
 *   exchange_create(plan, rank, world, my_blob)        my_blob: exchange_blob_bytes(plan) bytes to publish
 *   <all-gather the blobs between the processes by any host-side means, rank order>
 *   exchange_connect(ex, all_blobs)
 *   per step:  <host barrier>  exchange_post(ex, bed_host)  <host barrier>  exchange_pull(ex)
 *              plan_run / plan_run_to_host with bed_host == NULL          exchange_finish(ex)
 *   exchange_disconnect(ex) on every rank, <host barrier>, exchange_destroy(ex) (frees the buffer the peers had
 *   mapped and puts the plan back on its own), then plan_destroy.
 * The rows are exchanged in 32 fixed segments of the file (the same on every rank); the plans may chunk the SNP axis
 * differently (a band's chunk length depends on its row_end).  All plans must have the same n_file and m_file. */
#define GCTA_B200_IPC_HANDLE_BYTES 64
typedef struct gcta_b200_exchange gcta_b200_exchange;
/* geometry only (no device): number of segments for a file of m_file rows, and the rows of segment `seg` that `rank`
 * of `world` uploads (row_begin/row_end may be NULL) */
int64_t gcta_b200_exchange_slice(int64_t m_file, int32_t world, int32_t rank, int64_t seg, int64_t* row_begin, int64_t* row_end);
int64_t gcta_b200_exchange_blob_bytes(gcta_b200_plan* plan);
int gcta_b200_exchange_create(gcta_b200_plan* plan, int32_t rank, int32_t world, uint8_t* my_blob, gcta_b200_exchange** ex,
                              char* err, size_t errcap);
int gcta_b200_exchange_connect(gcta_b200_exchange* ex, const uint8_t* all_blobs, char* err, size_t errcap);
int gcta_b200_exchange_post(gcta_b200_exchange* ex, const uint8_t* bed_host, char* err, size_t errcap);
int gcta_b200_exchange_pull(gcta_b200_exchange* ex, char* err, size_t errcap);
int gcta_b200_exchange_finish(gcta_b200_exchange* ex, char* err, size_t errcap);
void gcta_b200_exchange_disconnect(gcta_b200_exchange* ex);
void gcta_b200_exchange_destroy(gcta_b200_exchange* ex);
/* inverse of plan_upload: copy the plan's HBM genotype buffer to host memory in .bed payload layout
 * (bench/test use: obtain the host copy of a device-generated data set) */
int gcta_b200_plan_download_bed(gcta_b200_plan* plan, uint8_t* bed_host, char* err, size_t errcap);
/* decode + SYRK + pair counts + finalize on the device; results stay in HBM */
int gcta_b200_plan_run(gcta_b200_plan* plan, gcta_b200_out* out, char* err, size_t errcap);
/* plan_upload + plan_run in one call, pipelined: the packed rows of SNP chunk c+1 are copied host -> HBM on a
 * copy stream while chunk c is standardised and contracted (use pinned host memory for full overlap).  This is what
 * gcta_b200_make_grm runs. */
int gcta_b200_plan_run_streamed(gcta_b200_plan* plan, const uint8_t* bed_host, gcta_b200_out* out, char* err, size_t errcap);
/* plan_run_streamed + plan_download in one call with the download pipelined too: when nothing is missing, the last
 * SNP chunk is contracted in a few segments, each followed by its finalize and an asynchronous device->host copy of
 * its rows, so A reaches out->A_lower (and N == m is filled on the host) behind the remaining tensor work.
 * bed_host == NULL: the genotypes are already resident (plan_upload / plan_upload_rows + all-gather). */
int gcta_b200_plan_run_to_host(gcta_b200_plan* plan, const uint8_t* bed_host, gcta_b200_out* out, char* err, size_t errcap);
/* copy this plan's row band of A (and N when N_lower != NULL) to host memory */
int gcta_b200_plan_download(gcta_b200_plan* plan, gcta_b200_out* out, char* err, size_t errcap);
/* device pointers of the finalised band (float A, int32 N; N is NULL when nothing is missing
 * and N == m everywhere), valid until destroy */
int gcta_b200_plan_device_results(gcta_b200_plan* plan, const float** dA, const int32_t** dN, int64_t* count);
void gcta_b200_plan_destroy(gcta_b200_plan* plan);

/* ------------------------------------------------------------------------------------------
 * Kernel-level entry points (device pointers, caller's stream as void* = cudaStream_t).
 * Used by the parity tests to check each stage against the oracle.
 * ---------------------------------------------------------------------------------------- */

/* K1 - fused 2-bit decode + per-SNP allele frequency + standardisation.
 * Replaces read_bedfile's unpack (data.cpp:171-185), mu_func (data.cpp:781-803), make_XMat
 * (data.cpp:1001-1007) and std_XMat (data.cpp:1030-1043) for included SNPs [k0, k0+kc).
 *   d_bed      pitched packed genotypes, row r at d_bed + r*pitch, pitch % 16 == 0
 *   d_keepmask uint32[pitch/4]: bit 2t of word w set iff file individual 16w+t is kept
 *   d_keep     int32[n] or NULL when every file individual is kept (then n == n_file)
 *   d_sexmale  uint32[pitch/4] like d_keepmask, set for kept males; NULL unless xchr
 *   d_include  int32[>= k0+kc] file row of each included SNP, NULL = identity
 *   d_mu_in    double[>= k0+kc] override of 2p, or NULL
 *   d_Z        out: 16-bit standardised genotypes, SNP-major [kc_pad][ldz], kc_pad = kc rounded
 *              up to 64, ldz % 256 == 0 and ldz >= n; padding rows / columns are written as 0
 *   d_mu, d_cnt, d_sum, d_zlut   out, indexed by included SNP k (absolute): 2p, non-missing
 *              count, allele-1 dosage sum, and float[8] = z for codes 0..3 (female/autosome)
 *              then codes 0..3 (male, xchr only)
 *   d_total_missing  uint64 accumulator (+= missing genotypes of these SNPs) */
int gcta_b200_bed_decode(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                         const uint32_t* d_keepmask, const int32_t* d_keep, int64_t n,
                         const uint32_t* d_sexmale, const int32_t* d_include, int64_t k0, int64_t kc,
                         const double* d_mu_in, int32_t xchr, int32_t alg, int32_t operand,
                         void* d_Z, int64_t ldz,
                         double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut,
                         unsigned long long* d_total_missing, void* stream, char* err, size_t errcap);

/* K1s + K1f - the same stage as two streaming passes, used by the plan whenever the SNPs are autosomal, alg == 0 and the
 * operand is one 16-bit plane (with a keep list: on the rows compacted by gcta_b200_compact_rows; the other cases stay on
 * gcta_b200_bed_decode + gcta_b200_diag_accum).
 *   gcta_b200_snp_stats        per included SNP in [k0, k0+kc): allele counts of the kept individuals, 2p, and the
 *                              doubly-rounded standardised values (calcu_mu / mu_func data.cpp:759-803; the per-SNP
 *                              constants of std_XMat data.cpp:1027-1040).  Outputs as for gcta_b200_bed_decode.
 *   gcta_b200_decode_emit_diag reads each packed row once more and writes the 16-bit Z chunk (make_XMat / std_XMat,
 *                              data.cpp:982-1056) AND, in the same pass, adds sum_k z_fk^2 (float64) and the non-missing
 *                              counts of file individuals [diag_f0, diag_f1) into d_diag / d_nii (grm.cpp:265-275);
 *                              d_diag == d_nii == NULL skips the diagonal.  d_Z 32-byte aligned, layout as above.
 *                              d_OK != NULL: the same pass also writes the 8-bit E4M3 non-missing indicators (1.0 / 0) of
 *                              every genotype as a matrix [kc rounded up to 128][ldz] bytes, what gcta_b200_nonmiss_matrix
 *                              with GCTA_B200_E4M3 produces (the operand of the counting pass, grm.cpp:270, 291). */
/* K1c - a keep list applied to the packed rows (keep_indi / remove_indi, data.cpp:657-673, as read_bedfile applies them,
 * data.cpp:171-185): the 2-bit codes of kept individuals 0..n-1 of included SNPs [k0, k0+kc) packed densely, row kk at
 * d_out + kk * pitch_out.  The plan runs it per SNP chunk so that the stages above never gather per individual. */
int gcta_b200_compact_rows(const uint8_t* d_bed, int64_t pitch, const int32_t* d_include, int64_t k0, int64_t kc,
                           const int32_t* d_keep, int64_t n, uint8_t* d_out, int64_t pitch_out,
                           void* stream, char* err, size_t errcap);
int gcta_b200_snp_stats(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_keepmask, int64_t n,
                        const int32_t* d_include, int64_t k0, int64_t kc, const double* d_mu_in, int32_t alg,
                        double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut,
                        unsigned long long* d_total_missing, void* stream, char* err, size_t errcap);
int gcta_b200_decode_emit_diag(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_include,
                               int64_t k0, int64_t kc, const float* d_zlut, int32_t operand, void* d_Z, int64_t ldz,
                               double* d_diag, int32_t* d_nii, int64_t diag_f0, int64_t diag_f1, void* d_OK,
                               void* stream, char* err, size_t errcap);

/* K1d - exact per-individual diagonal: diag[f] += sum_k z_fk^2 in float64 and nii[f] += number
 * of non-missing SNPs, over included SNPs [k0, k0+kc), per FILE individual f (grm.cpp:265-275). */
int gcta_b200_diag_accum(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                         const uint32_t* d_sexmale, const int32_t* d_include, int64_t k0, int64_t kc,
                         const float* d_zlut, double* d_diag, int32_t* d_nii,
                         void* stream, char* err, size_t errcap);

/* K1d without any allocation - NO COUNTERPART IN THE REFERENCE.
 * Same result as gcta_b200_bed_decode (for its per-SNP table) followed by gcta_b200_diag_accum, for
 * autosomal SNPs [k0, k0+kc): diag[f] += sum_k z_fk^2, nii[f] += non-missing SNPs, per FILE individual.
 * For callers that cannot allocate device memory - a host that gave all of HBM to its own pool, or a
 * loop that must stay out of the CUDA allocator because cudaMalloc / cudaFree stall the device.  The
 * per-SNP table lives in a 16 MiB array built into the library (held by every GPU that loads it,
 * until the process exits); the caller passes only inputs and outputs.
 *   - at most 524288 SNPs per call; a larger kc is refused with GCTA_B200_ERR_ARG - split the range.
 *   - the first call on each device looks the array's device address up (with that device current);
 *     make that first call outside any stream capture.
 *   - the call itself calls no cudaMalloc / cudaFree or allocating library routine.
 *   - all calls share the one array: calls that can overlap on the GPU (different streams, threads or
 *     processes' streams of the same context) race on it - serialise them.  Each call rewrites the
 *     part of the table it reads, so a previous call's leftovers are harmless.
 *   - d_keepmask as for gcta_b200_bed_decode; d_diag / d_nii are accumulated into (zero them first). */
int gcta_b200_diag_accum_static(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                                const uint32_t* d_keepmask, const int32_t* d_include, int64_t k0, int64_t kc,
                                double* d_diag, int32_t* d_nii, void* stream, char* err, size_t errcap);

/* K1m - missingness bit planes for K3: d_miss[w][i] (uint32, i < ldm) has bit b set iff kept
 * individual i is NON-missing at included SNP 32*w + b; w in [k0/32, (k0+kc)/32) (k0 % 32 == 0). */
int gcta_b200_nonmiss_planes(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                             const int32_t* d_keep, int64_t n, const int32_t* d_include,
                             int64_t k0, int64_t kc, uint32_t* d_planes, int64_t ldm,
                             void* stream, char* err, size_t errcap);

/* K2 - lower-triangle SYRK  S += Z^T Z  on tcgen05 tensor cores (grm.cpp:276-296 numerator).
 * Tiles: 256 rows x 256 columns of S.  d_tiles = int32[ntiles][2] (row block, column block);
 * tile t of the list is stored at d_S + t*256*256 floats in the layout
 * [col/4][row][col%4] (see gcta_b200_s_tile_index).  S must be zeroed by the caller before
 * the first chunk. */
int gcta_b200_syrk_tiles(const void* d_Z, int64_t ldz, int64_t kc_pad, int32_t operand,
                         const int32_t* d_tiles, int64_t ntiles, float* d_S,
                         void* stream, char* err, size_t errcap);

/* K2 with different row and column operands:  S += Za^T Zb  on the same tile list (two matrices of one shape).
 * Used for the cross terms hi'lo + lo'hi of the split-FP16 operand mode (GCTA_B200_FP16X2), where
 * gcta_b200_bed_decode writes the residual plane kc_pad*ldz elements behind the first. */
int gcta_b200_syrk_tiles_ab(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int32_t operand,
                            const int32_t* d_tiles, int64_t ntiles, float* d_S,
                            void* stream, char* err, size_t errcap);

/* K2 with the contraction cut into TMEM chains of chain_snps SNPs (a multiple of 128): each chain is one FP32 accumulation
 * on the tensor cores, added into S by a round-to-nearest reduction - long chains of same-sign products (related
 * individuals) otherwise drift by the accumulator's truncation (grm.cpp:288-289 accumulates in double).  All chains run
 * in one launch.  What the plan calls with its default chain of 32,768 SNPs. */
int gcta_b200_syrk_tiles_chained(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int64_t chain_snps, int32_t operand,
                                 const int32_t* d_tiles, int64_t ntiles, float* d_S,
                                 void* stream, char* err, size_t errcap);

/* K2 as CTA pairs (tcgen05.mma.cta_group::2): two SMs work on two vertically adjacent 256x256 tiles and share the
 * column operand, each staging only half of it (csrc/syrk2.cu).  Same tile storage as gcta_b200_syrk_tiles.
 *   gcta_b200_pair_list  host helper: from a band's tile list (gcta_b200_tile_list order = storage order) to the work
 *                        list int32[count][4] = {column block, first row block, storage index of the upper tile or -1,
 *                        storage index of the lower tile or -1}; returns count (pairs == NULL: count only)
 *   gcta_b200_syrk_pairs d_S (+)= Za^T Zb over the pair list; int_acc != 0: counting mode (d_S is int32 tiles)
 * Selected for the whole path with GCTA_B200_2CTA=1. */
int64_t gcta_b200_pair_list(const int32_t* tiles, int64_t ntiles, int32_t raster_rows, int32_t* pairs);
int gcta_b200_syrk_pairs(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int32_t operand,
                         const int32_t* d_pairs, int64_t npairs, void* d_S, int32_t int_acc,
                         void* stream, char* err, size_t errcap);

/* K3 - bit-exact pair counts N_ij = sum_k ok_ik ok_jk by AND + popc over the K1m bit planes
 * (grm.cpp:291), accumulated into int32 tiles laid out like K2's S tiles. */
int gcta_b200_paircount_tiles(const uint32_t* d_planes, int64_t ldm, int64_t nwords,
                              const int32_t* d_tiles, int64_t ntiles, int32_t* d_Ntiles,
                              void* stream, char* err, size_t errcap);

/* K1n + K2 in counting mode - the default way N is computed: the non-missing indicator of every genotype as a
 * 16-bit 0/1 matrix laid out like Z (gcta_b200_nonmiss_matrix), contracted by the SYRK kernel with an int32
 * epilogue (gcta_b200_count_tiles).  0/1 products and integer partial sums below 2^24 are exact in the FP32
 * tensor-core accumulator, so the result is bit-identical to K3's AND+popc (tests assert this) at the tensor
 * rate.  d_Ntiles as for gcta_b200_paircount_tiles.  grm.cpp:270, 291. */
int gcta_b200_nonmiss_matrix(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_keep, int64_t n,
                             const int32_t* d_include, int64_t k0, int64_t kc, int32_t operand, void* d_OK, int64_t ldz,
                             void* stream, char* err, size_t errcap);
int gcta_b200_count_tiles(const void* d_OK, int64_t ldz, int64_t kc_pad, int32_t operand, const int32_t* d_tiles,
                          int64_t ntiles, int32_t* d_Ntiles, void* stream, char* err, size_t errcap);

/* K4 - A = S / N with the reference's zero-count rules (grm.cpp:273-275, 294), exact diagonal
 * from K1d, packed to .grm.gz order.  d_Ntiles NULL means N == m_const for every pair. */
int gcta_b200_finalize_tiles(const float* d_S, const int32_t* d_Ntiles, int32_t m_const,
                             const int32_t* d_tiles, int64_t ntiles, int64_t n,
                             const int32_t* d_keep, const double* d_diag, const int32_t* d_nii,
                             int64_t row_begin, int64_t row_end,
                             float* d_A_lower, int32_t* d_N_lower,
                             void* stream, char* err, size_t errcap);

/* ---- --grm-alg 1 (grm.cpp:230-237, 257-264, 278-285): A_ij = sum_k z_ik z_jk / sum_k ok_ik ok_jk d_k with z centred
 * only and d_k = mean over non-missing individuals of z_k^2.  The "_ex" variants add the extra operands; passing NULL
 * for them is the plain call above.
 *   bed_decode_ex    d_dk      out  double[>= k0+kc]: d_k per included SNP (call with alg = 1)
 *   diag_accum_ex    d_dk in, d_dw out  double per FILE individual: dw[f] += sum_k ok_fk d_k (diagonal denominator)
 *   nonmiss_matrix_ex d_dk in  the matrix holds ok * sqrt(d_k) instead of ok, so that the SYRK kernel in float mode
 *                              (gcta_b200_syrk_tiles) yields the off-diagonal denominators D_ij
 *   finalize_tiles_ex d_Dtiles (float tiles like S, or NULL when nothing is missing and D_ij == d_const), d_dw */
int gcta_b200_bed_decode_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                            const uint32_t* d_keepmask, const int32_t* d_keep, int64_t n,
                            const uint32_t* d_sexmale, const int32_t* d_include, int64_t k0, int64_t kc,
                            const double* d_mu_in, int32_t xchr, int32_t alg, int32_t operand,
                            void* d_Z, int64_t ldz,
                            double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut, double* d_dk,
                            unsigned long long* d_total_missing, void* stream, char* err, size_t errcap);
int gcta_b200_diag_accum_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file,
                            const uint32_t* d_sexmale, const int32_t* d_include, int64_t k0, int64_t kc,
                            const float* d_zlut, double* d_diag, int32_t* d_nii, const double* d_dk, double* d_dw,
                            void* stream, char* err, size_t errcap);
int gcta_b200_nonmiss_matrix_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_keep, int64_t n,
                                const int32_t* d_include, int64_t k0, int64_t kc, int32_t operand, const double* d_dk,
                                void* d_OK, int64_t ldz, void* stream, char* err, size_t errcap);
int gcta_b200_finalize_tiles_ex(const float* d_S, const int32_t* d_Ntiles, int32_t m_const,
                                const float* d_Dtiles, double d_const, const double* d_dw,
                                const int32_t* d_tiles, int64_t ntiles, int64_t n,
                                const int32_t* d_keep, const double* d_diag, const int32_t* d_nii,
                                int64_t row_begin, int64_t row_end,
                                float* d_A_lower, int32_t* d_N_lower,
                                void* stream, char* err, size_t errcap);

/* index of element (r, c), r < 256, c < 256, inside one S / N tile */
static inline int64_t gcta_b200_s_tile_index(int r, int c) { return ((int64_t)(c >> 2) * 256 + r) * 4 + (c & 3); }

/* helpers */
const char* gcta_b200_version(void);
int gcta_b200_device_count(void); /* 0 when no driver / device */
/* row band [row_begin,row_end) owned by part_rank of part_world for n kept individuals
 * (boundaries are multiples of 256, bands balanced by tile count) */
int gcta_b200_partition(int64_t n, int32_t part_rank, int32_t part_world, int64_t* row_begin, int64_t* row_end);
/* number of 256x256 tiles of band [row_begin,row_end) and, when tiles != NULL, the list in
 * launch order (int32[count][2]) */
int64_t gcta_b200_tile_list(int64_t n, int64_t row_begin, int64_t row_end, int32_t* tiles);

#ifdef __cplusplus
}
#endif
#endif /* GCTA_B200_H */
