// decode.cu - K1 / K1d / K1m: the row-per-CTA kernels that read the packed 2-bit genotypes.
// Since round 2 the plan uses them for the X chromosome, --grm-alg 1 and the split-FP16 operand
// mode only; everything else goes through K1s / K1f (/ K1c) in decode2.cu.
//
// K1  bed_decode      : one pass per SNP row - count alleles (popc over bit planes, shuffle
//                       reduction), derive 2p and the reference's doubly-rounded standardised
//                       values as a 4-entry look-up table, and write the 16-bit SNP-major Z
//                       row that K2 consumes.  Replaces data.cpp:171-185 (unpack), 781-803
//                       (mu_func), 1001-1007 (make_XMat) and 1030-1043 (std_XMat).
// K1d diag_accum      : exact float64 diagonal sum_k z_ik^2 and per-individual non-missing
//                       counts N_ii (grm.cpp:265-275), column-wise over the packed bytes.
// K1m nonmiss_planes  : bit-packed non-missing indicators, 32 SNPs per word, individual-minor,
//                       for the AND+popc pair-count kernel K3.
//
// All three are HBM-bound byte/bit kernels: 128-bit loads of packed genotypes, 128-bit
// coalesced stores of Z.  Algorithmic bytes per SNP: ceil(n/4) read + 2*ldz written (K1).
#include <algorithm>
#include <atomic>
#include <vector>

#include "common.cuh"

namespace gb {

// code (2 bits, low bit first): 00 -> 2 copies of allele 1, 10b -> 1, 11 -> 0, 01b -> missing
// (data.cpp:176-181: _snp_2 = !low, _snp_1 = !high, missing iff _snp_1 && !_snp_2).
struct SnpCounts {
    int cnt, sum, cntM, sumM, het, hetM;   // non-missing, allele-1 dosage sum, the same for males, heterozygotes
};

template <bool XCHR>
__device__ __forceinline__ void count_word(uint32_t w, uint32_t km, uint32_t mm, SnpCounts& c)
{
    const uint32_t lo = w & 0x55555555u, hi = (w >> 1) & 0x55555555u;
    const uint32_t nm = ~(lo & ~hi) & km;  // kept and not missing (even bit positions)
    c.cnt += __popc(nm);
    c.sum += __popc(nm & ~hi) + __popc(nm & ~lo);
    c.het += __popc(nm & hi & ~lo);   // code 10b
    if (XCHR) {
        const uint32_t nmm = nm & mm;
        c.cntM += __popc(nmm);
        c.sumM += __popc(nmm & ~hi) + __popc(nmm & ~lo);
        c.hetM += __popc(nmm & hi & ~lo);
    }
}

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ uint32_t to16(float z, int bf16)
{
    return bf16 ? (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(z)) : (uint32_t)__half_as_ushort(__float2half_rn(z));
}

struct DecodeParams {
    const uint8_t* bed;
    int64_t pitch;
    int64_t n_file;
    const uint32_t* keepmask;
    const int32_t* keep;
    int64_t n, n_out;
    const uint32_t* sexmale;
    const int32_t* include;
    int64_t k0, kc, kc_pad;
    const double* mu_in;
    int divide, bf16;
    uint16_t* Z;
    uint16_t* Z2;   // optional residual plane (split FP16 mode)
    int64_t ldz;
    double* mu;
    int32_t* cnt;
    int32_t* sum;
    float* zlut;
    double* dk;   // optional: per-SNP mean of z^2 (--grm-alg 1)
    unsigned long long* total_missing;
};

// GROUP threads cooperate on one SNP row; 256/GROUP rows per block.
template <int GROUP, bool XCHR>
__global__ void __launch_bounds__(256) bed_decode_kernel(const DecodeParams p)
{
    constexpr int ROWS = 256 / GROUP;
    const int tid = threadIdx.x % GROUP;
    const int64_t kk = (int64_t)blockIdx.x * ROWS + threadIdx.x / GROUP;
    if (kk >= p.kc_pad) return;  // uniform per group (and per block when GROUP == 256)
    uint4* zrow = reinterpret_cast<uint4*>(p.Z + kk * p.ldz);
    const int nchunk = (int)(p.ldz / 8);
    if (kk >= p.kc) {  // padding row of the last 64-SNP block
        for (int c = tid; c < nchunk; c += GROUP) zrow[c] = make_uint4(0, 0, 0, 0);
        if (p.Z2) {
            uint4* zrow2 = reinterpret_cast<uint4*>(p.Z2 + kk * p.ldz);
            for (int c = tid; c < nchunk; c += GROUP) zrow2[c] = make_uint4(0, 0, 0, 0);
        }
        return;
    }
    const int64_t k = p.k0 + kk;
    const uint8_t* row = p.bed + (int64_t)(p.include ? p.include[k] : (int32_t)k) * p.pitch;

    // ---- pass 1: allele counts ----
    SnpCounts c = {0, 0, 0, 0};
    const int nvec = (int)(p.pitch / 16);
    const uint4* row16 = reinterpret_cast<const uint4*>(row);
    const uint4* km16 = reinterpret_cast<const uint4*>(p.keepmask);
    const uint4* mm16 = reinterpret_cast<const uint4*>(p.sexmale);
    for (int v = tid; v < nvec; v += GROUP) {
        const uint4 w = __ldg(row16 + v), km = __ldg(km16 + v);
        uint4 mm = make_uint4(0, 0, 0, 0);
        if (XCHR) mm = __ldg(mm16 + v);
        count_word<XCHR>(w.x, km.x, mm.x, c);
        count_word<XCHR>(w.y, km.y, mm.y, c);
        count_word<XCHR>(w.z, km.z, mm.z, c);
        count_word<XCHR>(w.w, km.w, mm.w, c);
    }
    c.cnt = warp_sum(c.cnt);
    c.sum = warp_sum(c.sum);
    c.het = warp_sum(c.het);
    if (XCHR) {
        c.cntM = warp_sum(c.cntM);
        c.sumM = warp_sum(c.sumM);
        c.hetM = warp_sum(c.hetM);
    }
    if (GROUP > 32) {
        __shared__ int red[8][6];
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) {
            red[w][0] = c.cnt; red[w][1] = c.sum; red[w][2] = c.cntM; red[w][3] = c.sumM; red[w][4] = c.het; red[w][5] = c.hetM;
        }
        __syncthreads();
        c = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 8; i++) {
            c.cnt += red[i][0]; c.sum += red[i][1]; c.cntM += red[i][2]; c.sumM += red[i][3]; c.het += red[i][4]; c.hetM += red[i][5];
        }
    }

    // ---- per-SNP constants: mu = 2p and the doubly-rounded z values (data.cpp:801, 1027-1040) ----
    double mu;
    if (p.mu_in) mu = p.mu_in[k];
    else if (XCHR) {
        // weights 0.5 (male) / 1.0 (female), data.cpp:765-767: (0.5 sM + sF)/(0.5 cM + cF) == (sM + 2 sF)/(cM + 2 cF)
        const int num = c.sumM + 2 * (c.sum - c.sumM), den = c.cntM + 2 * (c.cnt - c.cntM);
        mu = den > 0 ? __ddiv_rn((double)num, (double)den) : 0.0;
    } else
        mu = c.cnt > 0 ? __ddiv_rn((double)c.sum, (double)c.cnt) : 0.0;
    const double var = __dmul_rn(mu, __dsub_rn(1.0, __dmul_rn(0.5, mu)));
    const double isd = fabs(var) < 1.0e-50 ? 0.0 : __dsqrt_rn(__ddiv_rn(1.0, var));
    float zf[4], zm[4];
#pragma unroll
    for (int code = 0; code < 4; code++) {
        const int x = code == 0 ? 2 : (code == 2 ? 1 : 0);
        float z = __double2float_rn(__dsub_rn((double)x, mu));                 // X -= mu   (data.cpp:1039)
        if (p.divide) z = __double2float_rn(__dmul_rn((double)z, isd));        // X *= sd   (data.cpp:1040)
        if (code == 1) z = 0.0f;                                               // missing contributes nothing
        zf[code] = z;
        zm[code] = XCHR ? __double2float_rn(__dmul_rn((double)z, __dsqrt_rn(0.5))) : z;  // data.cpp:1048-1054
    }
    // 16-bit operand tables.  Split mode (Z2 != NULL): a second plane holds the FP16 residual z - fp16(z), so that
    // hi*hi + hi*lo + lo*hi reproduces the FP32 products to ~2^-22 (the "tight" operand mode)
    uint64_t lutF_hi = 0, lutM_hi = 0, lutF_lo = 0, lutM_lo = 0;
#pragma unroll
    for (int code = 0; code < 4; code++) {
        const uint32_t hf = to16(zf[code], p.bf16), hm = to16(zm[code], p.bf16);
        lutF_hi |= (uint64_t)hf << (16 * code);
        lutM_hi |= (uint64_t)hm << (16 * code);
        if (p.Z2) {
            const float rf = zf[code] - __half2float(__ushort_as_half((unsigned short)hf));
            const float rm = zm[code] - __half2float(__ushort_as_half((unsigned short)hm));
            lutF_lo |= (uint64_t)to16(rf, 0) << (16 * code);
            lutM_lo |= (uint64_t)to16(rm, 0) << (16 * code);
        }
    }
    if (tid == 0) {
        p.mu[k] = mu;
        p.cnt[k] = c.cnt;
        p.sum[k] = c.sum;
#pragma unroll
        for (int code = 0; code < 4; code++) {
            p.zlut[k * 8 + code] = zf[code];
            p.zlut[k * 8 + 4 + code] = zm[code];
        }
        atomicAdd(p.total_missing, (unsigned long long)(p.n - c.cnt));
        if (p.dk) {
            // --grm-alg 1: d_k = mean over non-missing individuals of z^2 (grm.cpp:230-237), from the genotype
            // counts: n2 (code 00) = (sum - het) / 2, n1 = het, n0 = cnt - n1 - n2, split by sex for the X chromosome
            const int n1m = c.hetM, n2m = (c.sumM - c.hetM) / 2, n0m = c.cntM - n1m - n2m;
            const int n1f = c.het - n1m, n2f = (c.sum - c.het) / 2 - n2m, n0f = (c.cnt - c.cntM) - n1f - n2f;
            double ss = (double)n0f * ((double)zf[3] * zf[3]) + (double)n1f * ((double)zf[2] * zf[2]) + (double)n2f * ((double)zf[0] * zf[0]);
            ss += (double)n0m * ((double)zm[3] * zm[3]) + (double)n1m * ((double)zm[2] * zm[2]) + (double)n2m * ((double)zm[0] * zm[0]);
            p.dk[k] = c.cnt > 0 ? ss / (double)c.cnt : 0.0;
        }
    }

    // ---- pass 2: write the standardised row (8 individuals = one 16-byte store per lane) ----
    const uint16_t* row2 = reinterpret_cast<const uint16_t*>(row);
    const int nplanes = p.Z2 ? 2 : 1;
    for (int plane = 0; plane < nplanes; plane++) {
    const uint64_t lutF = plane ? lutF_lo : lutF_hi, lutM = plane ? lutM_lo : lutM_hi;
    if (plane) zrow = reinterpret_cast<uint4*>(p.Z2 + kk * p.ldz);
    if (GROUP == 256 && !XCHR && !p.keep) {
        // Fast path (one CTA per row, everyone kept, one LUT for all): a 256-entry table "packed byte -> four
        // 16-bit values" in shared memory turns the emit loop into 2 LDS.64 + 1 STG.128 per 8 genotypes; the
        // per-genotype shift/mask/select version below is issue-bound (ncu: SM 91 %, HBM 61 %).
        __shared__ uint2 tab[256];
        if (plane) __syncthreads();   // everyone is done reading the previous plane's table
        {
            const uint32_t b = threadIdx.x;
            uint32_t v[4];
#pragma unroll
            for (int e = 0; e < 4; e++) v[e] = (uint32_t)(lutF >> (16 * ((b >> (2 * e)) & 3u))) & 0xFFFFu;
            tab[b] = make_uint2(v[0] | (v[1] << 16), v[2] | (v[3] << 16));
        }
        __syncthreads();
        const int nfull = (int)(p.n_out / 8);   // chunks entirely below n_out
        for (int ch = tid; ch < nchunk; ch += GROUP) {
            uint4 o = make_uint4(0, 0, 0, 0);
            if (ch < nfull) {
                const uint32_t w = __ldg(row2 + ch);
                const uint2 a = tab[w & 0xFFu], b = tab[w >> 8];
                o = make_uint4(a.x, a.y, b.x, b.y);
            } else if ((int64_t)ch * 8 < p.n_out) {   // the one chunk that straddles n_out
                const uint32_t w = __ldg(row2 + ch);
                uint32_t h[4] = {0, 0, 0, 0};
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    uint32_t v = (uint32_t)(lutF >> (16 * ((w >> (2 * e)) & 3u))) & 0xFFFFu;
                    if ((int64_t)ch * 8 + e >= p.n_out) v = 0;
                    h[e >> 1] |= v << (16 * (e & 1));
                }
                o = make_uint4(h[0], h[1], h[2], h[3]);
            }
            zrow[ch] = o;
        }
        continue;
    }
    const uint16_t* male2 = reinterpret_cast<const uint16_t*>(p.sexmale);
    for (int ch = tid; ch < nchunk; ch += GROUP) {
        const int64_t i0 = (int64_t)ch * 8;
        uint32_t o[4] = {0, 0, 0, 0};
        if (i0 < p.n_out) {
            if (!p.keep) {
                const uint32_t w = __ldg(row2 + ch);
                // sexmale marks individual 16w+t at bit 2t, i.e. the same 2-bit slots as the genotypes
                const uint32_t mb = XCHR ? (uint32_t)__ldg(male2 + ch) : 0u;
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    const uint32_t code = (w >> (2 * e)) & 3u;
                    const uint64_t lut = (XCHR && ((mb >> (2 * e)) & 1u)) ? lutM : lutF;
                    uint32_t v = (uint32_t)(lut >> (16 * code)) & 0xFFFFu;
                    if (i0 + e >= p.n_out) v = 0;
                    o[e >> 1] |= v << (16 * (e & 1));
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    uint32_t v = 0;
                    if (i0 + e < p.n_out) {
                        const int f = p.keep[i0 + e];
                        const uint32_t code = (row[f >> 2] >> ((f & 3) * 2)) & 3u;
                        const bool male = XCHR && ((p.sexmale[f >> 4] >> ((f & 15) * 2)) & 1u);
                        v = (uint32_t)((male ? lutM : lutF) >> (16 * code)) & 0xFFFFu;
                    }
                    o[e >> 1] |= v << (16 * (e & 1));
                }
            }
        }
        zrow[ch] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    }   // plane
}

// ------------------------------------------------------------------------------------------------
// K1d: thread = one 32-bit column (16 file individuals) x a slice of SNPs.
// The squares z^2 of the (at most) three values a SNP takes are staged per tile of 192 SNPs as a
// [SNP][code] float table in shared memory; per genotype the inner loop is shift, mask, one LDS
// (code -> z^2; a warp touches at most 4 banks) and one FADD.  FP32 partial sums run over the
// block's 768-SNP slice (keeps the kernel at 64 registers, 4 blocks/SM) and are folded into FP64
// by the final atomics; missing genotypes are counted bit-sliced (2-bit lanes -> 8-bit lanes every
// 3 SNPs -> 32-bit per tile), so counting costs ~0.3 instructions per genotype.
// ------------------------------------------------------------------------------------------------
constexpr int DIAG_TILE = 192;              // SNPs per shared-memory tile (multiple of 3, <= 255)
constexpr int DIAG_SPLIT = 4 * DIAG_TILE;   // SNPs per block slice
constexpr int DIAG_THREADS = 256;

// ALG1 (--grm-alg 1): additionally dw[f] += sum over non-missing SNPs of d_k, the diagonal's denominator
// (grm.cpp:257-264), through a second [SNP][code] table {d_k, 0, d_k, d_k}.
template <bool XCHR, bool ALG1 = false>
__global__ void __launch_bounds__(DIAG_THREADS) diag_accum_kernel(const uint8_t* __restrict__ bed, int64_t pitch, int64_t n_file,
                                                                   const uint32_t* __restrict__ sexmale,
                                                                   const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                                   const float* __restrict__ zlut, double* __restrict__ diag,
                                                                   int32_t* __restrict__ nii, const double* __restrict__ dk = nullptr,
                                                                   double* __restrict__ dw = nullptr)
{
    __shared__ __align__(16) float q_s[((XCHR ? 2 : 1) + (ALG1 ? 1 : 0)) * DIAG_TILE * 4];   // [sex][SNP][code] squares (+ d_k table)
    constexpr uint32_t DK_OFF = (XCHR ? 2 : 1) * DIAG_TILE * 16;   // byte offset of the d_k table
    __shared__ int64_t row_s[DIAG_TILE];                                   // byte offset of each SNP row
    const int64_t col = (int64_t)blockIdx.x * DIAG_THREADS + threadIdx.x;
    const bool active = col * 16 < n_file;
    const int64_t kbeg = (int64_t)blockIdx.y * DIAG_SPLIT, kend = min(kbeg + (int64_t)DIAG_SPLIT, kc);
    // byte offsets into q_s of the male table, per individual slot (0 for females / autosomes)
    uint32_t moff[16];
#pragma unroll
    for (int e = 0; e < 16; e++) moff[e] = 0;
    if (XCHR && active) {
        const uint32_t mm = sexmale[col];
#pragma unroll
        for (int e = 0; e < 16; e++) moff[e] = ((mm >> (2 * e)) & 1u) ? DIAG_TILE * 16 : 0;
    }
    // FP32 partial sums over the whole 768-SNP slice (rounding ~1e-6 relative, unbiased), folded into FP64 by the atomics
    float acc[16], acc2[ALG1 ? 16 : 1];
    int miss[16];
#pragma unroll
    for (int e = 0; e < 16; e++) { acc[e] = 0.0f; miss[e] = 0; if (ALG1) acc2[e] = 0.0f; }
    const char* q_bytes = reinterpret_cast<const char*>(q_s);

    for (int64_t t0 = kbeg; t0 < kend; t0 += DIAG_TILE) {
        const int tn = (int)min((int64_t)DIAG_TILE, kend - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < tn; i += DIAG_THREADS) {
            const int64_t k = k0 + t0 + i;
            row_s[i] = (include ? (int64_t)include[k] : k) * pitch;
            const float4 lf = __ldg(reinterpret_cast<const float4*>(zlut + k * 8));
            // exact squares of floats rounded once to float (relative 6e-8, unbiased)
            reinterpret_cast<float4*>(q_s)[i] = make_float4(lf.x * lf.x, 0.0f, lf.z * lf.z, lf.w * lf.w);
            if (XCHR) {
                const float4 lm = __ldg(reinterpret_cast<const float4*>(zlut + k * 8 + 4));
                reinterpret_cast<float4*>(q_s)[DIAG_TILE + i] = make_float4(lm.x * lm.x, 0.0f, lm.z * lm.z, lm.w * lm.w);
            }
            if (ALG1) {
                const float d = (float)dk[k];
                reinterpret_cast<float4*>(q_s)[(XCHR ? 2 : 1) * DIAG_TILE + i] = make_float4(d, 0.0f, d, d);
            }
        }
        __syncthreads();
        if (!active) continue;
        uint32_t c8[4] = {0, 0, 0, 0};
        uint32_t c2 = 0;
        int in3 = 0;
#pragma unroll 4
        for (int i = 0; i < tn; i++) {
            const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(bed + row_s[i]) + col);
            const char* qrow = q_bytes + i * 16;
#pragma unroll
            for (int e = 0; e < 16; e++) {
                // byte offset of the code inside the 16-byte row: ((w >> 2e) & 3) * 4
                const uint32_t off = e == 0 ? (w << 2) & 0xCu : (w >> (2 * e - 2)) & 0xCu;
                acc[e] += *reinterpret_cast<const float*>(qrow + off + (XCHR ? moff[e] : 0u));
                if (ALG1) acc2[e] += *reinterpret_cast<const float*>(qrow + off + DK_OFF);
            }
            c2 += (w & 0x55555555u) & ~(w >> 1);   // missing = code 01b, one count per 2-bit lane
            if (++in3 == 3) {
                in3 = 0;
#pragma unroll
                for (int g = 0; g < 4; g++) c8[g] += (c2 >> (2 * g)) & 0x03030303u;
                c2 = 0;
            }
        }
#pragma unroll
        for (int g = 0; g < 4; g++) c8[g] += (c2 >> (2 * g)) & 0x03030303u;
#pragma unroll
        for (int e = 0; e < 16; e++) {
            miss[e] += (int)((c8[e & 3] >> (8 * (e >> 2))) & 0xFFu);   // byte b of c8[g] = individual 4b + g
        }
    }
    if (!active) return;
    const int span = (int)(kend - kbeg);
#pragma unroll
    for (int e = 0; e < 16; e++) {
        const int64_t f = col * 16 + e;
        if (f < n_file) {
            atomicAdd(diag + f, (double)acc[e]);
            atomicAdd(nii + f, span - miss[e]);
            if (ALG1) atomicAdd(dw + f, (double)acc2[e]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K1m: non-missing bit planes.  thread = (plane word w, 16 individuals); identity keep only.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) nonmiss_planes_kernel(const uint8_t* __restrict__ bed, int64_t pitch, int64_t n,
                                                              const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                              uint32_t* __restrict__ planes, int64_t ldm)
{
    const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // 16-individual column
    if (col * 16 >= ldm) return;
    const int64_t w = k0 / 32 + blockIdx.y;
    uint32_t out[16];
#pragma unroll
    for (int e = 0; e < 16; e++) out[e] = 0;
    if (col * 4 < pitch) {
        for (int b = 0; b < 32; b++) {
            const int64_t k = w * 32 + b;
            if (k >= k0 + kc) break;
            const int64_t frow = include ? include[k] : k;
            const uint32_t g = __ldg(reinterpret_cast<const uint32_t*>(bed + frow * pitch) + col);
            const uint32_t nm = ~((g & 0x55555555u) & ~((g >> 1) & 0x55555555u)) & 0x55555555u;
#pragma unroll
            for (int e = 0; e < 16; e++) out[e] |= ((nm >> (2 * e)) & 1u) << b;
        }
    }
    uint4* dst = reinterpret_cast<uint4*>(planes + w * ldm + col * 16);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; e++) o[e] = (col * 16 + q * 4 + e < n) ? out[q * 4 + e] : 0u;
        dst[q] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// gather variant for a non-identity keep list: thread = (plane word, kept individual)
__global__ void __launch_bounds__(128) nonmiss_planes_gather_kernel(const uint8_t* __restrict__ bed, int64_t pitch,
                                                                     const int32_t* __restrict__ keep, int64_t n,
                                                                     const int32_t* __restrict__ include, int64_t k0,
                                                                     int64_t kc, uint32_t* __restrict__ planes, int64_t ldm)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ldm) return;
    const int64_t w = k0 / 32 + blockIdx.y;
    uint32_t out = 0;
    if (i < n) {
        const int f = keep[i];
        for (int b = 0; b < 32; b++) {
            const int64_t k = w * 32 + b;
            if (k >= k0 + kc) break;
            const int64_t frow = include ? include[k] : k;
            const uint32_t code = (bed[frow * pitch + (f >> 2)] >> ((f & 3) * 2)) & 3u;
            out |= (uint32_t)(code != 1u) << b;
        }
    }
    planes[w * ldm + i] = out;
}

}  // namespace gb

extern "C" int gcta_b200_bed_decode(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_keepmask,
                                    const int32_t* d_keep, int64_t n, const uint32_t* d_sexmale, const int32_t* d_include,
                                    int64_t k0, int64_t kc, const double* d_mu_in, int32_t xchr, int32_t alg, int32_t operand,
                                    void* d_Z, int64_t ldz, double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut,
                                    unsigned long long* d_total_missing, void* stream, char* err, size_t errcap)
{
    return gcta_b200_bed_decode_ex(d_bed, pitch, n_file, d_keepmask, d_keep, n, d_sexmale, d_include, k0, kc, d_mu_in, xchr, alg, operand,
                                   d_Z, ldz, d_mu, d_cnt, d_sum, d_zlut, nullptr, d_total_missing, stream, err, errcap);
}

extern "C" int gcta_b200_bed_decode_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_keepmask,
                                       const int32_t* d_keep, int64_t n, const uint32_t* d_sexmale, const int32_t* d_include,
                                       int64_t k0, int64_t kc, const double* d_mu_in, int32_t xchr, int32_t alg, int32_t operand,
                                       void* d_Z, int64_t ldz, double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut,
                                       double* d_dk, unsigned long long* d_total_missing, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_keepmask && d_Z && d_mu && d_cnt && d_sum && d_zlut && d_total_missing, "bed_decode: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0 && ((uintptr_t)d_bed & 15) == 0, "bed_decode: bed pitch and pointer must be multiples of 16 bytes");
    GB_ARG(pitch * 4 >= n_file && n_file > 0, "bed_decode: pitch too small for n_file");
    GB_ARG(n > 0 && n <= n_file && (d_keep || n == n_file), "bed_decode: n must equal n_file when keep is NULL");
    GB_ARG(ldz > 0 && ldz % 256 == 0, "bed_decode: ldz must be a positive multiple of 256");
    GB_ARG(((uintptr_t)d_Z & 15) == 0, "bed_decode: Z must be 16-byte aligned");
    GB_ARG(kc > 0 && k0 >= 0, "bed_decode: empty SNP range");
    GB_ARG(!xchr || d_sexmale, "bed_decode: xchr needs the male mask");
    GB_ARG(operand == GCTA_B200_FP16 || operand == GCTA_B200_BF16 || operand == GCTA_B200_FP16X2, "bed_decode: unknown operand type");
    DecodeParams p;
    p.bed = d_bed; p.pitch = pitch; p.n_file = n_file; p.keepmask = d_keepmask; p.keep = d_keep; p.n = n; p.n_out = n < ldz ? n : ldz;  // a Z narrower than n keeps only the first ldz individuals (row bands need no more)
    p.sexmale = d_sexmale; p.include = d_include; p.k0 = k0; p.kc = kc; p.kc_pad = (kc + 63) / 64 * 64;
    p.mu_in = d_mu_in; p.divide = alg == 0; p.bf16 = operand == GCTA_B200_BF16;
    p.Z = (uint16_t*)d_Z;
    p.Z2 = operand == GCTA_B200_FP16X2 ? (uint16_t*)d_Z + p.kc_pad * ldz : nullptr;   // split mode: the residual plane follows the first
    p.ldz = ldz; p.mu = d_mu; p.cnt = d_cnt; p.sum = d_sum; p.zlut = d_zlut; p.dk = d_dk;
    p.total_missing = d_total_missing;
    cudaStream_t st = (cudaStream_t)stream;
    const bool wide = pitch >= 4096;  // one CTA per row when a row is at least 4 KB, else one warp per row
    if (wide) {
        const unsigned grid = (unsigned)p.kc_pad;
        if (xchr) bed_decode_kernel<256, true><<<grid, 256, 0, st>>>(p);
        else bed_decode_kernel<256, false><<<grid, 256, 0, st>>>(p);
    } else {
        const unsigned grid = (unsigned)((p.kc_pad + 7) / 8);
        if (xchr) bed_decode_kernel<32, true><<<grid, 256, 0, st>>>(p);
        else bed_decode_kernel<32, false><<<grid, 256, 0, st>>>(p);
    }
    GB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int gcta_b200_diag_accum(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_sexmale,
                                    const int32_t* d_include, int64_t k0, int64_t kc, const float* d_zlut, double* d_diag,
                                    int32_t* d_nii, void* stream, char* err, size_t errcap)
{
    return gcta_b200_diag_accum_ex(d_bed, pitch, n_file, d_sexmale, d_include, k0, kc, d_zlut, d_diag, d_nii, nullptr, nullptr, stream, err, errcap);
}

extern "C" int gcta_b200_diag_accum_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_sexmale,
                                       const int32_t* d_include, int64_t k0, int64_t kc, const float* d_zlut, double* d_diag,
                                       int32_t* d_nii, const double* d_dk, double* d_dw, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_zlut && d_diag && d_nii, "diag_accum: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0, "diag_accum: pitch must be a multiple of 16");
    GB_ARG(kc > 0, "diag_accum: empty SNP range");
    const int64_t cols = (n_file + 15) / 16;
    dim3 grid((unsigned)((cols + DIAG_THREADS - 1) / DIAG_THREADS), (unsigned)((kc + DIAG_SPLIT - 1) / DIAG_SPLIT));
    GB_ARG((d_dk == nullptr) == (d_dw == nullptr), "diag_accum: d_dk and d_dw go together");
    cudaStream_t st = (cudaStream_t)stream;
    if (d_dk) {
        if (d_sexmale) diag_accum_kernel<true, true><<<grid, DIAG_THREADS, 0, st>>>(d_bed, pitch, n_file, d_sexmale, d_include, k0, kc, d_zlut, d_diag, d_nii, d_dk, d_dw);
        else diag_accum_kernel<false, true><<<grid, DIAG_THREADS, 0, st>>>(d_bed, pitch, n_file, nullptr, d_include, k0, kc, d_zlut, d_diag, d_nii, d_dk, d_dw);
    } else if (d_sexmale) diag_accum_kernel<true><<<grid, DIAG_THREADS, 0, st>>>(d_bed, pitch, n_file, d_sexmale, d_include, k0, kc, d_zlut, d_diag, d_nii);
    else diag_accum_kernel<false><<<grid, DIAG_THREADS, 0, st>>>(d_bed, pitch, n_file, nullptr, d_include, k0, kc, d_zlut, d_diag, d_nii);
    GB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int gcta_b200_nonmiss_planes(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_keep, int64_t n,
                                        const int32_t* d_include, int64_t k0, int64_t kc, uint32_t* d_planes, int64_t ldm,
                                        void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_planes, "nonmiss_planes: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0, "nonmiss_planes: pitch must be a multiple of 16");
    GB_ARG(k0 % 32 == 0 && kc > 0, "nonmiss_planes: k0 must be a multiple of 32");
    GB_ARG(ldm > 0 && ldm % 256 == 0, "nonmiss_planes: ldm must be a positive multiple of 256");
    GB_ARG(d_keep || n == n_file, "nonmiss_planes: n must equal n_file when keep is NULL");
    if (n > ldm) n = ldm;   // planes narrower than n keep only the first ldm individuals
    const unsigned words = (unsigned)((kc + 31) / 32);
    if (!d_keep) {
        dim3 grid((unsigned)((ldm / 16 + 127) / 128), words);
        nonmiss_planes_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(d_bed, pitch, n, d_include, k0, kc, d_planes, ldm);
    } else {
        dim3 grid((unsigned)((ldm + 127) / 128), words);
        nonmiss_planes_gather_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(d_bed, pitch, d_keep, n, d_include, k0, kc, d_planes, ldm);
    }
    GB_CUDA(cudaGetLastError());
    return 0;
}

// ================================================================================================
// Allocation-free variant of K1d (no counterpart in the reference; see include/gcta_b200.h).
//
// gcta_b200_diag_accum needs a per-SNP table of standardised values (8 floats per SNP) that
// some earlier step has to place in device memory; in the plan that table is cudaMalloc'ed.
// Some hosts cannot do that: an engine that reserved all of HBM for its own pool at start-up,
// or a service loop that must never enter the CUDA allocator because cudaMalloc/cudaFree
// synchronise the whole device.  For them the table lives in the fixed array below, which is
// part of the library image: every GPU that loads libgcta_b200.so carries these 16 MiB until
// the process exits, whether or not the entry point is ever called.
// ================================================================================================
namespace gb {

constexpr int64_t STATIC_LUT_MAX_SNPS = 524288;  // per call; 524288 SNPs x 8 floats = 16 MiB
static __device__ __align__(256) float g_static_zlut[STATIC_LUT_MAX_SNPS * 8];

// one warp per SNP: allele counts of the kept individuals -> 2p -> the 4 standardised values,
// written to lut[(k - k0) * 8 ..] in the layout diag_accum_kernel reads (autosomes only, so
// the "male" half repeats the first).  Same arithmetic as bed_decode_kernel's per-SNP block.
__global__ void __launch_bounds__(256) snp_lut_kernel(const uint8_t* __restrict__ bed, int64_t pitch,
                                                       const uint32_t* __restrict__ keepmask,
                                                       const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                       float* __restrict__ lut)
{
    const int lane = threadIdx.x & 31;
    const int64_t kk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (kk >= kc) return;
    const int64_t k = k0 + kk;
    const uint4* row16 = reinterpret_cast<const uint4*>(bed + (int64_t)(include ? include[k] : (int32_t)k) * pitch);
    const uint4* km16 = reinterpret_cast<const uint4*>(keepmask);
    SnpCounts c = {0, 0, 0, 0};
    for (int v = lane; v < (int)(pitch / 16); v += 32) {
        const uint4 w = __ldg(row16 + v), km = __ldg(km16 + v);
        count_word<false>(w.x, km.x, 0u, c);
        count_word<false>(w.y, km.y, 0u, c);
        count_word<false>(w.z, km.z, 0u, c);
        count_word<false>(w.w, km.w, 0u, c);
    }
    c.cnt = warp_sum(c.cnt);
    c.sum = warp_sum(c.sum);
    if (lane != 0) return;
    const double mu = c.cnt > 0 ? __ddiv_rn((double)c.sum, (double)c.cnt) : 0.0;
    const double var = __dmul_rn(mu, __dsub_rn(1.0, __dmul_rn(0.5, mu)));
    const double isd = fabs(var) < 1.0e-50 ? 0.0 : __dsqrt_rn(__ddiv_rn(1.0, var));
#pragma unroll
    for (int code = 0; code < 4; code++) {
        const int x = code == 0 ? 2 : (code == 2 ? 1 : 0);
        float z = __double2float_rn(__dsub_rn((double)x, mu));
        z = __double2float_rn(__dmul_rn((double)z, isd));
        if (code == 1) z = 0.0f;
        lut[kk * 8 + code] = z;
        lut[kk * 8 + 4 + code] = z;
    }
}

// Device address of g_static_zlut for the CURRENT device.  In host code the array's name is
// only a handle the runtime understands, not something a kernel may dereference, so the real
// address is asked for once per device (with that device current) and remembered.
static int static_lut_address(float** out, char* err, size_t errcap)
{
    // atomics: two host threads may make their first call at the same time; both then look the same address up
    static std::atomic<float*> cached[64] = {};
    int dev = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_ARG(dev >= 0 && dev < 64, "diag_accum_static: device ordinal above 63");
    float* p = cached[dev].load(std::memory_order_acquire);
    if (!p) {
        GB_CUDA(cudaGetSymbolAddress((void**)&p, g_static_zlut));
        cached[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

}  // namespace gb

extern "C" int gcta_b200_diag_accum_static(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_keepmask,
                                           const int32_t* d_include, int64_t k0, int64_t kc, double* d_diag, int32_t* d_nii,
                                           void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_keepmask && d_diag && d_nii, "diag_accum_static: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0 && ((uintptr_t)d_bed & 15) == 0, "diag_accum_static: bed pitch and pointer must be multiples of 16 bytes");
    GB_ARG(pitch * 4 >= n_file && n_file > 0, "diag_accum_static: pitch too small for n_file");
    GB_ARG(kc > 0 && k0 >= 0, "diag_accum_static: empty SNP range");
    if (kc > STATIC_LUT_MAX_SNPS)
        return set_err(err, errcap, GCTA_B200_ERR_ARG,
                       "Error: diag_accum_static: " + std::to_string(kc) + " SNPs in one call exceed the built-in table of " +
                           std::to_string(STATIC_LUT_MAX_SNPS) + " SNPs; split the range over several calls.");
    float* lut = nullptr;
    int rc = static_lut_address(&lut, err, errcap);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // every table entry of [0, kc) is rewritten here before the second kernel reads it, so nothing
    // left behind by an earlier call can leak into this one
    snp_lut_kernel<<<(unsigned)((kc + 7) / 8), 256, 0, st>>>(d_bed, pitch, d_keepmask, d_include, k0, kc, lut);
    GB_CUDA(cudaGetLastError());
    const int64_t cols = (n_file + 15) / 16;
    dim3 grid((unsigned)((cols + DIAG_THREADS - 1) / DIAG_THREADS), (unsigned)((kc + DIAG_SPLIT - 1) / DIAG_SPLIT));
    // diag_accum_kernel indexes its table by absolute SNP number; the table starts at k0
    diag_accum_kernel<false><<<grid, DIAG_THREADS, 0, st>>>(d_bed, pitch, n_file, nullptr, d_include, k0, kc, lut - k0 * 8, d_diag, d_nii);
    GB_CUDA(cudaGetLastError());
    return 0;
}

// ================================================================================================
// Allele counts only (no Z): the counting half of calcu_mu / mu_func (data.cpp:759-803) for the
// host's --freq, --maf and --max-maf, which run before / without make_grm (option.cpp:601-615).
// One warp per included SNP; optional male / female masks give the X-chromosome split.
// ================================================================================================
namespace gb {

__global__ void __launch_bounds__(256) allele_counts_kernel(const uint8_t* __restrict__ bed, int64_t pitch, int64_t row_base,
                                                             const uint32_t* __restrict__ keepmask,
                                                             const uint32_t* __restrict__ malemask,
                                                             const uint32_t* __restrict__ femalemask,
                                                             const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                             int32_t* __restrict__ out /* [6][kc]: cnt,sum,cntM,sumM,cntF,sumF */)
{
    const int lane = threadIdx.x & 31;
    const int64_t kk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (kk >= kc) return;
    const int64_t k = k0 + kk;
    const int64_t frow = (include ? (int64_t)include[k] : k) - row_base;
    const uint4* row16 = reinterpret_cast<const uint4*>(bed + frow * pitch);
    const uint4* km16 = reinterpret_cast<const uint4*>(keepmask);
    const uint4* mm16 = reinterpret_cast<const uint4*>(malemask);
    const uint4* fm16 = reinterpret_cast<const uint4*>(femalemask);
    SnpCounts a = {0, 0, 0, 0}, f = {0, 0, 0, 0};
    for (int v = lane; v < (int)(pitch / 16); v += 32) {
        const uint4 w = __ldg(row16 + v), km = __ldg(km16 + v);
        if (malemask) {
            const uint4 mm = __ldg(mm16 + v), fm = __ldg(fm16 + v);
            count_word<true>(w.x, km.x, mm.x, a); count_word<true>(w.y, km.y, mm.y, a);
            count_word<true>(w.z, km.z, mm.z, a); count_word<true>(w.w, km.w, mm.w, a);
            count_word<true>(w.x, km.x, fm.x, f); count_word<true>(w.y, km.y, fm.y, f);
            count_word<true>(w.z, km.z, fm.z, f); count_word<true>(w.w, km.w, fm.w, f);
        } else {
            count_word<false>(w.x, km.x, 0u, a); count_word<false>(w.y, km.y, 0u, a);
            count_word<false>(w.z, km.z, 0u, a); count_word<false>(w.w, km.w, 0u, a);
        }
    }
    const int v0 = warp_sum(a.cnt), v1 = warp_sum(a.sum), v2 = warp_sum(a.cntM), v3 = warp_sum(a.sumM);
    const int v4 = warp_sum(f.cntM), v5 = warp_sum(f.sumM);
    if (lane == 0) {
        out[0 * kc + kk] = v0; out[1 * kc + kk] = v1; out[2 * kc + kk] = v2;
        out[3 * kc + kk] = v3; out[4 * kc + kk] = v4; out[5 * kc + kk] = v5;
    }
}

}  // namespace gb

extern "C" int gcta_b200_allele_counts(const gcta_b200_args* a, int32_t* cnt, int32_t* sum, int32_t* cntM, int32_t* sumM,
                                       int32_t* cntF, int32_t* sumF, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(a && a->bed && cnt && sum, "allele_counts: null argument");
    GB_ARG(a->n_file > 0 && a->m_file > 0, "allele_counts: empty genotype file");
    const int64_t n = a->keep ? a->n : a->n_file, m = a->include ? a->m : a->m_file;
    if (m <= 0) return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: No SNP is retained for analysis.");
    if (n <= 0) return set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: No individual is retained for analysis.");
    if (gcta_b200_device_count() <= 0)
        return set_err(err, errcap, GCTA_B200_ERR_NO_DEVICE, "Error: no CUDA device is available (gcta_b200 has no CPU fallback).");
    const bool split = a->sex != nullptr && cntM && sumM && cntF && sumF;
    GB_CUDA(cudaSetDevice(a->device));
    const int64_t stride = (a->n_file + 3) / 4, pitch = (stride + 15) / 16 * 16;
    std::string km((size_t)pitch, '\0'), mm, fm;
    if (split) { mm.assign((size_t)pitch, '\0'); fm.assign((size_t)pitch, '\0'); }
    auto setbit = [](std::string& s, int64_t f) { s[(size_t)(f >> 2)] |= (char)(1 << ((f & 3) * 2)); };
    for (int64_t i = 0; i < n; i++) {
        const int64_t f = a->keep ? a->keep[i] : i;
        GB_ARG(f >= 0 && f < a->n_file, "allele_counts: keep index out of range");
        setbit(km, f);
        if (split && a->sex[f] == 1) setbit(mm, f);
        if (split && a->sex[f] == 2) setbit(fm, f);
    }
    // slabs of included SNPs whose file rows span at most ~1 GiB
    const int64_t slab_rows = std::max<int64_t>(1, (1ll << 30) / pitch);
    uint8_t* d_bed = nullptr; uint32_t *d_km = nullptr, *d_mm = nullptr, *d_fm = nullptr; int32_t *d_inc = nullptr, *d_out = nullptr;
    int rc = 0;
    cudaError_t e = cudaSuccess;
    auto fail = [&](cudaError_t ce, const char* what) {
        rc = set_err(err, errcap, ce == cudaErrorMemoryAllocation ? GCTA_B200_ERR_NOMEM : GCTA_B200_ERR_CUDA,
                     std::string("Error: CUDA call failed: ") + what + ": " + cudaGetErrorString(ce));
    };
#define AC_TRY(x) do { if (!rc && (e = (x)) != cudaSuccess) fail(e, #x); } while (0)
    AC_TRY(cudaMalloc(&d_bed, (size_t)pitch * std::min<int64_t>(slab_rows, a->m_file)));
    AC_TRY(cudaMalloc(&d_km, (size_t)pitch));
    AC_TRY(cudaMemcpy(d_km, km.data(), (size_t)pitch, cudaMemcpyHostToDevice));
    if (split) {
        AC_TRY(cudaMalloc(&d_mm, (size_t)pitch)); AC_TRY(cudaMalloc(&d_fm, (size_t)pitch));
        AC_TRY(cudaMemcpy(d_mm, mm.data(), (size_t)pitch, cudaMemcpyHostToDevice));
        AC_TRY(cudaMemcpy(d_fm, fm.data(), (size_t)pitch, cudaMemcpyHostToDevice));
    }
    if (a->include) {
        AC_TRY(cudaMalloc(&d_inc, (size_t)m * 4));
        AC_TRY(cudaMemcpy(d_inc, a->include, (size_t)m * 4, cudaMemcpyHostToDevice));
    }
    AC_TRY(cudaMalloc(&d_out, (size_t)6 * std::min<int64_t>(slab_rows, m) * 4));
    std::vector<int32_t> h;
    for (int64_t k0 = 0; !rc && k0 < m;) {
        const int64_t r0 = a->include ? a->include[k0] : k0;
        int64_t k1 = k0;
        while (k1 < m && k1 - k0 < slab_rows && (a->include ? a->include[k1] : k1) - r0 < slab_rows &&
               (a->include ? a->include[k1] : k1) >= r0)
            k1++;
        if (k1 == k0) { rc = set_err(err, errcap, GCTA_B200_ERR_ARG, "Error: allele_counts: include must be ascending"); break; }
        const int64_t r1 = (a->include ? a->include[k1 - 1] : k1 - 1) + 1, kc = k1 - k0;
        if (pitch != stride) AC_TRY(cudaMemset(d_bed, 0, (size_t)pitch * (r1 - r0)));
        AC_TRY(cudaMemcpy2D(d_bed, (size_t)pitch, a->bed + r0 * stride, (size_t)stride, (size_t)stride, (size_t)(r1 - r0), cudaMemcpyHostToDevice));
        if (!rc) {
            allele_counts_kernel<<<(unsigned)((kc + 7) / 8), 256>>>(d_bed, pitch, r0, d_km, split ? d_mm : nullptr, split ? d_fm : nullptr, d_inc, k0, kc, d_out);
            AC_TRY(cudaGetLastError());
        }
        h.resize((size_t)6 * kc);
        AC_TRY(cudaMemcpy(h.data(), d_out, (size_t)6 * kc * 4, cudaMemcpyDeviceToHost));
        if (!rc) {
            memcpy(cnt + k0, h.data(), (size_t)kc * 4);
            memcpy(sum + k0, h.data() + kc, (size_t)kc * 4);
            if (split) {
                memcpy(cntM + k0, h.data() + 2 * kc, (size_t)kc * 4); memcpy(sumM + k0, h.data() + 3 * kc, (size_t)kc * 4);
                memcpy(cntF + k0, h.data() + 4 * kc, (size_t)kc * 4); memcpy(sumF + k0, h.data() + 5 * kc, (size_t)kc * 4);
            }
        }
        k0 = k1;
    }
#undef AC_TRY
    cudaFree(d_bed); cudaFree(d_km); cudaFree(d_mm); cudaFree(d_fm); cudaFree(d_inc); cudaFree(d_out);
    return rc;
}
