// decode2.cu - K1s + K1f: the standardisation stage as two streaming passes over the packed genotypes.
//
// K1s snp_stats        : one warp per included SNP row - allele counts of the kept individuals
//                        (popc over bit planes, shuffle reduction), 2p, and the reference's
//                        doubly-rounded standardised values as 4 floats per SNP.  Reads the packed
//                        row once, writes ~50 bytes.  Replaces calcu_mu / mu_func
//                        (data.cpp:759-803) and the per-SNP constants of std_XMat (data.cpp:1027-1040).
// K1f decode_emit_diag : CTA = a slab of 2048 individuals x a run of SNP rows.  A thread owns one
//                        32-bit packed column (16 individuals) and walks down the rows: per row one
//                        4-byte load, eight nibble look-ups (two genotypes each) in a per-SNP table
//                        in shared memory that yield the two 16-bit z values and their two float
//                        squares, one 32-byte store of the Z row segment, and sixteen float adds into
//                        the thread's own diagonal sums.  Missing genotypes are counted bit-sliced.
//                        Replaces make_XMat / std_XMat (data.cpp:982-1056) and the diagonal of the
//                        triple loop (grm.cpp:265-275) in ONE pass: the separate diagonal kernel and
//                        its re-read of the packed rows are gone.
//
// Both are HBM-bound byte kernels.  Algorithmic bytes per SNP: ceil(n/4) (K1s) and
// ceil(n/4) + 2*ldz (K1f).  This is the path for the common case (autosomes, --grm-alg 0, one
// 16-bit plane; with a keep list after K1c below has compacted the chunk's rows to the kept
// individuals); the X chromosome, --grm-alg 1 and the split-FP16 mode stay on the row-per-CTA
// kernels of decode.cu.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace gb {

constexpr int F_THREADS = 128;            // threads per CTA = packed 32-bit columns per slab
constexpr int F_SLAB = F_THREADS * 16;    // individuals per slab
constexpr int F_TB = 48;                  // SNP rows per table block (multiple of 6)
constexpr int F_FLUSH_BLOCKS = 16;        // table blocks between flushes of the float partial sums (768 rows: the rounding of a
                                          // float running sum grows with its length; measured 5e-7 relative at 3072 rows)

__device__ __forceinline__ int warp_sum2(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------
// K1s
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 4) snp_stats_kernel(const uint8_t* __restrict__ bed, int64_t pitch, int64_t n_kept,
                                                         const uint32_t* __restrict__ keepmask,
                                                         const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                         const double* __restrict__ mu_in, int divide, double* __restrict__ mu_out,
                                                         int32_t* __restrict__ cnt_out, int32_t* __restrict__ sum_out,
                                                         float* __restrict__ zlut, unsigned long long* __restrict__ total_missing)
{
    __shared__ int miss_s[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int missing = 0;
    // grid-stride over the rows: the grid is one wave of resident CTAs, so no partly filled last wave
    for (int64_t kk = (int64_t)blockIdx.x * 8 + warp; kk < kc; kk += (int64_t)gridDim.x * 8) {
        const int64_t k = k0 + kk;
        const uint4* row16 = reinterpret_cast<const uint4*>(bed + (int64_t)(include ? include[k] : (int32_t)k) * pitch);
        const uint4* km16 = reinterpret_cast<const uint4*>(keepmask);
        int cnt = 0, sum = 0;
        const int nvec = (int)(pitch / 16);
        // 16 bytes = 64 genotypes per lane and step.  All masks live on the even bit positions, so two of them share one
        // POPC (a | b << 1): 6 POPCs per 16 bytes instead of 12 - the quarter-rate POPC unit would otherwise bound this
        // kernel at about the HBM rate.  Four row vectors are fetched before any is counted (memory-level parallelism:
        // one load in flight per warp reaches ~4.3 TB/s, measured).
        auto count16 = [&](const uint4& w, const uint4& km) {
            const uint32_t ww[4] = {w.x, w.y, w.z, w.w}, kw[4] = {km.x, km.y, km.z, km.w};
            uint32_t nm[4], x[4], y[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t lo = ww[q] & 0x55555555u, hi = (ww[q] >> 1) & 0x55555555u;
                nm[q] = ~(lo & ~hi) & kw[q];   // kept and not missing
                x[q] = nm[q] & ~hi;            // allele-1 dosage = [not hi] + [not lo] among those
                y[q] = nm[q] & ~lo;
            }
            cnt += __popc(nm[0] | (nm[1] << 1)) + __popc(nm[2] | (nm[3] << 1));
            sum += __popc(x[0] | (y[0] << 1)) + __popc(x[1] | (y[1] << 1)) + __popc(x[2] | (y[2] << 1)) + __popc(x[3] | (y[3] << 1));
        };
        int v = lane;
        for (; v + 96 < nvec; v += 128) {
            uint4 w[4], km[4];
#pragma unroll
            for (int u = 0; u < 4; u++) { w[u] = __ldg(row16 + v + 32 * u); km[u] = __ldg(km16 + v + 32 * u); }
#pragma unroll
            for (int u = 0; u < 4; u++) count16(w[u], km[u]);
        }
        for (; v < nvec; v += 32) count16(__ldg(row16 + v), __ldg(km16 + v));
        cnt = warp_sum2(cnt);
        sum = warp_sum2(sum);
        if (lane == 0) {
            // mu = 2p and the doubly-rounded z values (data.cpp:801, 1027-1040), same operation order as decode.cu
            const double mu = mu_in ? mu_in[k] : (cnt > 0 ? __ddiv_rn((double)sum, (double)cnt) : 0.0);
            const double var = __dmul_rn(mu, __dsub_rn(1.0, __dmul_rn(0.5, mu)));
            const double isd = fabs(var) < 1.0e-50 ? 0.0 : __dsqrt_rn(__ddiv_rn(1.0, var));
            float z[4];
#pragma unroll
            for (int code = 0; code < 4; code++) {
                const int x = code == 0 ? 2 : (code == 2 ? 1 : 0);
                float v = __double2float_rn(__dsub_rn((double)x, mu));              // X -= mu   (data.cpp:1039)
                if (divide) v = __double2float_rn(__dmul_rn((double)v, isd));       // X *= sd   (data.cpp:1040)
                z[code] = code == 1 ? 0.0f : v;                                     // missing contributes nothing
            }
            mu_out[k] = mu;
            cnt_out[k] = cnt;
            sum_out[k] = sum;
            reinterpret_cast<float4*>(zlut + k * 8)[0] = make_float4(z[0], z[1], z[2], z[3]);
            reinterpret_cast<float4*>(zlut + k * 8)[1] = make_float4(z[0], z[1], z[2], z[3]);   // "male" half: same on autosomes
            missing += (int)(n_kept - cnt);
        }
    }
    if (lane == 0) miss_s[warp] = missing;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) t += (unsigned long long)miss_s[i];
        if (t) atomicAdd(total_missing, t);
    }
}

// ------------------------------------------------------------------------------------------------
// K1f
// ------------------------------------------------------------------------------------------------
struct FusedParams {
    const uint8_t* bed;
    int64_t pitch, n_file;
    const int32_t* include;
    int64_t k0, kc, kc_pad;
    const float* zlut;
    int bf16;
    uint16_t* Z;
    int64_t ldz, n_out;
    double* diag;
    int32_t* nii;
    int64_t diag_f0, diag_f1;   // file individuals whose diagonal is wanted
    uint8_t* OK;                // optional: 8-bit E4M3 non-missing indicators [ok_pad][ldz] for the counting pass
    int64_t ok_pad;             // kc rounded up to 128 (the 8-bit K block of the SYRK kernel)
    int64_t rows_per_cta;
};

__device__ __forceinline__ uint32_t f2h16(float z, int bf16)
{
    return bf16 ? (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(z)) : (uint32_t)__half_as_ushort(__float2half_rn(z));
}

// no "memory" clobber on the stores: nothing in the kernel reads Z or OK, and the table look-ups of the next row may move above them
__device__ __forceinline__ void st_z256(uint16_t* dst, const uint32_t (&o)[8])
{
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]),
                 "r"(o[5]), "r"(o[6]), "r"(o[7]));
}

// One packed word = 16 genotypes of one SNP.  `row_addr` is the shared-memory address of the SNP's 256-byte table row
// (256-byte aligned): 16 nibble entries of 8 bytes - the two 16-bit z values of the nibble's two individuals, then their
// two E4M3 non-missing indicators - followed from byte 128 by 16 float2 entries with the squares of the z values.
// The eight nibble offsets come out of two masked copies of the word by PRMT, which also merges them into the table
// address; the squares are added two at a time by the packed FADD2 (acc2[j] = sums of individuals 2j and 2j+1).
template <bool DIAG, bool OKP>
__device__ __forceinline__ void emit_word(uint32_t w, uint32_t row_addr, uint16_t* zdst, uint8_t* okdst, int nvalid, unsigned long long (&acc2)[8])
{
    const uint32_t ev = (w << 3) & 0x78787878u;   // 8 x nibbles 0, 2, 4, 6 in bytes 0..3
    const uint32_t od = (w >> 1) & 0x78787878u;   // 8 x nibbles 1, 3, 5, 7
    uint32_t o[8], k[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint32_t a;
        const uint32_t src = (j & 1) ? od : ev;
        // byte 0 = the nibble's offset, bytes 1..3 = those of the row address
        if ((j >> 1) == 0) asm("prmt.b32 %0, %1, %2, 0x7650;" : "=r"(a) : "r"(src), "r"(row_addr));
        else if ((j >> 1) == 1) asm("prmt.b32 %0, %1, %2, 0x7651;" : "=r"(a) : "r"(src), "r"(row_addr));
        else if ((j >> 1) == 2) asm("prmt.b32 %0, %1, %2, 0x7652;" : "=r"(a) : "r"(src), "r"(row_addr));
        else asm("prmt.b32 %0, %1, %2, 0x7653;" : "=r"(a) : "r"(src), "r"(row_addr));
        if (OKP) asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(o[j]), "=r"(k[j]) : "r"(a));
        else asm volatile("ld.shared.u32 %0, [%1];" : "=r"(o[j]) : "r"(a));
        if (DIAG) {
            unsigned long long q;
            asm volatile("ld.shared.b64 %0, [%1+128];" : "=l"(q) : "r"(a));
            asm("add.rn.f32x2 %0, %0, %1;" : "+l"(acc2[j]) : "l"(q));
        }
    }
    uint32_t ok4[4];
    if (OKP) {
#pragma unroll
        for (int j = 0; j < 4; j++) ok4[j] = (k[2 * j] & 0xFFFFu) | (k[2 * j + 1] << 16);
    }
    if (nvalid >= 16) {
        st_z256(zdst, o);
        if (OKP) asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(okdst), "r"(ok4[0]), "r"(ok4[1]), "r"(ok4[2]), "r"(ok4[3]));
    } else {   // the one column that straddles n_out, and the zero padding behind it
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (2 * j >= nvalid) o[j] = 0;
            else if (2 * j + 1 >= nvalid) o[j] &= 0xFFFFu;
        }
        reinterpret_cast<uint4*>(zdst)[0] = make_uint4(o[0], o[1], o[2], o[3]);
        reinterpret_cast<uint4*>(zdst)[1] = make_uint4(o[4], o[5], o[6], o[7]);
        if (OKP) {
#pragma unroll
            for (int j = 0; j < 4; j++)
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (4 * j + e >= nvalid) ok4[j] &= ~(0xFFu << (8 * e));
            *reinterpret_cast<uint4*>(okdst) = make_uint4(ok4[0], ok4[1], ok4[2], ok4[3]);
        }
    }
}

template <bool DIAG, bool OKP>
__device__ __forceinline__ void run_rows(const FusedParams& p, int64_t r0, int64_t r1, int64_t col, bool in_z, bool can_load, int nvalid,
                                         uint8_t (*tab)[F_TB][256], int64_t (*row_off)[F_TB])
{
    const int tid = threadIdx.x;
    const int64_t i0 = col * 16;
    unsigned long long acc2[8];    // float pairs: the running sums of z^2 of individuals (2j, 2j+1)
    uint32_t m16[8];
#pragma unroll
    for (int g = 0; g < 8; g++) { m16[g] = 0; acc2[g] = 0ull; }
    const bool tdiag = DIAG && in_z && i0 < p.diag_f1 && i0 + 16 > p.diag_f0;
    int rows_acc = 0;

    auto build = [&](int buf, int64_t rbase) {
        const int nr = (int)min((int64_t)F_TB, r1 - rbase);
        for (int idx = tid; idx < nr * 16; idx += F_THREADS) {
            const int row = idx >> 4, nib = idx & 15;
            const int64_t k = p.k0 + rbase + row;
            const float za = __ldg(p.zlut + k * 8 + (nib & 3)), zb = __ldg(p.zlut + k * 8 + (nib >> 2));
            // E4M3 1.0 = 0x38 for a non-missing genotype (code != 01b), 0 for a missing one
            const uint32_t okp = ((nib & 3) != 1 ? 0x38u : 0u) | ((nib >> 2) != 1 ? 0x3800u : 0u);
            *reinterpret_cast<uint2*>(&tab[buf][row][8 * nib]) = make_uint2(f2h16(za, p.bf16) | (f2h16(zb, p.bf16) << 16), okp);
            if (DIAG) *reinterpret_cast<float2*>(&tab[buf][row][128 + 8 * nib]) = make_float2(za * za, zb * zb);   // exact products rounded once to float
        }
        for (int row = tid; row < nr; row += F_THREADS) {
            const int64_t k = p.k0 + rbase + row;
            row_off[buf][row] = (p.include ? (int64_t)p.include[k] : k) * p.pitch;
        }
    };
    auto flush = [&]() {
        if (tdiag) {
#pragma unroll
            for (int e = 0; e < 16; e++) {
                const int64_t f = i0 + e;
                if (f >= p.diag_f0 && f < p.diag_f1 && f < p.n_file) {
                    // m16[g]: low half = individual g, high half = individual 8+g; m16[4+g]: individuals 4+g, 12+g
                    const int g = e & 3, hi = (e >> 2) & 1, top = e >> 3;
                    const int miss = (int)((m16[g + 4 * hi] >> (16 * top)) & 0xFFFFu);
                    const float sum = __uint_as_float((uint32_t)(acc2[e >> 1] >> (32 * (e & 1))));
                    atomicAdd(p.diag + f, (double)sum);
                    atomicAdd(p.nii + f, rows_acc - miss);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < 8; g++) { m16[g] = 0; acc2[g] = 0ull; }
        rows_acc = 0;
    };

    const uint8_t* colp = p.bed + col * 4;
    uint32_t wn[6];            // the packed words of the next group of six rows
    bool have_next = false;
    auto load6 = [&](uint32_t (&dst)[6], int b, int64_t rb, int row) {
#pragma unroll
        for (int u = 0; u < 6; u++) {
            const int64_t off = p.include ? row_off[b][row + u] : (p.k0 + rb + row + u) * p.pitch;
            dst[u] = can_load ? __ldg(reinterpret_cast<const uint32_t*>(colp + off)) : 0u;
        }
    };
    int buf = 0, blocks = 0;
    build(0, r0);
    __syncthreads();
    for (int64_t rbase = r0; rbase < r1; rbase += F_TB, buf ^= 1) {
        if (rbase + F_TB < r1) build(buf ^ 1, rbase + F_TB);   // next block's tables while this one is consumed
        const int nr = (int)min((int64_t)F_TB, r1 - rbase);
        if (in_z) {
            uint16_t* zrow = p.Z + rbase * p.ldz + i0;
            uint8_t* okrow = OKP ? p.OK + rbase * p.ldz + i0 : nullptr;
            const uint32_t tab_addr = smem_u32(&tab[buf][0][0]);
            uint32_t c8[4] = {0, 0, 0, 0};
            int row = 0;
            if (!have_next && nr >= 6) load6(wn, buf, rbase, 0);
            for (; row + 6 <= nr; row += 6) {
                uint32_t w[6];
#pragma unroll
                for (int u = 0; u < 6; u++) w[u] = wn[u];
                // the next six rows' words are fetched before this group is processed, so that their latency hides behind
                // its ~250 instructions - also across a table block's end when the row addresses need no table
                have_next = false;
                if (row + 12 <= nr) { load6(wn, buf, rbase, row + 6); have_next = true; }
                else if (!p.include && nr == F_TB && row + 6 == nr && rbase + F_TB + 6 <= r1) { load6(wn, buf ^ 1, rbase + F_TB, 0); have_next = true; }
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    uint32_t c2 = 0;
#pragma unroll
                    for (int u = 3 * h; u < 3 * h + 3; u++) {
                        emit_word<DIAG, OKP>(w[u], tab_addr + (uint32_t)(row + u) * 256u, zrow + (int64_t)(row + u) * p.ldz,
                                             OKP ? okrow + (int64_t)(row + u) * p.ldz : nullptr, nvalid, acc2);
                        if (DIAG) c2 += (w[u] & 0x55555555u) & ~(w[u] >> 1);   // missing = code 01b, one count per 2-bit lane
                    }
                    if (DIAG) {
#pragma unroll
                        for (int g = 0; g < 4; g++) c8[g] += (c2 >> (2 * g)) & 0x03030303u;
                    }
                }
            }
            for (; row < nr; row++) {
                const int64_t off = p.include ? row_off[buf][row] : (p.k0 + rbase + row) * p.pitch;
                const uint32_t w = can_load ? __ldg(reinterpret_cast<const uint32_t*>(colp + off)) : 0u;
                emit_word<DIAG, OKP>(w, tab_addr + (uint32_t)row * 256u, zrow + (int64_t)row * p.ldz, OKP ? okrow + (int64_t)row * p.ldz : nullptr, nvalid, acc2);
                if (DIAG) {
                    const uint32_t c2 = (w & 0x55555555u) & ~(w >> 1);
#pragma unroll
                    for (int g = 0; g < 4; g++) c8[g] += (c2 >> (2 * g)) & 0x03030303u;
                }
            }
            if (DIAG) {
                // byte b of c8[g] = individual 4b + g (at most F_TB per byte) -> 16-bit lanes
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    m16[g] += c8[g] & 0x00FF00FFu;
                    m16[4 + g] += (c8[g] >> 8) & 0x00FF00FFu;
                }
                rows_acc += nr;
                if (++blocks == F_FLUSH_BLOCKS) { flush(); blocks = 0; }
            }
        }
        __syncthreads();   // the next block's tables are complete; this block's may be overwritten
    }
    if (DIAG) flush();
}

__global__ void __launch_bounds__(F_THREADS, 4) decode_emit_diag_kernel(const FusedParams p)
{
    __shared__ __align__(256) uint8_t tab[2][F_TB][256];   // per SNP row 256 bytes, see emit_word
    __shared__ int64_t row_off[2][F_TB];                   // byte offset of each SNP row
    const int64_t col = (int64_t)blockIdx.x * F_THREADS + threadIdx.x;   // packed 32-bit column = individuals [16 col, 16 col + 16)
    const int64_t i0 = col * 16;
    const bool in_z = i0 < p.ldz;                        // threads behind ldz only help with the tables and the barriers
    const bool can_load = in_z && col * 4 < p.pitch;
    const int nvalid = (int)max((int64_t)0, min((int64_t)16, p.n_out - i0));
    const int64_t r0 = (int64_t)blockIdx.y * p.rows_per_cta, r1 = min(r0 + p.rows_per_cta, p.kc);
    // does the slab hold individuals whose diagonal this band needs?  (uniform per CTA)
    const int64_t s0 = (int64_t)blockIdx.x * F_SLAB;
    const bool cta_diag = p.diag != nullptr && s0 < p.diag_f1 && s0 + F_SLAB > p.diag_f0;
    if (r0 < r1) {
        if (p.OK) {
            if (cta_diag) run_rows<true, true>(p, r0, r1, col, in_z, can_load, nvalid, tab, row_off);
            else run_rows<false, true>(p, r0, r1, col, in_z, can_load, nvalid, tab, row_off);
        } else {
            if (cta_diag) run_rows<true, false>(p, r0, r1, col, in_z, can_load, nvalid, tab, row_off);
            else run_rows<false, false>(p, r0, r1, col, in_z, can_load, nvalid, tab, row_off);
        }
    }
    // padding rows of the last K block: zero
    if (blockIdx.y == gridDim.y - 1 && in_z) {
        for (int64_t kk = p.kc; kk < p.kc_pad; kk++) {
            uint4* d = reinterpret_cast<uint4*>(p.Z + kk * p.ldz + i0);
            d[0] = make_uint4(0, 0, 0, 0);
            d[1] = make_uint4(0, 0, 0, 0);
        }
        if (p.OK)
            for (int64_t kk = p.kc; kk < p.ok_pad; kk++) *reinterpret_cast<uint4*>(p.OK + kk * p.ldz + i0) = make_uint4(0, 0, 0, 0);
    }
}

// ------------------------------------------------------------------------------------------------
// K1c: a keep list (--keep / --remove, data.cpp:657-673) on the packed rows.  One thread writes one 32-bit word of the
// output row = the 2-bit codes of 16 kept individuals, fetched from wherever the keep list says they sit in the file
// row.  The output is a packed chunk in which "every individual is kept", so K1s / K1f (and the indicator kernel) run
// their plain paths on it; the per-individual gathers of the row kernels in decode.cu are gone.
// ------------------------------------------------------------------------------------------------
constexpr int C_ROWS = 64;   // SNP rows per CTA of K1c: the 16 keep indices of a thread are fetched once and serve all of them

__global__ void __launch_bounds__(256) compact_rows_kernel(const uint8_t* __restrict__ bed, int64_t pitch, const int32_t* __restrict__ include,
                                                            int64_t k0, int64_t kc, const int32_t* __restrict__ keep, int64_t n,
                                                            uint32_t* __restrict__ out, int64_t words_out)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= words_out) return;
    const int64_t i0 = w * 16;
    const int cnt = (int)max((int64_t)0, min((int64_t)16, n - i0));
    // file positions of this thread's 16 kept individuals (the same for every SNP row); unused slots repeat the last one
    int f[16];
#pragma unroll
    for (int e = 0; e < 16; e++) f[e] = 0;
    if (cnt == 16) {
        const int4* kp = reinterpret_cast<const int4*>(keep + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int4 f4 = __ldg(kp + q);
            f[4 * q] = f4.x; f[4 * q + 1] = f4.y; f[4 * q + 2] = f4.z; f[4 * q + 3] = f4.w;
        }
    } else {
#pragma unroll
        for (int e = 0; e < 16; e++)
            if (e < cnt) f[e] = keep[i0 + e];
    }
    const int64_t kk_beg = (int64_t)blockIdx.y * C_ROWS, kk_end = min(kc, kk_beg + C_ROWS);
    // The kept individuals of a thread form a few RUNS of file neighbours (a keep list drops a few per cent of the
    // sample).  A run's codes are one funnel shift of two adjacent words of the file row, masked and moved to the run's
    // place in the output word: ~7 instructions per run and row instead of ~6 per individual.  Up to four runs are handled
    // this way (their descriptors are per-thread constants); a thread with more falls back to the walk below.
    int nruns = 0;
    int r_wi[4] = {0, 0, 0, 0}, r_sh[4] = {0, 0, 0, 0}, r_dst[4] = {0, 0, 0, 0}, r_len[4] = {0, 0, 0, 0};
    if (cnt == 16) {
#pragma unroll
        for (int e = 0; e < 16; e++) {
            const bool start = e == 0 || f[e] != f[e - 1] + 1;
            if (start) {
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (j == nruns) { r_wi[j] = f[e] >> 4; r_sh[j] = (f[e] & 15) * 2; r_dst[j] = 2 * e; }
                nruns++;
            }
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (j == nruns - 1) r_len[j]++;
        }
    }
    if (cnt == 16 && nruns <= 4) {
        const int last_word = (int)(pitch / 4) - 1;
        uint32_t r_mask[4];
        int r_wi2[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            r_mask[j] = r_len[j] >= 16 ? 0xFFFFFFFFu : ((1u << (2 * r_len[j])) - 1u);   // an unused run: length 0, mask 0
            r_wi2[j] = min(r_wi[j] + 1, last_word);   // the high word is only needed when the run crosses into it
        }
        for (int64_t kk = kk_beg; kk < kk_end; kk++) {
            const int64_t k = k0 + kk;
            const uint32_t* row32 = reinterpret_cast<const uint32_t*>(bed + (include ? (int64_t)include[k] : k) * pitch);
            uint32_t o = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t bits = __funnelshift_r(__ldg(row32 + r_wi[j]), __ldg(row32 + r_wi2[j]), r_sh[j]);
                o |= (bits & r_mask[j]) << r_dst[j];
            }
            out[kk * words_out + w] = o;
        }
        return;
    }
    for (int64_t kk = kk_beg; kk < kk_end; kk++) {
        const int64_t k = k0 + kk;
        const uint32_t* row32 = reinterpret_cast<const uint32_t*>(bed + (include ? (int64_t)include[k] : k) * pitch);
        // a word is only fetched when the next individual has left the current one
        uint32_t o = 0, curw = 0;
        int curi = -1;
#pragma unroll
        for (int e = 0; e < 16; e++) {
            const int wi = f[e] >> 4;
            if (wi != curi) { curw = __ldg(row32 + wi); curi = wi; }
            if (e < cnt) o |= ((curw >> ((f[e] & 15) * 2)) & 3u) << (2 * e);
        }
        out[kk * words_out + w] = o;   // pad individuals: code 00, masked out by the keep mask / n_out downstream
    }
}

}  // namespace gb

// K1s through the C ABI (device pointers).  d_mu / d_cnt / d_sum / d_zlut / d_total_missing as for gcta_b200_bed_decode.
extern "C" int gcta_b200_snp_stats(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const uint32_t* d_keepmask, int64_t n,
                                   const int32_t* d_include, int64_t k0, int64_t kc, const double* d_mu_in, int32_t alg,
                                   double* d_mu, int32_t* d_cnt, int32_t* d_sum, float* d_zlut,
                                   unsigned long long* d_total_missing, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_keepmask && d_mu && d_cnt && d_sum && d_zlut && d_total_missing, "snp_stats: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0 && ((uintptr_t)d_bed & 15) == 0, "snp_stats: bed pitch and pointer must be multiples of 16 bytes");
    GB_ARG(pitch * 4 >= n_file && n_file > 0 && n > 0 && n <= n_file, "snp_stats: pitch too small for n_file, or bad n");
    GB_ARG(kc > 0 && k0 >= 0, "snp_stats: empty SNP range");
    GB_ARG(((uintptr_t)d_zlut & 15) == 0, "snp_stats: zlut must be 16-byte aligned");
    int dev = 0, sms = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int per_sm = 0;   // resident CTAs per SM: the grid is exactly one wave
    GB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, snp_stats_kernel, 256, 0));
    const int64_t ctas = std::min<int64_t>((kc + 7) / 8, (int64_t)sms * std::max(per_sm, 1));
    snp_stats_kernel<<<(unsigned)ctas, 256, 0, (cudaStream_t)stream>>>(d_bed, pitch, n, d_keepmask, d_include, k0, kc, d_mu_in,
                                                                                  alg == 0, d_mu, d_cnt, d_sum, d_zlut, d_total_missing);
    GB_CUDA(cudaGetLastError());
    return 0;
}

// K1f through the C ABI (device pointers).  Every file individual is kept (n == n_file), autosomes, one 16-bit plane.
extern "C" int gcta_b200_decode_emit_diag(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_include,
                                          int64_t k0, int64_t kc, const float* d_zlut, int32_t operand, void* d_Z, int64_t ldz,
                                          double* d_diag, int32_t* d_nii, int64_t diag_f0, int64_t diag_f1, void* d_OK,
                                          void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_zlut && d_Z, "decode_emit_diag: null pointer");
    GB_ARG((d_diag == nullptr) == (d_nii == nullptr), "decode_emit_diag: d_diag and d_nii go together");
    GB_ARG(pitch > 0 && pitch % 16 == 0 && ((uintptr_t)d_bed & 15) == 0, "decode_emit_diag: bed pitch and pointer must be multiples of 16 bytes");
    GB_ARG(pitch * 4 >= n_file && n_file > 0, "decode_emit_diag: pitch too small for n_file");
    GB_ARG(ldz > 0 && ldz % 256 == 0, "decode_emit_diag: ldz must be a positive multiple of 256");
    GB_ARG(((uintptr_t)d_Z & 31) == 0 && ((uintptr_t)d_OK & 15) == 0, "decode_emit_diag: Z must be 32-byte aligned, the indicator plane 16-byte aligned");
    GB_ARG(kc > 0 && k0 >= 0, "decode_emit_diag: empty SNP range");
    GB_ARG(operand == GCTA_B200_FP16 || operand == GCTA_B200_BF16, "decode_emit_diag: operand must be FP16 or BF16");
    FusedParams p;
    p.bed = d_bed; p.pitch = pitch; p.n_file = n_file; p.include = d_include; p.k0 = k0; p.kc = kc; p.kc_pad = (kc + 63) / 64 * 64;
    p.zlut = d_zlut; p.bf16 = operand == GCTA_B200_BF16; p.Z = (uint16_t*)d_Z; p.ldz = ldz;
    p.n_out = n_file < ldz ? n_file : ldz;   // a Z narrower than n keeps only the first ldz individuals (row bands need no more)
    p.diag = d_diag; p.nii = d_nii;
    p.OK = (uint8_t*)d_OK; p.ok_pad = (kc + 127) / 128 * 128;
    p.diag_f0 = d_diag ? std::max<int64_t>(0, diag_f0) : 0;
    p.diag_f1 = d_diag ? std::min<int64_t>(n_file, diag_f1) : 0;
    int dev = 0, sms = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int64_t slabs = (ldz / 16 + F_THREADS - 1) / F_THREADS;
    // Rows per CTA.  All CTAs do the same work (except those of a partly filled last slab), so the launch should be a whole
    // number of waves of the resident capacity (4 CTAs per SM): groups of rows = (capacity / slabs) x waves, with about
    // 384 rows per CTA - long enough to amortise the table set-up and the final atomics, short enough for a few waves.
    const int64_t rows_env = getenv("GCTA_B200_DECODE_ROWS") ? atoll(getenv("GCTA_B200_DECODE_ROWS")) : 0;
    const int64_t per_wave = std::max<int64_t>(1, (int64_t)sms * 4 / slabs);
    const int64_t waves = std::max<int64_t>(1, (kc + per_wave * 384 - 1) / (per_wave * 384));
    int64_t groups = per_wave * waves;
    int64_t rows = rows_env > 0 ? rows_env : (kc + groups - 1) / groups;
    rows = std::max<int64_t>((rows + 5) / 6 * 6, 6);
    groups = (kc + rows - 1) / rows;
    GB_ARG(groups <= 65535, "decode_emit_diag: SNP range too long for one launch");
    p.rows_per_cta = rows;
    dim3 grid((unsigned)slabs, (unsigned)groups);
    decode_emit_diag_kernel<<<grid, F_THREADS, 0, (cudaStream_t)stream>>>(p);
    GB_CUDA(cudaGetLastError());
    return 0;
}

// K1c through the C ABI (device pointers): packed rows of included SNPs [k0, k0+kc) restricted to the kept individuals,
// row kk of the output at d_out + kk * pitch_out (pitch_out % 16 == 0, >= ceil(n/4)); d_keep 16-byte aligned.
extern "C" int gcta_b200_compact_rows(const uint8_t* d_bed, int64_t pitch, const int32_t* d_include, int64_t k0, int64_t kc,
                                      const int32_t* d_keep, int64_t n, uint8_t* d_out, int64_t pitch_out, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_keep && d_out, "compact_rows: null pointer");
    GB_ARG(pitch > 0 && n > 0 && kc > 0 && k0 >= 0, "compact_rows: empty range");
    GB_ARG(pitch_out % 16 == 0 && pitch_out * 4 >= n && ((uintptr_t)d_out & 15) == 0 && ((uintptr_t)d_keep & 15) == 0,
           "compact_rows: output pitch must be a multiple of 16 bytes that holds n individuals; d_out and d_keep 16-byte aligned");
    const int64_t words = pitch_out / 4;
    for (int64_t r0 = 0; r0 < kc; r0 += 65535ll * C_ROWS) {   // grid.y limit
        const int64_t rows = std::min<int64_t>(65535ll * C_ROWS, kc - r0);
        dim3 grid((unsigned)((words + 255) / 256), (unsigned)((rows + C_ROWS - 1) / C_ROWS));
        compact_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_bed, pitch, d_include, k0 + r0, rows, d_keep, n,
                                                                   reinterpret_cast<uint32_t*>(d_out + r0 * pitch_out), words);
    }
    GB_CUDA(cudaGetLastError());
    return 0;
}
