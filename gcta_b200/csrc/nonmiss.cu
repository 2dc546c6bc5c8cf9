// nonmiss.cu - K1n: the non-missing indicator matrix OK[k][i] in {0, 1} as 16-bit floats, SNP-major
// [kc_pad][ldz] like Z, so that the SYRK kernel can contract it:  N_ij = sum_k OK_ki * OK_kj
// (grm.cpp:270, 291).  Every product is 0 or 1 and every partial sum an integer below 2^24, so the
// FP32 tensor-core accumulation is exact and N comes out bit-identical to the AND+popc kernel (K3),
// at the tensor rate instead of the popc issue rate (8x faster on the 2 %-missing configuration).
//
// HBM-bound: ceil(n/4) bytes read and 2*ldz bytes written per SNP; one warp per SNP row, 8 individuals
// (one 16-byte store) per lane and iteration.
#include "common.cuh"

namespace gb {

__global__ void __launch_bounds__(256) nonmiss_matrix_kernel(const uint8_t* __restrict__ bed, int64_t pitch,
                                                              const int32_t* __restrict__ keep, int64_t n,
                                                              const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                              int64_t kc_pad, uint32_t one16, const double* __restrict__ dk, int bf16,
                                                              uint16_t* __restrict__ OK, int64_t ldz)
{
    const int lane = threadIdx.x & 31;
    const int64_t kk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (kk >= kc_pad) return;
    uint4* orow = reinterpret_cast<uint4*>(OK + kk * ldz);
    const int nchunk = (int)(ldz / 8);
    if (kk >= kc) {  // padding row of the last 64-SNP block
        for (int c = lane; c < nchunk; c += 32) orow[c] = make_uint4(0, 0, 0, 0);
        return;
    }
    const int64_t k = k0 + kk;
    if (dk) {   // --grm-alg 1: indicator scaled by sqrt(d_k), so that the contraction gives sum_k ok ok d_k
        const float w = sqrtf((float)dk[k]);
        one16 = bf16 ? (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(w)) : (uint32_t)__half_as_ushort(__float2half_rn(w));
    }
    const uint8_t* row = bed + (include ? (int64_t)include[k] : k) * pitch;
    const uint16_t* row2 = reinterpret_cast<const uint16_t*>(row);
    for (int ch = lane; ch < nchunk; ch += 32) {
        const int64_t i0 = (int64_t)ch * 8;
        uint32_t o[4] = {0, 0, 0, 0};
        if (i0 < n) {
#pragma unroll
            for (int e = 0; e < 8; e++) {
                uint32_t code = 1;  // treated as missing beyond n
                if (i0 + e < n) {
                    if (!keep) code = ((uint32_t)__ldg(row2 + ch) >> (2 * e)) & 3u;
                    else {
                        const int f = keep[i0 + e];
                        code = (row[f >> 2] >> ((f & 3) * 2)) & 3u;
                    }
                }
                o[e >> 1] |= (code != 1u ? one16 : 0u) << (16 * (e & 1));   // 01b = missing (data.cpp:176-181, 792)
            }
        }
        orow[ch] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// 8-bit variant: E4M3 1.0 (0x38) / 0 per genotype, [kc_pad][ldz] bytes, kc_pad a multiple of 128 (the K block of the
// 8-bit SYRK); one warp per SNP row, 16 individuals (one 16-byte store) per lane and iteration.
__global__ void __launch_bounds__(256) nonmiss_matrix8_kernel(const uint8_t* __restrict__ bed, int64_t pitch,
                                                               const int32_t* __restrict__ keep, int64_t n,
                                                               const int32_t* __restrict__ include, int64_t k0, int64_t kc,
                                                               int64_t kc_pad, uint8_t* __restrict__ OK, int64_t ldz)
{
    const int lane = threadIdx.x & 31;
    const int64_t kk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (kk >= kc_pad) return;
    uint4* orow = reinterpret_cast<uint4*>(OK + kk * ldz);
    const int nchunk = (int)(ldz / 16);
    if (kk >= kc) {
        for (int c = lane; c < nchunk; c += 32) orow[c] = make_uint4(0, 0, 0, 0);
        return;
    }
    const int64_t k = k0 + kk;
    const uint8_t* row = bed + (include ? (int64_t)include[k] : k) * pitch;
    const uint32_t* row4 = reinterpret_cast<const uint32_t*>(row);
    for (int ch = lane; ch < nchunk; ch += 32) {
        const int64_t i0 = (int64_t)ch * 16;
        uint32_t o[4] = {0, 0, 0, 0};
        if (i0 < n) {
            const uint32_t w = keep ? 0u : __ldg(row4 + ch);
#pragma unroll
            for (int e = 0; e < 16; e++) {
                uint32_t code = 1;
                if (i0 + e < n) {
                    if (!keep) code = (w >> (2 * e)) & 3u;
                    else {
                        const int f = keep[i0 + e];
                        code = (row[f >> 2] >> ((f & 3) * 2)) & 3u;
                    }
                }
                o[e >> 2] |= (code != 1u ? 0x38u : 0u) << (8 * (e & 3));
            }
        }
        orow[ch] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

}  // namespace gb

extern "C" int gcta_b200_nonmiss_matrix(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_keep, int64_t n,
                                        const int32_t* d_include, int64_t k0, int64_t kc, int32_t operand, void* d_OK, int64_t ldz,
                                        void* stream, char* err, size_t errcap)
{
    return gcta_b200_nonmiss_matrix_ex(d_bed, pitch, n_file, d_keep, n, d_include, k0, kc, operand, nullptr, d_OK, ldz, stream, err, errcap);
}

extern "C" int gcta_b200_nonmiss_matrix_ex(const uint8_t* d_bed, int64_t pitch, int64_t n_file, const int32_t* d_keep, int64_t n,
                                           const int32_t* d_include, int64_t k0, int64_t kc, int32_t operand, const double* d_dk,
                                           void* d_OK, int64_t ldz, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_bed && d_OK, "nonmiss_matrix: null pointer");
    GB_ARG(pitch > 0 && pitch % 16 == 0, "nonmiss_matrix: pitch must be a multiple of 16");
    GB_ARG(kc > 0 && k0 >= 0, "nonmiss_matrix: empty SNP range");
    GB_ARG(ldz > 0 && ldz % 256 == 0, "nonmiss_matrix: ldz must be a positive multiple of 256");
    GB_ARG(d_keep || n == n_file, "nonmiss_matrix: n must equal n_file when keep is NULL");
    if (n > ldz) n = ldz;   // a matrix narrower than n keeps only the first ldz individuals
    if (operand == GCTA_B200_E4M3) {   // 8-bit indicators for the kind::f8f6f4 counting pass: rows padded to 128
        GB_ARG(!d_dk, "nonmiss_matrix: the 8-bit matrix holds plain 0/1 indicators only");
        const int64_t kc_pad8 = (kc + 127) / 128 * 128;
        nonmiss_matrix8_kernel<<<(unsigned)((kc_pad8 + 7) / 8), 256, 0, (cudaStream_t)stream>>>(d_bed, pitch, d_keep, n, d_include, k0, kc, kc_pad8,
                                                                                               (uint8_t*)d_OK, ldz);
        GB_CUDA(cudaGetLastError());
        return 0;
    }
    GB_ARG(operand == GCTA_B200_FP16 || operand == GCTA_B200_BF16, "nonmiss_matrix: unknown operand type");
    const int64_t kc_pad = (kc + 63) / 64 * 64;
    const uint32_t one16 = operand == GCTA_B200_BF16 ? 0x3F80u : 0x3C00u;   // 1.0 in bf16 / fp16
    nonmiss_matrix_kernel<<<(unsigned)((kc_pad + 7) / 8), 256, 0, (cudaStream_t)stream>>>(d_bed, pitch, d_keep, n, d_include, k0, kc, kc_pad,
                                                                                         one16, d_dk, operand == GCTA_B200_BF16, (uint16_t*)d_OK, ldz);
    GB_CUDA(cudaGetLastError());
    return 0;
}
