// finalize.cu - K4: A = S / N with the reference's zero-count rules, packed to .grm.gz order.
//
// grm.cpp:273-275 (diagonal: A_ii = S_ii / N_ii, 1.0 when N_ii == 0) and grm.cpp:294
// (off-diagonal: A_ij = S_ij / N_ij when N_ij > 0, else the zero A was initialised with).
// The diagonal uses the exact float64 sum of squares from K1d instead of the 16-bit tensor
// core result, so the mean of the diagonal does not depend on the operand type.
//
// One CTA per 128x256 tile; 16 tile rows at a time are transposed through shared memory so
// that both the tile reads (512 B runs of the [col/4][row][4] layout) and the packed
// lower-triangle writes (runs along j) are coalesced.  HBM-bound: 4 B (+4 B for N) read and
// 4 B (+4 B) written per pair.
#include "common.cuh"

namespace gb {

constexpr int FIN_ROWS = 16;

__global__ void __launch_bounds__(256) finalize_tiles_kernel(const float* __restrict__ S, const int32_t* __restrict__ Nt,
                                                              int32_t m_const, const int2* __restrict__ tiles, int64_t n,
                                                              const int32_t* __restrict__ keep, const double* __restrict__ diag,
                                                              const int32_t* __restrict__ nii, int64_t row_begin,
                                                              int64_t row_end, float* __restrict__ A_lower,
                                                              int32_t* __restrict__ N_lower, int alg1,
                                                              const float* __restrict__ Dt, double d_const,
                                                              const double* __restrict__ dw)
{
    __shared__ float a_s[FIN_ROWS][TILE_N + 1];
    __shared__ int32_t n_s[FIN_ROWS][TILE_N + 1];
    const int2 tl = tiles[blockIdx.x];
    const int64_t row0 = (int64_t)tl.x * TILE_M, col0 = (int64_t)tl.y * TILE_N;
    const float4* St = reinterpret_cast<const float4*>(S + (size_t)blockIdx.x * (TILE_M * TILE_N));
    const int4* Ntile = Nt ? reinterpret_cast<const int4*>(Nt + (size_t)blockIdx.x * (TILE_M * TILE_N)) : nullptr;
    const float4* Dtile = Dt ? reinterpret_cast<const float4*>(Dt + (size_t)blockIdx.x * (TILE_M * TILE_N)) : nullptr;
    const int64_t base = row_begin * (row_begin + 1) / 2;  // outputs are relative to the band start

    for (int r0 = 0; r0 < TILE_M; r0 += FIN_ROWS) {
        if (row0 + r0 >= n || row0 + r0 >= row_end) break;
        if (row0 + r0 + FIN_ROWS <= row_begin || col0 > row0 + r0 + FIN_ROWS - 1) continue;  // nothing of the triangle here
#pragma unroll
        for (int it = 0; it < FIN_ROWS * (TILE_N / 4) / 256; it++) {
            const int u = threadIdx.x + it * 256;
            const int c4 = u / FIN_ROWS, rr = u % FIN_ROWS;
            const float4 s = St[(size_t)c4 * TILE_M + r0 + rr];
            int4 cnt = make_int4(m_const, m_const, m_const, m_const);
            if (Ntile) cnt = Ntile[(size_t)c4 * TILE_M + r0 + rr];
            const float sv[4] = {s.x, s.y, s.z, s.w};
            const int cv[4] = {cnt.x, cnt.y, cnt.z, cnt.w};
            float4 dd = make_float4((float)d_const, (float)d_const, (float)d_const, (float)d_const);
            if (Dtile) dd = Dtile[(size_t)c4 * TILE_M + r0 + rr];
            const float dv[4] = {dd.x, dd.y, dd.z, dd.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                // off-diagonal rule, grm.cpp:294 (alg 1: divide by sum_k ok ok d_k instead of the count, grm.cpp:285);
                // the diagonal is overwritten below
                const double den = alg1 ? (Dtile ? (double)dv[e] : d_const) : (double)cv[e];
                a_s[rr][c4 * 4 + e] = cv[e] > 0 ? (float)((double)sv[e] / den) : 0.0f;
                n_s[rr][c4 * 4 + e] = cv[e];
            }
        }
        __syncthreads();
        for (int rr = 0; rr < FIN_ROWS; rr++) {
            const int64_t i = row0 + r0 + rr;
            if (i >= n || i < row_begin || i >= row_end) continue;
            const int64_t j = col0 + threadIdx.x;
            if (j > i) continue;
            float a = a_s[rr][threadIdx.x];
            int32_t cn = n_s[rr][threadIdx.x];
            if (j == i) {
                const int64_t f = keep ? keep[i] : i;
                cn = nii[f];
                a = cn > 0 ? (float)(diag[f] / (alg1 ? dw[f] : (double)cn)) : 1.0f;  // grm.cpp:264, 273-275
            }
            const int64_t o = i * (i + 1) / 2 + j - base;
            A_lower[o] = a;
            if (N_lower) N_lower[o] = cn;
        }
        __syncthreads();
    }
}

}  // namespace gb

extern "C" int gcta_b200_finalize_tiles(const float* d_S, const int32_t* d_Ntiles, int32_t m_const, const int32_t* d_tiles,
                                        int64_t ntiles, int64_t n, const int32_t* d_keep, const double* d_diag,
                                        const int32_t* d_nii, int64_t row_begin, int64_t row_end, float* d_A_lower,
                                        int32_t* d_N_lower, void* stream, char* err, size_t errcap)
{
    return gcta_b200_finalize_tiles_ex(d_S, d_Ntiles, m_const, nullptr, 0.0, nullptr, d_tiles, ntiles, n, d_keep, d_diag, d_nii, row_begin,
                                       row_end, d_A_lower, d_N_lower, stream, err, errcap);
}

extern "C" int gcta_b200_finalize_tiles_ex(const float* d_S, const int32_t* d_Ntiles, int32_t m_const, const float* d_Dtiles,
                                           double d_const, const double* d_dw, const int32_t* d_tiles, int64_t ntiles, int64_t n,
                                           const int32_t* d_keep, const double* d_diag, const int32_t* d_nii, int64_t row_begin,
                                           int64_t row_end, float* d_A_lower, int32_t* d_N_lower, void* stream, char* err,
                                           size_t errcap)
{
    using namespace gb;
    GB_ARG(d_S && d_tiles && d_diag && d_nii && d_A_lower, "finalize: null pointer");
    GB_ARG(row_begin >= 0 && row_begin <= row_end && row_end <= n, "finalize: bad row band");
    if (ntiles == 0) return 0;
    finalize_tiles_kernel<<<(unsigned)ntiles, 256, 0, (cudaStream_t)stream>>>(d_S, d_Ntiles, m_const,
                                                                             reinterpret_cast<const int2*>(d_tiles), n, d_keep,
                                                                             d_diag, d_nii, row_begin, row_end, d_A_lower, d_N_lower, d_dw != nullptr,
                                                                             d_Dtiles, d_const, d_dw);
    GB_CUDA(cudaGetLastError());
    return 0;
}
