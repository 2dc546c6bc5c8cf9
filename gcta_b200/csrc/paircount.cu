// paircount.cu - K3: bit-exact per-pair non-missing counts  N_ij = sum_k ok_ik * ok_jk.
//
// Replaces the `A_N[i][j]++` of the reference's triple loop (grm.cpp:270, 291) with
// AND + popc over the non-missing bit planes written by K1m (32 SNPs per uint32 word,
// individual-minor: planes[w][i]).  Same 256x256 tile list and tile storage layout as the
// SYRK kernel, int32 accumulators.  The kernel is bound by the popc issue rate, not by
// memory: each CTA stages 8 plane words x (128 + 256) individuals in shared memory and every
// thread keeps an 8 x 16 block of pair counters in registers.
//
// Skipped entirely by the plan when no kept genotype is missing (then N_ij == M, which is
// exactly what the reference produces).
#include "common.cuh"

namespace gb {

constexpr int PC_WORDS = 8;  // plane words staged per step
constexpr int PC_M = 128;    // rows of pairs per CTA (half an S tile)

__global__ void __launch_bounds__(256) paircount_tiles_kernel(const uint32_t* __restrict__ planes, int64_t ldm, int nwords,
                                                               const int2* __restrict__ tiles, int64_t n, int32_t* __restrict__ Nt)
{
    __shared__ __align__(16) uint32_t a_s[PC_WORDS][PC_M];
    __shared__ __align__(16) uint32_t b_s[PC_WORDS][TILE_N];
    const int tile = blockIdx.x / (TILE_M / PC_M), half = blockIdx.x % (TILE_M / PC_M);
    const int2 tl = tiles[tile];
    const int64_t row0 = (int64_t)tl.x * TILE_M + half * PC_M, col0 = (int64_t)tl.y * TILE_N;
    if (row0 >= n || row0 + PC_M - 1 < col0) return;  // beyond the last individual or wholly above the diagonal
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // thread block: rows ty*8.., columns tx*16..
    int acc[8][16];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 16; j++) acc[i][j] = 0;

    for (int w0 = 0; w0 < nwords; w0 += PC_WORDS) {
        // stage: 8 x 128 + 8 x 256 words, coalesced along the individual axis
        for (int idx = threadIdx.x; idx < PC_WORDS * (PC_M + TILE_N); idx += 256) {
            const int ww = idx / (PC_M + TILE_N), o = idx % (PC_M + TILE_N);
            const int w = w0 + ww;
            uint32_t v = 0;
            if (w < nwords) v = o < PC_M ? __ldg(planes + (int64_t)w * ldm + row0 + o) : __ldg(planes + (int64_t)w * ldm + col0 + (o - PC_M));
            if (o < PC_M) a_s[ww][o] = v; else b_s[ww][o - PC_M] = v;
        }
        __syncthreads();
#pragma unroll
        for (int ww = 0; ww < PC_WORDS; ww++) {
            uint32_t a[8], b[16];
            const uint4* ap = reinterpret_cast<const uint4*>(&a_s[ww][ty * 8]);
            const uint4* bp = reinterpret_cast<const uint4*>(&b_s[ww][tx * 16]);
#pragma unroll
            for (int q = 0; q < 2; q++) { const uint4 t = ap[q]; a[4 * q] = t.x; a[4 * q + 1] = t.y; a[4 * q + 2] = t.z; a[4 * q + 3] = t.w; }
#pragma unroll
            for (int q = 0; q < 4; q++) { const uint4 t = bp[q]; b[4 * q] = t.x; b[4 * q + 1] = t.y; b[4 * q + 2] = t.z; b[4 * q + 3] = t.w; }
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 16; j++) acc[i][j] += __popc(a[i] & b[j]);
        }
        __syncthreads();
    }
    // accumulate into the tile: element (r, c) at [(c/4)*TILE_M + r]*4 + c%4
    int4* T = reinterpret_cast<int4*>(Nt + (size_t)tile * (TILE_M * TILE_N));
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int g = 0; g < 4; g++) {
            int4* p = T + (size_t)(tx * 4 + g) * TILE_M + (half * PC_M + ty * 8 + i);
            int4 o = *p;
            o.x += acc[i][4 * g]; o.y += acc[i][4 * g + 1]; o.z += acc[i][4 * g + 2]; o.w += acc[i][4 * g + 3];
            *p = o;
        }
}

}  // namespace gb

extern "C" int gcta_b200_paircount_tiles(const uint32_t* d_planes, int64_t ldm, int64_t nwords, const int32_t* d_tiles,
                                         int64_t ntiles, int32_t* d_Ntiles, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(d_planes && d_tiles && d_Ntiles, "paircount: null pointer");
    GB_ARG(ldm > 0 && ldm % TILE_N == 0, "paircount: ldm must be a positive multiple of 256");
    GB_ARG(nwords > 0 && nwords < (1ll << 31), "paircount: bad word count");
    if (ntiles == 0) return 0;
    paircount_tiles_kernel<<<(unsigned)(ntiles * (TILE_M / PC_M)), 256, 0, (cudaStream_t)stream>>>(
        d_planes, ldm, (int)nwords, reinterpret_cast<const int2*>(d_tiles), ldm, d_Ntiles);
    GB_CUDA(cudaGetLastError());
    return 0;
}
