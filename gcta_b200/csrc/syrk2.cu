// syrk2.cu - K2 as CTA pairs: tcgen05.mma.cta_group::2 (two SMs on one 512 x 256 block of S).
//
// Same contraction, tile storage and epilogue as syrk.cu (S += Za^T Zb on the lower-triangle tiles), but the two
// CTAs of a cluster work on two vertically adjacent 256 x 256 tiles that need the SAME 256 columns of Z:
// each CTA stages its own 256 rows (A, 32 KB per 64-SNP stage) and only HALF of the shared columns (B, 16 KB);
// a cta_group::2 MMA (M = 256 = 128 rows per CTA, N = 256) reads the B halves of both CTAs.  Per stage and SM that
// is 48 KB from L2 instead of 64 KB, and half the B bytes read from shared memory by the tensor core.
//   both CTAs, warp 0 (one lane) : TMA producer; completion bytes of BOTH CTAs land on the leader's mbarrier
//   leader CTA, warp 1 (one lane): issues 2 x 4 UMMA 256x256x16 per stage (accumulators 0 and 1 of both CTAs),
//                                  commits multicast to both CTAs' barriers
//   both CTAs, warps 2-9         : epilogue of their own TMEM accumulators (own 256 x 256 tile)
#include <cuda.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <map>
#include <vector>

#include "common.cuh"

namespace gb {

constexpr int P_STAGES = 4;
constexpr int P_GROUP_MN = 64;
constexpr int P_GROUP_BYTES = BLOCK_K * P_GROUP_MN * 2;            // 8192
constexpr int P_A_GROUPS = TILE_M / P_GROUP_MN;                    // 4 (this CTA's 256 rows)
constexpr int P_B_GROUPS = TILE_N / 2 / P_GROUP_MN;                // 2 (this CTA's half of the 256 columns)
constexpr int P_STAGE_BYTES = (P_A_GROUPS + P_B_GROUPS) * P_GROUP_BYTES;   // 49152
constexpr int P_EPI_WARPS = 8;
constexpr int P_THREADS = 64 + 32 * P_EPI_WARPS;                   // 320
constexpr size_t P_SMEM = (size_t)P_STAGES * P_STAGE_BYTES + 1024 + 256;

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load of one CTA of a pair; the completion bytes are credited to the mbarrier at the same offset in the
// EVEN CTA of the pair (peer bit of the shared::cluster address cleared)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const void* tmap, uint64_t* leader_bar_local, int c0, int c1)
{
    const uint32_t bar = smem_u32(leader_bar_local) & 0xFEFFFFFFu;
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_result, uint32_t ncols)
{
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols)
{
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma2_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the mbarrier at this offset in both CTAs of the pair once the MMAs issued so far have completed
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ uint64_t make_mn_desc2(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(P_GROUP_BYTES >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}

// pairs[p] = {bj, bi0, t0, t1}: column block, first of the two row blocks, storage index of tile (bi0, bj) and of
// tile (bi0 + 1, bj) in the tile-packed S (-1: that tile is not part of the band / triangle - computed, not stored)
template <bool INT_ACC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(P_THREADS, 1)
syrk_pairs_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                  const int4* __restrict__ pairs, int npairs, int nkb, float* __restrict__ S, uint32_t idesc,
                  unsigned int* wave_counter)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + P_STAGES * P_STAGE_BYTES);
    uint64_t* empty_bar = full_bar + P_STAGES;
    uint64_t* tfull_bar = empty_bar + P_STAGES;
    uint64_t* tempty_bar = tfull_bar + 1;
    uint32_t* tmem_slot = (uint32_t*)(tempty_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int ncl = (int)gridDim.x / 2, cl = (int)blockIdx.x / 2;
    const int nwaves = (npairs + ncl - 1) / ncl;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapB);
        for (int s = 0; s < P_STAGES; s++) {
            mbar_init(&full_bar[s], 1);    // leader: its own arrive.expect_tx; bytes come from both CTAs
            mbar_init(&empty_bar[s], 1);   // one multicast commit from the leader's MMA thread
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 2 * P_EPI_WARPS);   // leader's copy collects the epilogue warps of both CTAs
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc2(tmem_slot, 512);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------- TMA producer (both CTAs) ----------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int w = 0; w < nwaves; w++) {
                const int p = cl + w * ncl;
                if (wave_counter) {   // wave barrier over the whole grid, see syrk.cu
                    atomicAdd(wave_counter, 1u);
                    const unsigned int target = gridDim.x * (unsigned int)(w + 1);
                    while (*reinterpret_cast<volatile unsigned int*>(wave_counter) < target) {}
                }
                if (p >= npairs) break;
                const int4 it = pairs[p];
                const int row0 = (it.y + (int)rank) * TILE_M, colh = it.x * TILE_N + (int)rank * (TILE_N / 2);
                for (int kb = 0; kb < nkb; kb++) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = smem + stage * P_STAGE_BYTES;
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * P_STAGE_BYTES);
#pragma unroll
                    for (int g = 0; g < P_A_GROUPS; g++)
                        tma_load_2d_pair(st + g * P_GROUP_BYTES, &tmapA, &full_bar[stage], row0 + g * P_GROUP_MN, kb * BLOCK_K);
#pragma unroll
                    for (int g = 0; g < P_B_GROUPS; g++)
                        tma_load_2d_pair(st + (P_A_GROUPS + g) * P_GROUP_BYTES, &tmapB, &full_bar[stage], colh + g * P_GROUP_MN,
                                         kb * BLOCK_K);
                    if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---------------- MMA issuer (leader CTA only) ----------------
        if (lane == 0 && rank == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int w = 0; w < nwaves; w++) {
                const int p = cl + w * ncl;
                if (p >= npairs) break;
                mbar_wait(tempty_bar, (w & 1) ^ 1);   // both CTAs have drained the previous accumulators
                tc_fence_after();
                for (int kb = 0; kb < nkb; kb++) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * P_STAGE_BYTES);
                    const uint64_t a0_desc = make_mn_desc2(a_addr);                                           // rows   0..127 of each CTA
                    const uint64_t a1_desc = make_mn_desc2(a_addr + (UMMA_M / P_GROUP_MN) * P_GROUP_BYTES);   // rows 128..255 of each CTA
                    const uint64_t b_desc = make_mn_desc2(a_addr + P_A_GROUPS * P_GROUP_BYTES);               // 128 columns per CTA
#pragma unroll
                    for (int k4 = 0; k4 < BLOCK_K / 16; k4++) {
                        const uint64_t adv = (uint64_t)((k4 * 2048) >> 4);
                        const uint32_t acc = (kb | k4) != 0 ? 1u : 0u;
                        umma2_f16(tmem_base, a0_desc + adv, b_desc + adv, idesc, acc);
                        umma2_f16(tmem_base + TILE_N, a1_desc + adv, b_desc + adv, idesc, acc);
                    }
                    umma2_commit_both(&empty_bar[stage]);
                    if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
                }
                umma2_commit_both(tfull_bar);
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue (warps 2..9 of both CTAs) ----------------
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int r = half * UMMA_M + q * 32 + lane;
        const uint32_t leader_tempty = mapa_u32(smem_u32(tempty_bar), 0);
        for (int w = 0; w < nwaves; w++) {
            const int p = cl + w * ncl;
            if (p >= npairs) break;
            const int4 it = pairs[p];
            const int t = rank ? it.w : it.z;
            mbar_wait(tfull_bar, w & 1);
            tc_fence_after();
            if (t >= 0) {
                float4* Stile = reinterpret_cast<float4*>(S + (size_t)t * (TILE_M * TILE_N));
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * TILE_N;
#pragma unroll 1
                for (int c0 = 0; c0 < TILE_N; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c0, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 8; g++) {
                        float4* ptr = Stile + (size_t)(c0 / 4 + g) * TILE_M + r;
                        if (INT_ACC) {
                            int4* pi = reinterpret_cast<int4*>(ptr);
                            int4 o = *pi;
                            o.x += __float2int_rn(__uint_as_float(v[4 * g + 0]));
                            o.y += __float2int_rn(__uint_as_float(v[4 * g + 1]));
                            o.z += __float2int_rn(__uint_as_float(v[4 * g + 2]));
                            o.w += __float2int_rn(__uint_as_float(v[4 * g + 3]));
                            *pi = o;
                        } else {
                            float4 o = *ptr;
                            o.x += __uint_as_float(v[4 * g + 0]);
                            o.y += __uint_as_float(v[4 * g + 1]);
                            o.z += __uint_as_float(v[4 * g + 2]);
                            o.w += __uint_as_float(v[4 * g + 3]);
                            *ptr = o;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_tempty);
        }
    }

    tc_fence_before();
    cluster_sync_all();   // nobody leaves (or frees TMEM) while the peer may still signal or read this CTA
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
}

typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int WAVE_SLOTS2 = 64;
static __device__ unsigned int g_wave_counter2[WAVE_SLOTS2];

}  // namespace gb

// Pair list of a row band: entries {bj, bi0, t0, t1} in launch order (raster groups like gcta_b200_tile_list).
// `tiles` is the band's tile list (gcta_b200_tile_list) whose order defines the storage indices t0 / t1.
extern "C" int64_t gcta_b200_pair_list(const int32_t* tiles, int64_t ntiles, int32_t raster_rows, int32_t* pairs)
{
    if (!tiles || ntiles <= 0) return 0;
    std::map<std::pair<int, int>, int> index;
    int b0 = tiles[0], b1 = tiles[0];
    for (int64_t t = 0; t < ntiles; t++) {
        index[{tiles[2 * t], tiles[2 * t + 1]}] = (int)t;
        b0 = std::min(b0, tiles[2 * t]);
        b1 = std::max(b1, tiles[2 * t]);
    }
    if (raster_rows < 2) raster_rows = 2;
    raster_rows &= ~1;
    int64_t cnt = 0;
    for (int g0 = b0; g0 <= b1; g0 += raster_rows) {
        const int g1 = std::min(g0 + raster_rows, b1 + 1);
        int maxcb = 0;
        for (auto& kv : index)
            if (kv.first.first >= g0 && kv.first.first < g1) maxcb = std::max(maxcb, kv.first.second + 1);
        for (int bj = 0; bj < maxcb; bj++)
            for (int bi = g0; bi < g1; bi += 2) {
                auto i0 = index.find({bi, bj}), i1 = index.find({bi + 1, bj});
                const int t0 = i0 == index.end() ? -1 : i0->second, t1 = i1 == index.end() ? -1 : i1->second;
                if (t0 < 0 && t1 < 0) continue;
                if (pairs) { pairs[4 * cnt] = bj; pairs[4 * cnt + 1] = bi; pairs[4 * cnt + 2] = t0; pairs[4 * cnt + 3] = t1; }
                cnt++;
            }
    }
    return cnt;
}

extern "C" int gcta_b200_syrk_pairs(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int32_t operand,
                                    const int32_t* d_pairs, int64_t npairs, void* d_S, int32_t int_acc, void* stream, char* err,
                                    size_t errcap)
{
    using namespace gb;
    GB_ARG(d_Za && d_Zb && d_pairs && d_S, "syrk_pairs: null pointer");
    GB_ARG(ldz > 0 && ldz % TILE_N == 0, "syrk_pairs: ldz must be a positive multiple of 256");
    GB_ARG(kc_pad > 0 && kc_pad % BLOCK_K == 0, "syrk_pairs: kc_pad must be a positive multiple of 64");
    GB_ARG(((uintptr_t)d_Za & 127) == 0 && ((uintptr_t)d_Zb & 127) == 0, "syrk_pairs: Z must be 128-byte aligned");
    GB_ARG(operand == GCTA_B200_FP16 || operand == GCTA_B200_BF16, "syrk_pairs: unknown operand type");
    GB_ARG(!int_acc || kc_pad < (1 << 24), "syrk_pairs: more than 2^24 SNPs per call would not be exact in FP32");
    if (npairs == 0) return 0;
    int dev = 0, sms = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    static EncodeTiledFn2 enc = nullptr;
    if (!enc) {
        void* fp = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            enc = (EncodeTiledFn2)fp;
    }
    if (!enc) return set_err(err, errcap, GCTA_B200_ERR_NO_DEVICE, "Error: cuTensorMapEncodeTiled is not available from the driver.");
    CUtensorMap tmA, tmB;
    cuuint64_t gdim[2] = {(cuuint64_t)ldz, (cuuint64_t)kc_pad};
    cuuint64_t gstride[1] = {(cuuint64_t)ldz * 2};
    cuuint32_t box[2] = {P_GROUP_MN, BLOCK_K};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = operand == GCTA_B200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    CUresult cr = enc(&tmA, dt, 2, const_cast<void*>(d_Za), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr == CUDA_SUCCESS)
        cr = enc(&tmB, dt, 2, const_cast<void*>(d_Zb), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS)
        return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: cuTensorMapEncodeTiled failed with code " + std::to_string((int)cr));
    // kind::f16 instruction descriptor as in syrk.cu, but M = 256 (128 rows in each CTA of the pair)
    const uint32_t fmt = operand == GCTA_B200_BF16 ? 1u : 0u;
    uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(TILE_N >> 3) << 17) |
                     ((uint32_t)(256 >> 4) << 24);
    static bool attr_set[64] = {};
    if (dev < 64 && !attr_set[dev]) {
        GB_CUDA(cudaFuncSetAttribute(syrk_pairs_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P_SMEM));
        GB_CUDA(cudaFuncSetAttribute(syrk_pairs_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P_SMEM));
        attr_set[dev] = true;
    }
    int ncl = (int)std::min<int64_t>(npairs, sms / 2);
    const int4* pl = reinterpret_cast<const int4*>(d_pairs);
    int np = (int)npairs, nkb = (int)(kc_pad / BLOCK_K);
    float* Sp = (float*)d_S;
    unsigned int* counter = nullptr;
    static const bool wave_sync = !(getenv("GCTA_B200_WAVESYNC") && atoi(getenv("GCTA_B200_WAVESYNC")) == 0);
    cudaLaunchAttribute attrs[2];
    int nattrs = 0;
    attrs[nattrs].id = cudaLaunchAttributeClusterDimension;
    attrs[nattrs].val.clusterDim.x = 2; attrs[nattrs].val.clusterDim.y = 1; attrs[nattrs].val.clusterDim.z = 1;
    nattrs++;
    if (wave_sync && npairs > ncl) {
        static std::atomic<unsigned int*> cached[64] = {};
        const int slot = wave_slot_of(dev, stream);   // one counter per (device, stream), like syrk.cu
        if (slot >= 0) {
            unsigned int* base = cached[dev].load();
            if (!base) {
                GB_CUDA(cudaGetSymbolAddress((void**)&base, g_wave_counter2));
                cached[dev].store(base);
            }
            counter = base + slot;
        }
        if (counter) {
            GB_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), (cudaStream_t)stream));
            attrs[nattrs].id = cudaLaunchAttributeCooperative;   // the barrier needs every CTA resident
            attrs[nattrs].val.cooperative = 1;
            nattrs++;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * ncl);
    cfg.blockDim = dim3(P_THREADS);
    cfg.dynamicSmemBytes = P_SMEM;
    cfg.stream = (cudaStream_t)stream;
    cfg.attrs = attrs;
    cfg.numAttrs = nattrs;
    if (int_acc) GB_CUDA(cudaLaunchKernelEx(&cfg, syrk_pairs_kernel<true>, tmA, tmB, pl, np, nkb, Sp, idesc, counter));
    else GB_CUDA(cudaLaunchKernelEx(&cfg, syrk_pairs_kernel<false>, tmA, tmB, pl, np, nkb, Sp, idesc, counter));
    GB_CUDA(cudaGetLastError());
    return 0;
}
