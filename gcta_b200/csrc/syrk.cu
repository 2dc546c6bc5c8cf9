// syrk.cu - K2: lower-triangle SYRK  S += Z^T Z  on the 5th-generation tensor cores.
//
// Replaces the numerator of the reference's triple loop (grm.cpp:276-296):
//     S_ij = sum_k z_ik * z_jk
// Z is the 16-bit standardised genotype chunk written by K1, SNP-major [kc_pad][ldz]
// (rows = SNPs = the contraction axis, columns = individuals), so both MMA operands are
// "MN-major" tiles of the same matrix and no transpose is ever materialised.
//
// One persistent CTA per SM walks a static list of 256x256 output tiles.  A tile is two
// UMMA M=128 halves sharing one B operand, each with its own 256-column TMEM accumulator
// (all 512 columns in use), so every 64-SNP stage moves 64 KB for 2 x (128x256x64) MACs -
// a third less L2->SM traffic per flop than a 128x256 tile, and a 2x larger wave footprint
// for L2 reuse:
//   warp 0 (one lane) : TMA producer - 8 boxes of [64 SNPs][64 individuals] per stage into a
//                       3-stage ring of 128B-swizzled shared memory
//   warp 1 (one lane) : tcgen05.mma issuer - UMMA 128x256x16, kind::f16, FP32 accumulate
//   warps 2-9         : epilogue - tcgen05.ld the accumulators, add them into the FP32 S tile
//                       in HBM (S carries the sum over SNP chunks)
#include <cuda.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace gb {

constexpr int NUM_STAGES = 3;
constexpr int STAGE_BYTES = 65536;   // 256 A rows + 256 B columns x one K block, for either element size
// Element-size dependent geometry.  EB = 2: FP16 / BF16 operands, kind::f16, UMMA K = 16.  EB = 1: E4M3 operands
// (0/1 indicators of the counting pass), kind::f8f6f4, UMMA K = 32 - twice the MMA rate on half the bytes.
// A 128-byte swizzle row holds 128/EB individuals; a K block is sized so that one group is 128 B x KB rows.
template <int EB> struct Geo {
    static constexpr int GROUP_MN = 128 / EB;                     // individuals per swizzle atom row
    static constexpr int KB = 64 * (2 / EB);                      // SNPs per pipeline stage: 64 (16-bit) or 128 (8-bit)
    static constexpr int GROUP_BYTES = KB * 128;                  // 8192 or 16384
    static constexpr int A_GROUPS = TILE_M / GROUP_MN;            // 4 or 2
    static constexpr int B_GROUPS = TILE_N / GROUP_MN;
    static constexpr int UK = 32 / EB;                            // SNPs per tcgen05.mma
    static constexpr int K_STEP_BYTES = UK * 128;                 // 2048 or 4096
    static_assert((A_GROUPS + B_GROUPS) * GROUP_BYTES == STAGE_BYTES, "stage size");
};
constexpr int EPI_WARPS = 8;
constexpr int SYRK_THREADS = 64 + 32 * EPI_WARPS;             // 320
constexpr int TMEM_COLS = 512;
constexpr size_t SYRK_SMEM = (size_t)NUM_STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

// Shared-memory matrix descriptor for an MN-major, 128B-swizzled operand whose 64-individual
// groups are GROUP_BYTES apart: canonical layout ((8,8,m),(8,k)) : ((1,8,LBO),(64,SBO)) in
// 16-bit elements, LBO = byte distance between 64-wide MN groups, SBO = byte distance between
// groups of 8 SNP rows (8 x 128 B).  Bits: [0,14) addr>>4, [16,30) LBO>>4, [32,46) SBO>>4,
// [46,48) version = 1 (Blackwell), [61,64) layout = 2 (SWIZZLE_128B).
template <int EB>
__device__ __forceinline__ uint64_t make_mn_desc(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(Geo<EB>::GROUP_BYTES >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}

// D[tmem] (+)= A[smem] * B[smem], kind::f8f6f4 (E4M3 in, FP32 accumulate); SASS UTCQMMA
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// INT_ACC = false: S (float) += accumulator.  INT_ACC = true: the operands are 0/1 indicators, every
// partial sum is an integer below 2^24 and therefore exact in FP32, and the epilogue adds it into an
// int32 tile - the pair counts N_ij on the tensor cores, bit-exact.
template <bool INT_ACC, int EB = 2>
__global__ void __launch_bounds__(SYRK_THREADS, 1)
syrk_tiles_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                  const int2* __restrict__ tiles, int ntiles, int nkb, float* __restrict__ S, uint32_t idesc,
                  unsigned int* wave_counter, int epi_red, int ckb)
{
    using G = Geo<EB>;
    // Work items = (chain, tile), chain-major: the K blocks of the launch are cut into chains of `ckb` blocks, each of
    // which is one accumulation in TMEM followed by an epilogue into S (ckb == nkb: one chain, items == tiles).  Running
    // all chains of a Z chunk in ONE launch leaves one partly filled last wave per chunk instead of one per chain.
    const int nchains = (nkb + ckb - 1) / ckb;
    const int nitems = ntiles * nchains;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + NUM_STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + NUM_STAGES;
    uint64_t* tfull_bar = empty_bar + NUM_STAGES;
    uint64_t* tempty_bar = tfull_bar + 1;
    uint32_t* tmem_slot = (uint32_t*)(tempty_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapB);
        for (int s = 0; s < NUM_STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, EPI_WARPS);
        mbar_fence_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            const int nwaves = (nitems + (int)gridDim.x - 1) / (int)gridDim.x;
            for (int w = 0; w < nwaves; w++) {
                const int u = blockIdx.x + w * (int)gridDim.x;
                if (wave_counter) {
                    // Wave barrier (cooperative launch): every CTA starts wave w together.  CTAs that share
                    // an operand panel then stream it within the few k-blocks of history L2 keeps per CTA, instead of
                    // drifting apart over the launch and re-fetching it from HBM.
                    atomicAdd(wave_counter, 1u);
                    const unsigned int target = gridDim.x * (unsigned int)(w + 1);
                    while (*reinterpret_cast<volatile unsigned int*>(wave_counter) < target) {}
                }
                if (u >= nitems) break;
                const int chain = u / ntiles;
                const int2 tl = tiles[u - chain * ntiles];
                const int row0 = tl.x * TILE_M, col0 = tl.y * TILE_N;
                const int kb0 = chain * ckb, kb1 = min(nkb, kb0 + ckb);
                for (int kb = kb0; kb < kb1; kb++) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = smem + stage * STAGE_BYTES;
                    mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
#pragma unroll
                    for (int g = 0; g < G::A_GROUPS; g++)
                        tma_load_2d(st + g * G::GROUP_BYTES, &tmapA, &full_bar[stage], row0 + g * G::GROUP_MN, kb * G::KB);
#pragma unroll
                    for (int g = 0; g < G::B_GROUPS; g++)
                        tma_load_2d(st + (G::A_GROUPS + g) * G::GROUP_BYTES, &tmapB, &full_bar[stage], col0 + g * G::GROUP_MN,
                                    kb * G::KB);
                    if (++stage == NUM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int u = blockIdx.x; u < nitems; u += gridDim.x, it++) {
                mbar_wait(tempty_bar, (it & 1) ^ 1);  // epilogue has drained the previous tile
                tc_fence_after();
                const int kb0 = (u / ntiles) * ckb, kb1 = min(nkb, kb0 + ckb);
                for (int kb = kb0; kb < kb1; kb++) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t a0_desc = make_mn_desc<EB>(a_addr);                                               // rows   0..127
                    const uint64_t a1_desc = make_mn_desc<EB>(a_addr + (UMMA_M / G::GROUP_MN) * G::GROUP_BYTES);     // rows 128..255
                    const uint64_t b_desc = make_mn_desc<EB>(a_addr + G::A_GROUPS * G::GROUP_BYTES);
#pragma unroll
                    for (int k4 = 0; k4 < G::KB / G::UK; k4++) {
                        // one MMA consumes UK SNPs = UK/8 swizzle atoms of 8 rows x 128 B further along K
                        const uint64_t adv = (uint64_t)((k4 * G::K_STEP_BYTES) >> 4);
                        const uint32_t acc = (kb != kb0 || k4 != 0) ? 1u : 0u;   // a chain starts from zero
                        if (EB == 2) {
                            umma_f16(tmem_base, a0_desc + adv, b_desc + adv, idesc, acc);
                            umma_f16(tmem_base + TILE_N, a1_desc + adv, b_desc + adv, idesc, acc);
                        } else {
                            umma_f8(tmem_base, a0_desc + adv, b_desc + adv, idesc, acc);
                            umma_f8(tmem_base + TILE_N, a1_desc + adv, b_desc + adv, idesc, acc);
                        }
                    }
                    umma_commit(&empty_bar[stage]);  // frees the smem stage when these MMAs retire
                    if (++stage == NUM_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(tfull_bar);  // both accumulators complete -> epilogue
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue (warps 2..9) ----------------
        const int q = warp & 3;              // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;    // which accumulator / 128-row half of the tile
        const int r = half * UMMA_M + q * 32 + lane;
        int it = 0;
        for (int u = blockIdx.x; u < nitems; u += gridDim.x, it++) {
            mbar_wait(tfull_bar, it & 1);
            tc_fence_after();
            const int t = u % ntiles;
            float4* Stile = reinterpret_cast<float4*>(S + (size_t)t * (TILE_M * TILE_N));
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * TILE_N;
            if (!INT_ACC && epi_red) {
                // Fire-and-forget vector reductions in L2 (red.global.add.v4.f32): every element of S belongs to exactly one
                // CTA of a launch and launches are stream-ordered, so the sum is the same round-to-nearest add as a
                // read-modify-write, without the epilogue waiting for S to arrive from HBM.  The TMEM reads are ping-ponged
                // (the next 32 columns are in flight while the current ones go out), and the accumulators are handed back to
                // the MMA issuer as soon as the last columns sit in registers - before their reductions are issued.
                auto red8 = [&](int c0, const uint32_t (&v)[32]) {
#pragma unroll
                    for (int g = 0; g < 8; g++) {
                        float4* p = Stile + (size_t)(c0 / 4 + g) * TILE_M + r;
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v[4 * g + 0]), "r"(v[4 * g + 1]),
                                     "r"(v[4 * g + 2]), "r"(v[4 * g + 3])
                                     : "memory");
                    }
                };
                uint32_t va[32], vb[32];
                tmem_ld_32x32(taddr, va);
#pragma unroll 1
                for (int c0 = 0; c0 < TILE_N; c0 += 64) {
                    tmem_ld_wait();                        // va = columns c0 .. c0+31
                    tmem_ld_32x32(taddr + c0 + 32, vb);
                    red8(c0, va);
                    tmem_ld_wait();                        // vb = columns c0+32 .. c0+63
                    if (c0 + 64 < TILE_N) tmem_ld_32x32(taddr + c0 + 64, va);
                    else {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty_bar);
                    }
                    red8(c0 + 32, vb);
                }
                continue;
            }
#pragma unroll 1
            for (int c0 = 0; c0 < TILE_N; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(taddr + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int g = 0; g < 8; g++) {
                    float4* p = Stile + (size_t)(c0 / 4 + g) * TILE_M + r;
                    if (INT_ACC) {
                        int4* pi = reinterpret_cast<int4*>(p);
                        int4 o = *pi;
                        o.x += __float2int_rn(__uint_as_float(v[4 * g + 0]));
                        o.y += __float2int_rn(__uint_as_float(v[4 * g + 1]));
                        o.z += __float2int_rn(__uint_as_float(v[4 * g + 2]));
                        o.w += __float2int_rn(__uint_as_float(v[4 * g + 3]));
                        *pi = o;
                    } else {
                        float4 o = *p;
                        o.x += __uint_as_float(v[4 * g + 0]);
                        o.y += __uint_as_float(v[4 * g + 1]);
                        o.z += __uint_as_float(v[4 * g + 2]);
                        o.w += __uint_as_float(v[4 * g + 3]);
                        *p = o;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// arrival counters of the wave barrier: one slot per (device, stream).  Launches on one stream run in order, and each
// resets its slot by a memset on that stream, so a slot is never shared by two launches in flight; a process that uses
// more than WAVE_SLOTS streams per device runs the extra ones without the barrier.
constexpr int WAVE_SLOTS = 64;
static __device__ unsigned int g_wave_counter[WAVE_SLOTS];

namespace {
std::mutex g_slot_mu;
std::map<std::pair<int, void*>, int> g_slots;
std::vector<int> g_free_slots[64];
int g_slots_used[64] = {};
}  // namespace

int wave_slot_of(int dev, void* stream)
{
    std::lock_guard<std::mutex> lock(g_slot_mu);
    auto it = g_slots.find({dev, stream});
    if (it != g_slots.end()) return it->second;
    if (dev < 0 || dev >= 64) return -1;
    int slot = -1;
    if (!g_free_slots[dev].empty()) { slot = g_free_slots[dev].back(); g_free_slots[dev].pop_back(); }
    else if (g_slots_used[dev] < WAVE_SLOTS) slot = g_slots_used[dev]++;
    if (slot >= 0) g_slots[{dev, stream}] = slot;
    return slot;
}

// a stream is about to be destroyed (plan_destroy): its slot goes back to the pool, so a long-lived process that creates
// many plans never runs out of counters
void wave_slot_release(int dev, void* stream)
{
    std::lock_guard<std::mutex> lock(g_slot_mu);
    auto it = g_slots.find({dev, stream});
    if (it == g_slots.end()) return;
    if (dev >= 0 && dev < 64) g_free_slots[dev].push_back(it->second);
    g_slots.erase(it);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

}  // namespace gb

static int syrk_launch(bool int_acc, const void* d_Z, const void* d_Zb, int64_t ldz, int64_t kc_pad, int32_t operand, const int32_t* d_tiles,
                       int64_t ntiles, float* d_S, void* stream, char* err, size_t errcap, int64_t chain_snps = 0)
{
    using namespace gb;
    GB_ARG(d_Z && d_Zb && d_tiles && d_S, "syrk: null pointer");
    GB_ARG(ldz > 0 && ldz % TILE_N == 0, "syrk: ldz must be a positive multiple of 256");
    const int eb = operand == GCTA_B200_E4M3 ? 1 : 2;      // bytes per operand element
    const int kblock = eb == 1 ? Geo<1>::KB : Geo<2>::KB;  // SNPs per pipeline stage
    GB_ARG(kc_pad > 0 && kc_pad % kblock == 0, eb == 1 ? "syrk: kc_pad must be a positive multiple of 128 for 8-bit operands"
                                                        : "syrk: kc_pad must be a positive multiple of 64");
    GB_ARG(((uintptr_t)d_Z & 127) == 0 && ((uintptr_t)d_Zb & 127) == 0, "syrk: Z must be 128-byte aligned");
    GB_ARG(operand == GCTA_B200_FP16 || operand == GCTA_B200_BF16 || (operand == GCTA_B200_E4M3 && int_acc),
           "syrk: unknown operand type (E4M3 is for the counting mode only)");
    if (ntiles == 0) return 0;
    int dev = 0, sms = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return set_err(err, errcap, GCTA_B200_ERR_NO_DEVICE, "Error: cuTensorMapEncodeTiled is not available from the driver.");

    CUtensorMap tmap, tmapB;
    cuuint64_t gdim[2] = {(cuuint64_t)ldz, (cuuint64_t)kc_pad};
    cuuint64_t gstride[1] = {(cuuint64_t)ldz * eb};
    cuuint32_t box[2] = {(cuuint32_t)(128 / eb), (cuuint32_t)kblock};   // 128 bytes of individuals x one K block
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = eb == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                           : (operand == GCTA_B200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
    CUresult cr = enc(&tmap, dt, 2, const_cast<void*>(d_Z), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr == CUDA_SUCCESS)   // the B operand may come from another matrix of the same shape (split-FP16 cross terms)
        cr = enc(&tmapB, dt, 2,
                 const_cast<void*>(d_Zb), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS)
        return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: cuTensorMapEncodeTiled failed with code " + std::to_string((int)cr));

    // instruction descriptor: D=F32 [4,6), A fmt [7,10), B fmt [10,13) (kind::f16: 0 = F16, 1 = BF16; kind::f8f6f4: 0 = E4M3),
    // A MN-major bit 15, B MN-major bit 16, N>>3 [17,23), M>>4 [24,29)
    const uint32_t fmt = operand == GCTA_B200_BF16 ? 1u : 0u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(TILE_N >> 3) << 17) |
                           ((uint32_t)(UMMA_M >> 4) << 24);

    static bool attr_set[64] = {};
    if (dev < 64 && !attr_set[dev]) {
        GB_CUDA(cudaFuncSetAttribute(syrk_tiles_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SYRK_SMEM));
        GB_CUDA(cudaFuncSetAttribute(syrk_tiles_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SYRK_SMEM));
        GB_CUDA(cudaFuncSetAttribute(syrk_tiles_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SYRK_SMEM));
        attr_set[dev] = true;
    }
    const int2* tl = reinterpret_cast<const int2*>(d_tiles);
    int nt = (int)ntiles, nkb = (int)(kc_pad / kblock);
    uint32_t id = idesc;
    typedef void (*KernelFn)(const CUtensorMap, const CUtensorMap, const int2*, int, int, float*, uint32_t, unsigned int*, int, int);
    const KernelFn kfn = eb == 1 ? (KernelFn)syrk_tiles_kernel<true, 1> : (int_acc ? (KernelFn)syrk_tiles_kernel<true> : (KernelFn)syrk_tiles_kernel<false>);
    // GCTA_B200_EPI_RED=0: epilogue as read-modify-write instead of red.global.add.v4.f32
    static const int epi_red_env = !(getenv("GCTA_B200_EPI_RED") && atoi(getenv("GCTA_B200_EPI_RED")) == 0);
    int epi_red = epi_red_env && !int_acc;
    // chains inside one launch need the reduction epilogue (two chains of a tile are added into S by different CTAs)
    int ckb = nkb;
    if (chain_snps > 0 && chain_snps < kc_pad) {
        GB_ARG(epi_red && chain_snps % kblock == 0, "syrk: chains inside a launch need the reduction epilogue and a chain length that is a multiple of the K block");
        ckb = (int)(chain_snps / kblock);
    }
    const int64_t nitems = (int64_t)nt * ((nkb + ckb - 1) / ckb);
    GB_ARG(nitems < (1ll << 31), "syrk: too many work items for one launch");
    int grid = (int)(nitems < sms ? nitems : sms);
    unsigned int* counter = nullptr;
    // Wave barrier on by default (cfg3: +10 % whole-job, HBM reads of K2 -40 %; profiles/README.md); GCTA_B200_WAVESYNC=0 disables
    static const bool wave_sync = !(getenv("GCTA_B200_WAVESYNC") && atoi(getenv("GCTA_B200_WAVESYNC")) == 0);
    if (wave_sync && nitems > grid) {
        static std::atomic<unsigned int*> cached[64] = {};
        const int slot = wave_slot_of(dev, stream);
        if (slot >= 0) {
            unsigned int* base = cached[dev].load();
            if (!base) {
                GB_CUDA(cudaGetSymbolAddress((void**)&base, g_wave_counter));
                cached[dev].store(base);
            }
            counter = base + slot;
        }
    }
    if (counter) {
        // the barrier needs all CTAs resident at once: cooperative launch guarantees it (or fails loudly)
        GB_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), (cudaStream_t)stream));
        void* args[] = {&tmap, &tmapB, &tl, &nt, &nkb, &d_S, &id, &counter, &epi_red, &ckb};
        cudaError_t ce = cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid), dim3(SYRK_THREADS), args, SYRK_SMEM, (cudaStream_t)stream);
        if (ce == cudaErrorCooperativeLaunchTooLarge || ce == cudaErrorNotSupported || ce == cudaErrorLaunchOutOfResources) {
            // e.g. part of the GPU is held by another context: run the same kernel without the barrier
            cudaGetLastError();
            counter = nullptr;
            kfn<<<grid, SYRK_THREADS, SYRK_SMEM, (cudaStream_t)stream>>>(tmap, tmapB, tl, nt, nkb, d_S, id, counter, epi_red, ckb);
        } else
            GB_CUDA(ce);
    } else
        kfn<<<grid, SYRK_THREADS, SYRK_SMEM, (cudaStream_t)stream>>>(tmap, tmapB, tl, nt, nkb, d_S, id, counter, epi_red, ckb);
    GB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int gcta_b200_syrk_tiles(const void* d_Z, int64_t ldz, int64_t kc_pad, int32_t operand, const int32_t* d_tiles,
                                    int64_t ntiles, float* d_S, void* stream, char* err, size_t errcap)
{
    return syrk_launch(false, d_Z, d_Z, ldz, kc_pad, operand, d_tiles, ntiles, d_S, stream, err, errcap);
}

// S += Za^T Zb on the same tile list: rows of a tile come from Za, columns from Zb (two matrices of one shape)
extern "C" int gcta_b200_syrk_tiles_ab(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int32_t operand,
                                       const int32_t* d_tiles, int64_t ntiles, float* d_S, void* stream, char* err, size_t errcap)
{
    return syrk_launch(false, d_Za, d_Zb, ldz, kc_pad, operand, d_tiles, ntiles, d_S, stream, err, errcap);
}

// S += Za^T Zb with the contraction cut into chains of chain_snps SNPs (see the kernel): one launch when the reduction
// epilogue is on (the default), else one launch per chain
extern "C" int gcta_b200_syrk_tiles_chained(const void* d_Za, const void* d_Zb, int64_t ldz, int64_t kc_pad, int64_t chain_snps, int32_t operand,
                                            const int32_t* d_tiles, int64_t ntiles, float* d_S, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(chain_snps > 0 && chain_snps % 128 == 0, "syrk_tiles_chained: chain_snps must be a positive multiple of 128");
    const bool red = !(getenv("GCTA_B200_EPI_RED") && atoi(getenv("GCTA_B200_EPI_RED")) == 0);
    const bool one_launch = !(getenv("GCTA_B200_CHAIN_LAUNCHES") && atoi(getenv("GCTA_B200_CHAIN_LAUNCHES")) != 0);
    if (chain_snps >= kc_pad) return syrk_launch(false, d_Za, d_Zb, ldz, kc_pad, operand, d_tiles, ntiles, d_S, stream, err, errcap);
    // Two chains of one tile are added into S by different CTAs.  With at least four waves of tiles they run waves apart
    // (items are chain-major and every wave starts at a grid-wide barrier), so the additions keep their order and the
    // result is reproducible bit for bit; smaller tile lists get one launch per chain.
    int dev = 0, sms = 0;
    GB_CUDA(cudaGetDevice(&dev));
    GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (red && one_launch && ntiles >= 4 * (int64_t)sms)
        return syrk_launch(false, d_Za, d_Zb, ldz, kc_pad, operand, d_tiles, ntiles, d_S, stream, err, errcap, chain_snps);
    for (int64_t ks = 0; ks < kc_pad; ks += chain_snps) {
        int rc = syrk_launch(false, (const uint16_t*)d_Za + ks * ldz, (const uint16_t*)d_Zb + ks * ldz, ldz, std::min<int64_t>(chain_snps, kc_pad - ks),
                             operand, d_tiles, ntiles, d_S, stream, err, errcap);
        if (rc) return rc;
    }
    return 0;
}

extern "C" int gcta_b200_count_tiles(const void* d_OK, int64_t ldz, int64_t kc_pad, int32_t operand, const int32_t* d_tiles,
                                     int64_t ntiles, int32_t* d_Ntiles, void* stream, char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(kc_pad < (1 << 24), "count_tiles: more than 2^24 SNPs per call would not be exact in FP32");
    return syrk_launch(true, d_OK, d_OK, ldz, kc_pad, operand, d_tiles, ntiles, reinterpret_cast<float*>(d_Ntiles), stream, err, errcap);
}
