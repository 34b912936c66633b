// Dense projection GEMM on the 5th-generation tensor cores:  C[M,N] = act(A[M,K] * W[N,K]^T + bias[N]).
//
// This is the Linear / 1x1-convolution shape of the path: value_proj, sampling_offsets,
// attention_weights, output_proj of MSDeformAttn (OPS/modules/ms_deform_attn.py:98-124), the encoder FFN
// (MSMF/modeling/pixel_decoder/msdeformattn.py:116-131), and the 1x1 convolutions of FusionLayers /
// the RGB-D level fusion in channels-last form (maskrefiner/modeling/mask_refiner/model.py:430-452,
// maskrefiner/modeling/backbone/resnet.py:473-485).  PyTorch stores a Linear weight as [N,K] row-major,
// i.e. BOTH operands are K-major — exactly what tcgen05.mma wants from shared memory.
//
// Structure (persistent: one CTA per SM walks 128 x BN output tiles, BN in {64, 128, 256}; the accumulator is
// double-buffered in TMEM so that the epilogue of one tile overlaps the MMAs of the next — with K = 256 a tile's main loop
// is only four k-blocks and the GEMM is bound by the HBM traffic of A and C, not by the tensor pipe):
//   warp 0   one lane: TMA producer — A and W tiles (128B-swizzled, K-major) into a 3-4-stage ring that keeps running
//            across tiles, mbarrier complete_tx; rows / columns / k past the matrix edge arrive as zeros (TMA fill)
//   warp 1   allocates TMEM (2 x BN fp32 columns x 128 lanes); one lane issues tcgen05.mma (cta_group::1,
//            kind::f16 for bf16 or kind::tf32 for fp32 inputs, M = 128, N = BN, K = 32 bytes per instruction),
//            tcgen05.commit frees the smem stage / signals the epilogue
//   warps 2-5  epilogue: tcgen05.ld (32 lanes x 32 columns per instruction) -> + bias -> activation ->
//            128B-swizzled staging -> TMA store of 32-row x 128-byte boxes; hand the TMEM buffer back
//
// fp32-accurate mode (SPLIT, "3xTF32"): the reference runs these layers in fp32 (`autocast(enabled=False)`,
// msdeformattn.py:314) and a single TF32 product only carries 10 mantissa bits.  Every fp32 operand is split into
// hi = its top 19 bits (what the tensor core reads of an fp32 word) and lo = x - hi (exact in fp32), and
//   A W^T ~= A_hi W_hi^T + A_hi W_lo^T + A_lo W_hi^T          (the lo * lo term is 2^-22 relative and dropped)
// is accumulated in the same fp32 TMEM tile by three tcgen05.mma per k-step.  W_hi / W_lo are prepared once per weight
// (quber_split_tf32); A is split IN SHARED MEMORY by the otherwise idle epilogue warps between the TMA's arrival and the
// MMA issue (the transform is elementwise, so it is oblivious of the 128B swizzle), so A still crosses HBM once.
// These GEMMs sit at 85 FLOP/B (N = K = 256), far below the ridge, which is why three times the MMAs cost little.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace quber {

constexpr int G_BM = 128;
constexpr int G_THREADS = 192;          // producer, MMA issuer, 4 epilogue warps (SPLIT kernels; + 4 converter warps)
constexpr int G_THREADS8 = 320;         // ... 8 epilogue warps (all other kernels)
constexpr int G_KBYTES = 128;          // bytes of K per stage row = one 128B swizzle atom

__device__ __forceinline__ uint32_t g_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void g_mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(g_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void g_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(g_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ unsigned int g_gemm_timeouts = 0;
// false = gave up (a lost TMA / MMA must fail the call, never hang the GPU)
__device__ __forceinline__ bool g_mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    for (int spin = 0; !done; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(g_smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1 << 22)) { atomicAdd(&g_gemm_timeouts, 1u); return false; }
    }
    return true;
}
__device__ __forceinline__ void g_tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(g_smem_u32(dst)), "l"(map), "r"(g_smem_u32(bar)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(g_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate, int tf32) {
    if (tf32)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, 128-byte rows, SWIZZLE_128B: 8-row atoms of 1024 bytes (SBO), LBO unused (= 1),
// descriptor version 1 (Blackwell), layout type 2 (cute/arch/mma_sm100_desc.hpp: SmemDescriptor).
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fffu);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// Persistent kernel: one CTA per SM walks the output tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... (N tiles fastest, so
// the CTAs working on one row panel at the same time share its A tile in L2).  The accumulator is double-buffered in
// TMEM (2 x BN fp32 columns), so the MMAs of tile t + 1 run while the epilogue warps drain tile t — with K = 256 a tile is
// only four k-blocks of MMA and a one-tile-per-CTA kernel spends most of its life in the prologue / epilogue.
//   warp 0       TMA producer (A, W [, W_lo] into an NST-stage ring; the ring runs on across tile boundaries)
//   warp 1       TMEM allocation, MMA issue: waits tmem_empty[buf], issues the tile's MMAs, commits tmem_full[buf]
//   warps 2-5    epilogue: tcgen05.ld -> + bias -> activation -> swizzled staging -> TMA store; arrive on tmem_empty[buf]
//   warps 6-9    (SPLIT only) operand split of the A stage in shared memory, between the TMA's arrival and the MMA issue
template <int BN, bool TF32, typename OT, bool SPLIT = false>
__global__ void __launch_bounds__(SPLIT ? G_THREADS + 128 : G_THREADS8)
gemm_bias_act_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                     const __grid_constant__ CUtensorMap map_wlo, const __grid_constant__ CUtensorMap map_c,
                     const float *__restrict__ bias, int M, int N, int K, int act,
                     const __grid_constant__ CUtensorMap map_a2, int kb_split) {
    static_assert(!SPLIT || TF32, "the split mode is an fp32 mode");
    static_assert(BN == 64 || BN == 128 || BN == 256, "tile width");
    constexpr int NST = SPLIT ? 2 : (BN == 256 ? 3 : 4);     // SPLIT stages hold A_hi | A_lo | W_hi | W_lo: twice the bytes
    extern __shared__ __align__(1024) unsigned char gsm_raw[];
    // SWIZZLE_128B tiles must sit on 1024-byte boundaries of the SHARED address space
    unsigned char *gsm = gsm_raw + ((1024u - (g_smem_u32(gsm_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full[4], empty[4], conv[4], tmem_full[2], tmem_empty[2];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float bias_s[BN];
    __shared__ int abort_s;
    constexpr int A_BYTES = G_BM * G_KBYTES, W_BYTES = BN * G_KBYTES, STAGE = (SPLIT ? 2 : 1) * (A_BYTES + W_BYTES);
    constexpr int W_OFF = (SPLIT ? 2 : 1) * A_BYTES;    // stage layout: A [| A_lo] | W [| W_lo]
    constexpr int KELEMS = TF32 ? 32 : 64;              // elements of K per stage
    // Epilogue warps: a warp reads the TMEM lanes 32 * (warp % 4) .. + 31 only, so four warps cover a tile's rows.  With
    // K = 256 a tile's MMAs take ~1.8 us while four warps need ~3.5 us to drain, convert and stage 128 x 256 outputs: the
    // kernels that are not tensor-bound run EIGHT epilogue warps, two per lane quarter, each taking half of the columns.
    constexpr int NEPI = SPLIT ? 4 : 8;
    unsigned char *stg_base = gsm + (size_t)NST * STAGE;      // epilogue staging: NEPI warps x 2 buffers x 4 KB, after the ring
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = (K + KELEMS - 1) / KELEMS;
    const int tiles_n = (N + BN - 1) / BN, tiles_m = (M + G_BM - 1) / G_BM;
    const int ntiles = tiles_n * tiles_m;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 4; ++s) { g_mbar_init(&full[s], 1); g_mbar_init(&empty[s], 1); g_mbar_init(&conv[s], 128); }
        for (int s = 0; s < 2; ++s) { g_mbar_init(&tmem_full[s], 1); g_mbar_init(&tmem_empty[s], NEPI); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        abort_s = 0;
    }
    if (warp == 1) {                                     // TMEM: two accumulators of BN fp32 columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(g_smem_u32(&tmem_base_s)), "n"(2 * BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int it = 0;
            for (int t = blockIdx.x; t < ntiles && !abort_s; t += gridDim.x) {
                const int m0 = (t / tiles_n) * G_BM, n0 = (t % tiles_n) * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % NST;
                    if (it >= NST && !g_mbar_wait(&empty[s], ((it / NST) - 1) & 1)) { abort_s = 1; break; }
                    unsigned char *sa = gsm + (size_t)s * STAGE, *sw = sa + W_OFF;
                    g_mbar_expect_tx(&full[s], A_BYTES + (SPLIT ? 2 : 1) * W_BYTES);
                    // kb_split > 0: A is the K-concatenation [A1 | A2] of two matrices that is never materialised
                    // (quber_gemm_bias_act_cat2): k-blocks below the split come from map_a, the others from map_a2
                    if (kb_split > 0 && kb >= kb_split) g_tma_load_2d(sa, &map_a2, &full[s], (kb - kb_split) * KELEMS, m0);
                    else g_tma_load_2d(sa, &map_a, &full[s], kb * KELEMS, m0);
                    g_tma_load_2d(sw, &map_w, &full[s], kb * KELEMS, n0);
                    if (SPLIT) g_tma_load_2d(sw + W_BYTES, &map_wlo, &full[s], kb * KELEMS, n0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            // instruction descriptor (cute/arch/mma_sm100_desc.hpp: InstrDescriptor): D = F32, A/B = BF16 or TF32,
            // both K-major, N >> 3 at bit 17, M >> 4 at bit 24
            const uint32_t fmt = TF32 ? 2u : 1u;
            const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(G_BM >> 4) << 24);
            int it = 0, tc = 0;
            bool ok = true;
            for (int t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++tc) {
                const int buf = tc & 1;
                if (tc >= 2 && !g_mbar_wait(&tmem_empty[buf], ((tc >> 1) - 1) & 1)) { abort_s = 1; break; }   // the epilogue has drained it
                tc_fence_after();
                const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % NST;
                    ok = g_mbar_wait(SPLIT ? &conv[s] : &full[s], (it / NST) & 1);      // SPLIT: the converter warps have written A_hi / A_lo
                    if (!ok) { abort_s = 1; break; }
                    tc_fence_after();
                    const uint32_t sa = g_smem_u32(gsm + (size_t)s * STAGE), sw = sa + W_OFF;
                    const uint64_t ad = umma_desc_k128(sa), wd = umma_desc_k128(sw);
#pragma unroll
                    for (int k = 0; k < G_KBYTES / 32; ++k) {    // 32 bytes of K per instruction: +2 in the 16-byte address field
                        if (SPLIT) {                             // small terms first, then hi * hi
                            const uint64_t ald = umma_desc_k128(sa + A_BYTES), wld = umma_desc_k128(sw + W_BYTES);
                            tc_mma(acc, ald + 2 * k, wd + 2 * k, idesc, (kb | k) != 0, 1);
                            tc_mma(acc, ad + 2 * k, wld + 2 * k, idesc, 1, 1);
                            tc_mma(acc, ad + 2 * k, wd + 2 * k, idesc, 1, 1);
                        } else {
                            tc_mma(acc, ad + 2 * k, wd + 2 * k, idesc, (kb | k) != 0, TF32);
                        }
                    }
                    tc_commit(&empty[s]);                        // frees the stage when these MMAs have read it
                }
                tc_commit(&tmem_full[buf]);                      // accumulator complete (or abort: release the epilogue)
            }
        }
    } else if (warp < 2 + NEPI) {
        // ===== epilogue: warps 2 .. 1 + NEPI own TMEM lanes 32 * (warp % 4) .. +31 =====
        // Each warp stages its 32 rows x 128 bytes of output per chunk in its own shared-memory buffers, 128B-swizzled so
        // that the 32 row-owning lanes hit distinct banks, and hands it to a TMA store: full 128-byte row segments leave
        // the SM, rows / columns past the matrix edge are clipped by the hardware.  Two staging buffers per warp: the store
        // of chunk c overlaps the TMEM read of chunk c + 1.
        const int q = warp & 3;
        constexpr int CW = 128 / (int)sizeof(OT);            // output columns per 128-byte chunk: 32 fp32 / 64 bf16
        constexpr int NCH = BN / CW;                         // chunks per tile; this warp's share: [ch0, ch1)
        const int part = (warp - 2) >> 2;                    // 0, or 1 for the second set of four warps
        const int ch0 = NEPI == 8 ? (part == 0 ? 0 : (NCH + 1) / 2) : 0, ch1 = NEPI == 8 ? (part == 0 ? (NCH + 1) / 2 : NCH) : NCH;
        unsigned char *stg = stg_base + (size_t)(warp - 2) * 2 * 4096;
        const uint32_t sw = (uint32_t)(lane & 7);            // swizzle phase of this lane's staging row
        int tc = 0, ci = 0, last_n0 = -1;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tc) {
            const int buf = tc & 1;
            const int m0 = (t / tiles_n) * G_BM, n0 = (t % tiles_n) * BN;
            if (n0 != last_n0) {                             // the tile's bias slice (changes only when tiles_n > 1)
                asm volatile("bar.sync 1, %0;" ::"n"(NEPI * 32) : "memory");              // nobody still reads the previous slice
                for (int i = threadIdx.x - 64; i < BN; i += NEPI * 32) bias_s[i] = (bias && n0 + i < N) ? __ldg(bias + n0 + i) : 0.f;
                asm volatile("bar.sync 1, %0;" ::"n"(NEPI * 32) : "memory");
                last_n0 = n0;
            }
            const bool ok = g_mbar_wait(&tmem_full[buf], (tc >> 1) & 1);
            tc_fence_after();
            if (!ok || abort_s) break;
            const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
            if (ch0 >= ch1) {                                // nothing to drain (one-chunk tiles): hand the accumulator back at once
                tc_fence_before();
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(g_smem_u32(&tmem_empty[buf])) : "memory");
            }
#pragma unroll 1
            for (int c0 = ch0 * CW; c0 < ch1 * CW; c0 += CW, ++ci) {
                unsigned char *bufp = stg + (ci & 1) * 4096;
                if (ci >= 2) {
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    __syncwarp();
                }
                unsigned char *myrow = bufp + lane * 128;
#pragma unroll
                for (int h = 0; h < CW / 32; ++h) {
                    uint32_t r[32];
                    const uint32_t taddr = acc + ((uint32_t)(32 * q) << 16) + (uint32_t)(c0 + 32 * h);
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    float v[32];
                    const float4 *b4 = reinterpret_cast<const float4 *>(bias_s + c0 + 32 * h);   // 8 broadcast LDS.128, not 32 LDS.32
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 bb = b4[j4];
                        const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                        for (int t2 = 0; t2 < 4; ++t2) {
                            float x = __uint_as_float(r[4 * j4 + t2]) + bv[t2];
                            if (act == 1) x = fmaxf(x, 0.f);
                            v[4 * j4 + t2] = x;
                        }
                    }
                    if constexpr (sizeof(OT) == 4) {
#pragma unroll
                        for (int j = 0; j < 8; ++j)              // 16-byte chunk j of the row lands at chunk j ^ (row & 7)
                            *reinterpret_cast<float4 *>(myrow + ((((uint32_t)j) ^ sw) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            uint32_t w[4];
#pragma unroll
                            for (int t2 = 0; t2 < 4; ++t2) {
                                const __nv_bfloat162 hh = __floats2bfloat162_rn(v[8 * j + 2 * t2], v[8 * j + 2 * t2 + 1]);
                                w[t2] = *reinterpret_cast<const uint32_t *>(&hh);
                            }
                            *reinterpret_cast<uint4 *>(myrow + ((((uint32_t)(4 * h + j)) ^ sw) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                        }
                    }
                }
                if (c0 + CW >= ch1 * CW) {                   // this warp's last TMEM read of the tile: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(g_smem_u32(&tmem_empty[buf])) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0 && n0 + c0 < N) {
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                                 ::"l"(&map_c), "r"(g_smem_u32(bufp)), "r"(n0 + c0), "r"(m0 + 32 * q) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                } else if (lane == 0) {
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");      // keep the group count in step with `ci`
                }
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // smem must outlive the stores that read it
    } else if (SPLIT) {
        // ===== operand split (SPLIT kernels have 4 epilogue warps: warps 6..9 = 128 threads): A -> A_hi (in place), A_lo =====
        const int tdx = threadIdx.x - G_THREADS;
        int it = 0;
        bool ok = true;
        for (int t = blockIdx.x; t < ntiles && ok; t += gridDim.x) {
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % NST;
                if (!g_mbar_wait(&full[s], (it / NST) & 1)) { abort_s = 1; ok = false; break; }
                uint4 *pa = reinterpret_cast<uint4 *>(gsm + (size_t)s * STAGE), *pl = pa + A_BYTES / 16;
#pragma unroll
                for (int i = 0; i < A_BYTES / 16 / 128; ++i) {
                    const uint4 v = pa[i * 128 + tdx];
                    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        hi[j] = w[j] & 0xffffe000u;                                       // the 19 bits kind::tf32 reads
                        lo[j] = __float_as_uint(__fsub_rn(__uint_as_float(w[j]), __uint_as_float(hi[j])));   // exact
                    }
                    pa[i * 128 + tdx] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    pl[i * 128 + tdx] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");             // generic-proxy writes -> tensor-core reads
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(g_smem_u32(&conv[s])) : "memory");
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(2 * BN) : "memory");
}

// ------------------------------------------------------------------------------------------------
// W-resident variant for bf16 operands, BN = 256 and K <= 256 (the MSDeformAttn projections and FFN1: K = d_model).
// ncu on the kernel above: after the epilogue was widened, 35 % of all warp samples are epilogue warps waiting for
// tmem_full — the MMAs wait for their operands, and two thirds of the operand bytes are the W tile (256 x 256 bf16 =
// 128 KB) that every output tile fetches again from L2: 148 SMs x 192 KB per 3.2 us = 8.9 TB/s, ~70 % of what the L2
// delivers chip-wide.  Here a CTA keeps ONE W tile in shared memory for its whole life (CTA c owns the n-tile c mod
// tiles_n and walks m-tiles), so a tile costs 64 KB of A from L2 instead of 192 KB.
//   shared memory: W 4 x 32 KB | A ring 4 x 16 KB | staging 8 warps x 4 KB (single buffers: the ring took their twins)
// Roles as above: warp 0 producer, warp 1 MMA issuer, warps 2-9 epilogue.
template <typename OT>
__global__ void __launch_bounds__(G_THREADS8)
gemm_wres_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_c, const float *__restrict__ bias, int M, int N, int K, int act) {
    constexpr int BN = 256, NSTA = 4, KELEMS = 64;
    constexpr int A_BYTES = G_BM * G_KBYTES, WKB_BYTES = BN * G_KBYTES;      // 16 KB, 32 KB per k-block
    extern __shared__ __align__(1024) unsigned char gsm_raw[];
    unsigned char *gsm = gsm_raw + ((1024u - (g_smem_u32(gsm_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full[NSTA], empty[NSTA], w_full, tmem_full[2], tmem_empty[2];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float bias_s[BN];
    __shared__ int abort_s;
    unsigned char *w_res = gsm, *a_ring = gsm + 4 * WKB_BYTES, *stg_base = a_ring + NSTA * A_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = (K + KELEMS - 1) / KELEMS;                               // <= 4
    const int tiles_n = (N + BN - 1) / BN, tiles_m = (M + G_BM - 1) / G_BM;
    const int nt = blockIdx.x % tiles_n, g = blockIdx.x / tiles_n, G = gridDim.x / tiles_n;   // host: gridDim.x % tiles_n == 0
    const int n0 = nt * BN;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTA; ++s) { g_mbar_init(&full[s], 1); g_mbar_init(&empty[s], 1); }
        g_mbar_init(&w_full, 1);
        for (int s = 0; s < 2; ++s) { g_mbar_init(&tmem_full[s], 1); g_mbar_init(&tmem_empty[s], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        abort_s = 0;
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(g_smem_u32(&tmem_base_s)), "n"(2 * BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer: the W tile once, then the A tiles of this CTA's m-tiles =====
            g_mbar_expect_tx(&w_full, (uint32_t)nkb * WKB_BYTES);
            for (int kb = 0; kb < nkb; ++kb) g_tma_load_2d(w_res + (size_t)kb * WKB_BYTES, &map_w, &w_full, kb * KELEMS, n0);
            int it = 0;
            for (int mt = g; mt < tiles_m && !abort_s; mt += G) {
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % NSTA;
                    if (it >= NSTA && !g_mbar_wait(&empty[s], ((it / NSTA) - 1) & 1)) { abort_s = 1; break; }
                    g_mbar_expect_tx(&full[s], A_BYTES);
                    g_tma_load_2d(a_ring + (size_t)s * A_BYTES, &map_a, &full[s], kb * KELEMS, mt * G_BM);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(G_BM >> 4) << 24);
            int it = 0, tc = 0;
            bool ok = g_mbar_wait(&w_full, 0);
            if (!ok) abort_s = 1;
            tc_fence_after();
            for (int mt = g; mt < tiles_m && ok; mt += G, ++tc) {
                const int buf = tc & 1;
                if (tc >= 2 && !g_mbar_wait(&tmem_empty[buf], ((tc >> 1) - 1) & 1)) { abort_s = 1; break; }
                tc_fence_after();
                const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % NSTA;
                    ok = g_mbar_wait(&full[s], (it / NSTA) & 1);
                    if (!ok) { abort_s = 1; break; }
                    tc_fence_after();
                    const uint64_t ad = umma_desc_k128(g_smem_u32(a_ring + (size_t)s * A_BYTES));
                    const uint64_t wd = umma_desc_k128(g_smem_u32(w_res + (size_t)kb * WKB_BYTES));
#pragma unroll
                    for (int k = 0; k < G_KBYTES / 32; ++k) tc_mma(acc, ad + 2 * k, wd + 2 * k, idesc, (kb | k) != 0, 0);
                    tc_commit(&empty[s]);
                }
                tc_commit(&tmem_full[buf]);
            }
        }
    } else {
        // ===== epilogue: warps 2..9, two per TMEM lane quarter, half of the columns each =====
        const int q = warp & 3, part = (warp - 2) >> 2;
        constexpr int CW = 128 / (int)sizeof(OT), NCH = BN / CW;
        const int ch0 = part == 0 ? 0 : NCH / 2, ch1 = part == 0 ? NCH / 2 : NCH;
        unsigned char *bufp = stg_base + (size_t)(warp - 2) * 4096;
        const uint32_t sw = (uint32_t)(lane & 7);
        for (int i = threadIdx.x - 64; i < BN; i += 256) bias_s[i] = (bias && n0 + i < N) ? __ldg(bias + n0 + i) : 0.f;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        unsigned char *myrow = bufp + lane * 128;
        int tc = 0, ci = 0;
        for (int mt = g; mt < tiles_m; mt += G, ++tc) {
            const int buf = tc & 1, m0 = mt * G_BM;
            const bool ok = g_mbar_wait(&tmem_full[buf], (tc >> 1) & 1);
            tc_fence_after();
            if (!ok || abort_s) break;
            const uint32_t acc = tmem_d + (uint32_t)(buf * BN) + ((uint32_t)(32 * q) << 16);
#pragma unroll 1
            for (int c0 = ch0 * CW; c0 < ch1 * CW; c0 += CW, ++ci) {
                uint32_t outw[32];                                   // the chunk's 128 bytes of this lane's row
#pragma unroll
                for (int h = 0; h < CW / 32; ++h) {
                    uint32_t r[32];
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                        : "r"(acc + (uint32_t)(c0 + 32 * h)));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const float4 *b4 = reinterpret_cast<const float4 *>(bias_s + c0 + 32 * h);   // 8 broadcast LDS.128, not 32 LDS.32
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 bb = b4[j4];
                        const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                        for (int t2 = 0; t2 < 4; ++t2) {
                            float x = __uint_as_float(r[4 * j4 + t2]) + bv[t2];
                            if (act == 1) x = fmaxf(x, 0.f);
                            r[4 * j4 + t2] = __float_as_uint(x);
                        }
                    }
                    if constexpr (sizeof(OT) == 4) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) outw[j] = r[j];
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
                            outw[16 * h + j] = *reinterpret_cast<const uint32_t *>(&hh);
                        }
                    }
                }
                if (c0 + CW >= ch1 * CW) {                           // last TMEM read of the tile by this warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(g_smem_u32(&tmem_empty[buf])) : "memory");
                }
                if (ci >= 1) {                                       // single staging buffer: the previous store has read it
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    __syncwarp();
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)                          // 16-byte chunk j of the row lands at chunk j ^ (row & 7)
                    *reinterpret_cast<uint4 *>(myrow + ((((uint32_t)j) ^ sw) << 4)) =
                        make_uint4(outw[4 * j], outw[4 * j + 1], outw[4 * j + 2], outw[4 * j + 3]);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    if (n0 + c0 < N)
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(&map_c), "r"(g_smem_u32(bufp)), "r"(n0 + c0), "r"(m0 + 32 * q) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(2 * BN) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn gemm_encode_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}
// row-major [rows, K] matrix with row pitch `ld` elements, viewed through (128 bytes of K) x box_rows boxes, 128B swizzle
static bool gemm_map(CUtensorMap *m, const void *base, int rows, int K, int ld, int box_rows, int es) {
    EncodeTiledFn enc = gemm_encode_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * es};
    const cuuint32_t box[2] = {(cuuint32_t)(G_KBYTES / es), (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, es == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides,
               box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int BN, bool TF32, typename OT, bool SPLIT = false>
static int gemm_launch(const CUtensorMap &ma, const CUtensorMap &mw, const CUtensorMap &mwlo, const CUtensorMap &mc, const float *bias,
                       int M, int N, int K, int act, cudaStream_t st, const CUtensorMap *ma2 = nullptr, int kb_split = 0) {
    constexpr int NST = SPLIT ? 2 : (BN == 256 ? 3 : 4);
    const size_t smem = (size_t)NST * (SPLIT ? 2 : 1) * (G_BM + BN) * G_KBYTES + (SPLIT ? 4 : 8) * 2 * 4096 + 1024;
    QUBER_CUDA_TRY(cudaFuncSetAttribute(gemm_bias_act_kernel<BN, TF32, OT, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (long long)cdiv(M, G_BM) * cdiv(N, BN);
    const int grid = (int)(ntiles < sm_count() ? ntiles : sm_count());       // persistent: one CTA per SM
    gemm_bias_act_kernel<BN, TF32, OT, SPLIT><<<grid, SPLIT ? G_THREADS + 128 : G_THREADS8, smem, st>>>(ma, mw, mwlo, mc, bias, M, N, K, act,
                                                                                                     ma2 ? *ma2 : ma, kb_split);
    return check_launch("gemm_bias_act_kernel");
}

template <typename OT>
static int gemm_wres_launch(const CUtensorMap &ma, const CUtensorMap &mw, const CUtensorMap &mc, const float *bias, int M, int N, int K,
                            int act, cudaStream_t st) {
    const size_t smem = 4 * (size_t)256 * G_KBYTES + 4 * (size_t)G_BM * G_KBYTES + 8 * 4096 + 1024;
    QUBER_CUDA_TRY(cudaFuncSetAttribute(gemm_wres_kernel<OT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int tiles_n = cdiv(N, 256), tiles_m = cdiv(M, G_BM);
    int per = sm_count() / tiles_n;                               // CTAs per n-tile
    if (per > tiles_m) per = tiles_m;
    gemm_wres_kernel<OT><<<per * tiles_n, G_THREADS8, smem, st>>>(ma, mw, mc, bias, M, N, K, act);
    return check_launch("gemm_wres_kernel");
}

// w -> hi = w with the 13 low mantissa bits cleared (what kind::tf32 reads), lo = w - hi (exact)
__global__ void __launch_bounds__(256)
split_tf32_kernel(const float *__restrict__ w, size_t n, float *__restrict__ hi, float *__restrict__ lo) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
        const float x = w[i];
        const float h = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
        hi[i] = h;
        lo[i] = __fsub_rn(x, h);
    }
}

}  // namespace quber

extern "C" int quber_gemm_timeouts(void) {
    unsigned int n = 0;
    if (cudaMemcpyFromSymbol(&n, quber::g_gemm_timeouts, sizeof(n)) != cudaSuccess) return -1;
    return (int)n;
}

extern "C" int quber_gemm_bias_act(const void *A, const void *W, const float *bias, void *C, int M, int N, int K, int lda, int ldw,
                                   int ldc, int in_dtype, int out_dtype, int act, void *stream) {
    using namespace quber;
    if (!A || !W || !C || M <= 0 || N <= 0 || K <= 0) return QUBER_E_INVALID;
    if ((in_dtype != 1 && in_dtype != 2) || (out_dtype != 0 && out_dtype != 1) || (act != 0 && act != 1)) return QUBER_E_INVALID;
    const bool tf32 = in_dtype == 2;
    const int es = tf32 ? 4 : 2, os = out_dtype == 0 ? 4 : 2;
    if (lda < K || ldw < K || ldc < N) return QUBER_E_INVALID;
    if (!aligned16(A) || !aligned16(W) || !aligned16(C) || ((size_t)lda * es) % 16 || ((size_t)ldw * es) % 16 || ((size_t)ldc * os) % 16)
        return QUBER_E_ALIGN;
    const int BN = (N % 256 == 0) ? 256 : ((N % 128 == 0 || N > 192) ? 128 : 64);    // 256: the whole row panel, A read once
    CUtensorMap ma, mw, mc;                 // C: 128-byte x 32-row boxes, one per epilogue warp and chunk
    if (!gemm_map(&ma, A, M, K, lda, G_BM, es) || !gemm_map(&mw, W, N, K, ldw, BN, es) || !gemm_map(&mc, C, M, N, ldc, 32, os)) {
        set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled");
        return QUBER_E_CUDA;
    }
    cudaStream_t st = (cudaStream_t)stream;
    // W tile resident in shared memory: measured for bf16 output (FFN1 138 -> 121 us, value projection 46 -> 42 us); with
    // fp32 output its single staging buffers cost what the saved L2 traffic returns (59 -> 62 us), so those stay above
    if (!tf32 && out_dtype == 1 && BN == 256 && K <= 256 && cdiv(N, 256) <= sm_count())
        return gemm_wres_launch<__nv_bfloat16>(ma, mw, mc, bias, M, N, K, act, st);
#define QG(BN_, TF_, OT_) return gemm_launch<BN_, TF_, OT_>(ma, mw, mw, mc, bias, M, N, K, act, st)
    if (BN == 256) {
        if (tf32) { if (out_dtype == 0) QG(256, true, float); else QG(256, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG(256, false, float); else QG(256, false, __nv_bfloat16); }
    } else if (BN == 128) {
        if (tf32) { if (out_dtype == 0) QG(128, true, float); else QG(128, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG(128, false, float); else QG(128, false, __nv_bfloat16); }
    } else {
        if (tf32) { if (out_dtype == 0) QG(64, true, float); else QG(64, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG(64, false, float); else QG(64, false, __nv_bfloat16); }
    }
#undef QG
}

// C = act([A1 | A2] W^T + bias): the 1x1 convolution that follows a channel concatenation (RGB-D "concat" fusion,
// maskrefiner/modeling/backbone/resnet.py:466-479,497-501: torch.cat([rgb, depth], 1) -> Conv2d(2c, c, 1)) on channels-last
// rows, without materialising the concatenation: the TMA producer takes the k-blocks below K1 from A1 and the rest from A2.
extern "C" int quber_gemm_bias_act_cat2(const void *A1, const void *A2, const void *W, const float *bias, void *C, int M, int N,
                                        int K1, int K2, int lda1, int lda2, int ldw, int ldc, int in_dtype, int out_dtype, int act,
                                        void *stream) {
    using namespace quber;
    if (!A1 || !A2 || !W || !C || M <= 0 || N <= 0 || K1 <= 0 || K2 <= 0) return QUBER_E_INVALID;
    if ((in_dtype != 1 && in_dtype != 2) || (out_dtype != 0 && out_dtype != 1) || (act != 0 && act != 1)) return QUBER_E_INVALID;
    const bool tf32 = in_dtype == 2;
    const int es = tf32 ? 4 : 2, os = out_dtype == 0 ? 4 : 2, kelems = tf32 ? 32 : 64, K = K1 + K2;
    if (lda1 < K1 || lda2 < K2 || ldw < K || ldc < N) return QUBER_E_INVALID;
    if (K1 % kelems != 0) return QUBER_E_UNSUPPORTED;                 // the split must fall on a k-block boundary
    if (!aligned16(A1) || !aligned16(A2) || !aligned16(W) || !aligned16(C) || ((size_t)lda1 * es) % 16 || ((size_t)lda2 * es) % 16 ||
        ((size_t)ldw * es) % 16 || ((size_t)ldc * os) % 16)
        return QUBER_E_ALIGN;
    const int BN = (N % 256 == 0) ? 256 : ((N % 128 == 0 || N > 192) ? 128 : 64);
    CUtensorMap ma, ma2, mw, mc;
    if (!gemm_map(&ma, A1, M, K1, lda1, G_BM, es) || !gemm_map(&ma2, A2, M, K2, lda2, G_BM, es) || !gemm_map(&mw, W, N, K, ldw, BN, es) ||
        !gemm_map(&mc, C, M, N, ldc, 32, os)) {
        set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled");
        return QUBER_E_CUDA;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int ks = K1 / kelems;
#define QG2(BN_, TF_, OT_) return gemm_launch<BN_, TF_, OT_>(ma, mw, mw, mc, bias, M, N, K, act, st, &ma2, ks)
    if (BN == 256) {
        if (tf32) { if (out_dtype == 0) QG2(256, true, float); else QG2(256, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG2(256, false, float); else QG2(256, false, __nv_bfloat16); }
    } else if (BN == 128) {
        if (tf32) { if (out_dtype == 0) QG2(128, true, float); else QG2(128, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG2(128, false, float); else QG2(128, false, __nv_bfloat16); }
    } else {
        if (tf32) { if (out_dtype == 0) QG2(64, true, float); else QG2(64, true, __nv_bfloat16); }
        else      { if (out_dtype == 0) QG2(64, false, float); else QG2(64, false, __nv_bfloat16); }
    }
#undef QG2
}

extern "C" int quber_split_tf32(const float *w, size_t n, float *hi, float *lo, void *stream) {
    using namespace quber;
    if (!w || !hi || !lo || n == 0) return QUBER_E_INVALID;
    const unsigned blocks = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
    split_tf32_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w, n, hi, lo);
    return check_launch("split_tf32_kernel");
}

extern "C" int quber_gemm_bias_act_x3(const float *A, const float *W_hi, const float *W_lo, const float *bias, void *C, int M, int N,
                                      int K, int lda, int ldw, int ldc, int out_dtype, int act, void *stream) {
    using namespace quber;
    if (!A || !W_hi || !W_lo || !C || M <= 0 || N <= 0 || K <= 0) return QUBER_E_INVALID;
    if ((out_dtype != 0 && out_dtype != 1) || (act != 0 && act != 1)) return QUBER_E_INVALID;
    const int os = out_dtype == 0 ? 4 : 2;
    if (lda < K || ldw < K || ldc < N) return QUBER_E_INVALID;
    if (!aligned16(A) || !aligned16(W_hi) || !aligned16(W_lo) || !aligned16(C) || ((size_t)lda * 4) % 16 || ((size_t)ldw * 4) % 16 ||
        ((size_t)ldc * os) % 16)
        return QUBER_E_ALIGN;
    const int BN = (N % 256 == 0) ? 256 : ((N % 128 == 0 || N > 192) ? 128 : 64);    // 256: A is read AND split once per row panel
    CUtensorMap ma, mw, mwl, mc;
    if (!gemm_map(&ma, A, M, K, lda, G_BM, 4) || !gemm_map(&mw, W_hi, N, K, ldw, BN, 4) || !gemm_map(&mwl, W_lo, N, K, ldw, BN, 4) ||
        !gemm_map(&mc, C, M, N, ldc, 32, os)) {
        set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled");
        return QUBER_E_CUDA;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (BN == 256) {
        if (out_dtype == 0) return gemm_launch<256, true, float, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
        return gemm_launch<256, true, __nv_bfloat16, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
    }
    if (BN == 128) {
        if (out_dtype == 0) return gemm_launch<128, true, float, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
        return gemm_launch<128, true, __nv_bfloat16, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
    }
    if (out_dtype == 0) return gemm_launch<64, true, float, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
    return gemm_launch<64, true, __nv_bfloat16, true>(ma, mw, mwl, mc, bias, M, N, K, act, st);
}
