// K2: per-instance boundary extraction by erosion + union / foreground maps + quadruple maps.
//
// Reference: explicit_error_estimation/util.py:62-99 (masks_to_fg_mask, mask_to_boundary,
// masks_to_boundary), tools/ours/panoptic2eee.py:110-123 (TP/TN/FP/FN).
//
// `d` iterations of a 3x3 erosion on a 1-px zero-padded image equal ONE zero-padded
// (2d+1)x(2d+1) uint8 min filter; that is what runs here, separably and four pixels per
// instruction (__vminu4 on packed bytes) with window doubling:
//     F_1 = x,  F_2p[i] = min(F_p[i], F_p[i+p]),  win_w[i] = min(F_p[i], F_p[i+w-p])  (p = 2^k <= w)
// first along rows at byte granularity (__byte_perm realigns the shifted words), then along
// columns on whole words.  One CTA owns one (image, tile) and walks all K mask planes, so the
// union (`boundary`) and foreground (`fg`) sums live in registers and are written once.
// A tile (+halo) whose bytes are all equal has no boundary: it is detected on load and skipped.
#include <cuda.h>

#include "common.cuh"

namespace quber {

constexpr int B_TW = 128;           // tile width in pixels (32 packed words)
constexpr int B_TWW = B_TW / 4;     // 32
constexpr int B_TH = 64;            // tile height
constexpr int B_THREADS = 256;
constexpr int B_RPT = B_TH * B_TWW / B_THREADS;  // central words per thread (8)
constexpr int B_MAXK_OCC = 512;     // the occupancy short-cut handles up to this many planes per image

__device__ __forceinline__ uint32_t shift_bytes(uint32_t lo, uint32_t hi, int s) {
    // bytes [s, s+4) of the little-endian 8-byte string lo:hi, s in 0..3
    return __byte_perm(lo, hi, 0x3210 + 0x1111 * s);
}

// ------------------------------------------------------------------------------------------------
// The (tile + halo) box of every mask plane is fetched by ONE
// cp.async.bulk.tensor.3d per plane — out-of-image bytes arrive as zeros, which IS the erosion's
// zero padding — two planes ahead of the compute through two mbarriers, and the per-instance
// result leaves through a TMA store (clipped at the image edge by the hardware).  No thread issues
// a global load; the only per-thread global stores are the two union maps at the end.
//
// Measured on B200 (scripts/tma_probe.cu): a tiled load faults (illegal instruction) unless the
// innermost start coordinate is a multiple of 16 BYTES; negative / past-the-edge load coordinates
// are fine (zero fill); a store must not start at a negative coordinate.  Hence the horizontal halo
// is rounded up to 16 pixels (B_HX) and stores always start at the tile origin.
//
// Per plane a CTA takes one of three routes, decided from one pass over the staged bytes:
//   empty   every byte 0                         -> nothing to erode; a zero tile is stored
//   binary  every byte in {0, v} for a single v  -> the tile is packed to ONE BIT per pixel (16 pixels
//           per 128-bit shared load), eroded with funnel shifts and ANDs on 32-pixel words, and expanded
//           back to bytes only for the 128x64 central tile.  This is the route every real mask takes
//           (the reference feeds 0/255 or 0/1 masks).
//   bytes   anything else                        -> the exact uint8 min filter (window doubling on
//           packed bytes, see the header), so arbitrary planes still match cv2 bit for bit.
// The union maps are uint8 WRAPPING sums over the planes (util.py:96-98 sums uint8 arrays), kept as
// packed byte accumulators in registers in every route.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// false = timed out (a lost TMA must fail the call, never hang the GPU)
__device__ __forceinline__ bool mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    for (int spin = 0; !done; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1 << 22)) return false;
    }
    return true;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, uint64_t *bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, const void *src, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(map), "r"(smem_u32(src)), "r"(x), "r"(y), "r"(z) : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// packed-byte helpers that stay cheap in SASS (sm_100 has no SIMD-in-word integer min/add: __vminu4 and
// friends expand to LOP3/IADD sequences, which is why the binary route avoids them)
__device__ __forceinline__ uint32_t add_bytes(uint32_t a, uint32_t b) {          // per-byte wrapping add
    return ((a & 0x7f7f7f7fu) + (b & 0x7f7f7f7fu)) ^ ((a ^ b) & 0x80808080u);
}
__device__ __forceinline__ uint32_t nonzero_bytes01(uint32_t w) {                // 0x01 per non-zero byte
    return ((((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) & 0x80808080u) >> 7;
}
__device__ __forceinline__ uint32_t nibble_to_bytes01(uint32_t n) {              // bit i of n -> 0x01 in byte i
    return (n * 0x00204081u) & 0x01010101u;
}

// staged geometry shared by the kernel and the host launcher
struct BGeom {
    int hx, SW, WS, WS16, R, RW32;
    uint32_t stage_bytes, stage_stride, v1_bytes, bits_bytes, hb_bytes;
    size_t smem;
    __host__ __device__ explicit BGeom(int d) {
        hx = (d + 15) & ~15;                  // horizontal halo in pixels: TMA start coordinate % 16 == 0
        SW = B_TW + 2 * hx;                   // staged row, bytes (multiple of 32)
        WS = SW >> 2;
        WS16 = SW >> 4;
        R = B_TH + 2 * d;
        RW32 = (SW >> 5) + 1;                 // bit-row words + one zero word read by the last funnel shift
        stage_bytes = (uint32_t)R * SW;
        stage_stride = (stage_bytes + 127u) & ~127u;
        v1_bytes = ((uint32_t)R * B_TWW * 4 + 127u) & ~127u;
        bits_bytes = ((uint32_t)R * RW32 * 4 + 127u) & ~127u;
        hb_bytes = ((uint32_t)R * 4 * 4 + 127u) & ~127u;
        smem = 2 * (size_t)stage_stride + v1_bytes + 3 * (size_t)B_TH * B_TWW * 4 + bits_bytes + hb_bytes + 128;
    }
};

__device__ unsigned int g_tma_timeouts = 0;     // diagnostic: mbarrier waits that gave up (quber_tma_timeouts())

__global__ void __launch_bounds__(B_THREADS)
boundary_tma_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map, int K, int H, int W,
                    int d, int has_per, uint8_t *__restrict__ boundary, uint8_t *__restrict__ fg, int32_t *__restrict__ status,
                    const uint8_t *__restrict__ occ, int occ_ty, int occ_tx) {
    extern __shared__ __align__(128) unsigned char sm_raw[];
    __shared__ __align__(8) uint64_t full[2];
    __shared__ uint32_t red[6];                 // 3 rotating (OR, AND) slots of the per-plane byte census
    __shared__ int s_abort;
    __shared__ uint16_t s_live[B_MAXK_OCC];     // planes with a set pixel near this tile, in plane order
    __shared__ int s_nlive;
    const BGeom g(d);
    const int hx = g.hx, WS = g.WS, WS16 = g.WS16, R = g.R, RW32 = g.RW32, hw = hx >> 2;
    auto S = [&](int i) { return reinterpret_cast<uint32_t *>(sm_raw + (size_t)(i & 1) * g.stage_stride); };
    uint32_t *V1 = reinterpret_cast<uint32_t *>(sm_raw + 2 * (size_t)g.stage_stride);       // [R][32]   (byte route)
    uint32_t *O0 = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(V1) + g.v1_bytes);   // [64][32] x2
    uint32_t *Z = O0 + 2 * B_TH * B_TWW;                                                   // zero tile
    uint32_t *bits = Z + B_TH * B_TWW;                                                     // [R][RW32] one bit per staged pixel
    uint32_t *Hb = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(bits) + g.bits_bytes);   // [R][4]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.z;
    const int ty0 = blockIdx.y * B_TH, tx0 = blockIdx.x * B_TW;
    const size_t HW = (size_t)H * W;
    const int w = 2 * d + 1;
    int p_top = 1;
    while (p_top * 2 <= w) p_top *= 2;
    const int lead = hx - d;

    for (int i = tid; i < B_TH * B_TWW; i += B_THREADS) Z[i] = 0u;
    for (int i = tid; i < R * RW32; i += B_THREADS) bits[i] = 0u;      // the pad word of every bit row stays 0
    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        red[0] = red[2] = red[4] = 0u;
        red[1] = red[3] = red[5] = 0xffffffffu;
        s_abort = 0;
    }
    // Occupancy short-cut.  `occ` (written by the encode pass that read every mask byte anyway, csrc/encode.cu) says for
    // every plane and every 128 x 16 pixel tile whether the tile holds a set pixel.  A plane with no set pixel within d of
    // this tile has neither boundary nor foreground here: it is never loaded, its per-instance tile is the zero tile.
    // Conservative (tile granularity), so the census below still sees some empty boxes.  Warp 0 builds the list of live
    // planes in plane order.
    const bool use_occ = occ != nullptr && K <= B_MAXK_OCC;
    if (warp == 0) {
        int n = 0;
        if (use_occ) {
            const int r0 = max(ty0 - d, 0) / QUBER_OCC_TILE_H, r1 = min((ty0 + B_TH - 1 + d) / QUBER_OCC_TILE_H, occ_ty - 1);
            const int c0 = max(tx0 - d, 0) / QUBER_OCC_TILE_W, c1 = min((tx0 + B_TW - 1 + d) / QUBER_OCC_TILE_W, occ_tx - 1);
            for (int k0 = 0; k0 < K; k0 += 32) {
                const int k = k0 + lane;
                int any = 0;
                if (k < K) {
                    const uint8_t *o = occ + ((size_t)b * K + k) * occ_ty * occ_tx;
                    for (int r = r0; r <= r1; ++r)
                        for (int c = c0; c <= c1; ++c) any |= __ldg(o + r * occ_tx + c);
                }
                const uint32_t m = __ballot_sync(0xffffffffu, any != 0);
                if (any) s_live[n + __popc(m & ((1u << lane) - 1u))] = (uint16_t)k;
                n += __popc(m);
            }
        }
        if (lane == 0) s_nlive = use_occ ? n : K;
    }
    fence_proxy_async();         // Z was written by the generic proxy, TMA stores read it
    __syncthreads();
    const int n_live = s_nlive;
    auto plane_of = [&](int j) { return use_occ ? (int)s_live[j] : j; };
    auto issue = [&](int j) {    // thread 0 only: fetch the j-th live plane into stage j & 1
        fence_proxy_async();
        mbar_expect_tx(&full[j & 1], g.stage_bytes);
        tma_load_3d(S(j), &in_map, &full[j & 1], tx0 - hx, ty0 - d, b * K + plane_of(j));
    };
    if (tid == 0) {
        if (n_live > 0) issue(0);
        if (n_live > 1) issue(1);
    }
    if (use_occ && has_per && tid == 32) {       // the zero tiles of the planes that are not live: one thread, its own bulk group
        int j = 0;
        for (int k = 0; k < K; ++k) {
            if (j < n_live && (int)s_live[j] == k) { ++j; continue; }
            tma_store_3d(&out_map, Z, tx0, ty0, b * K + k);
        }
        tma_commit();
    }

    const int row = tid >> 2, jw = tid & 3;     // this thread's 32 output pixels: tile row, 32-pixel column group
    uint32_t uni_acc[8], fg_acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) uni_acc[q] = fg_acc[q] = 0;
    int o_slot = 0;

    for (int k = 0; k < n_live; ++k) {          // k counts LIVE planes from here on; the plane itself is plane_of(k)
        uint32_t *cur = S(k);
        if (!mbar_wait(&full[k & 1], (k >> 1) & 1)) {            // a lost TMA: this CTA's outputs are invalid — say so per call
            s_abort = 1;
            if (tid == 0) { atomicAdd(&g_tma_timeouts, 1u); if (status) atomicOr(status, 1); }
        }
        // ---- pass A: is the staged tile (+ halo) all zero?  ~87 % of the (tile, plane) pairs at C2 are, and they
        //      leave after this OR (8 instructions per 16 pixels) without the census / bit packing below ----
        const int slot = (k % 3) * 2;
        const uint4 *c4 = reinterpret_cast<const uint4 *>(cur);
        const int n16 = R * WS16;
        {
            uint32_t o = 0u;
            for (int idx = tid; idx < n16; idx += B_THREADS) {
                const uint4 x = c4[idx];
                o |= x.x | x.y | x.z | x.w;
            }
            o = __reduce_or_sync(0xffffffffu, o);
            if (lane == 0 && o) atomicOr(&red[slot], o);
        }
        __syncthreads();                                    // S1a: OR census complete
        if (s_abort) break;
        const uint32_t oc = red[slot];
        const uint32_t v = (oc | (oc >> 8) | (oc >> 16) | (oc >> 24)) & 0xffu;          // OR of all bytes
        const bool empty = v == 0u;
        if (tid == 0) {
            const int s2 = ((k + 2) % 3) * 2;               // next written two iterations from now, last read one ago
            red[s2] = 0u;
            red[s2 + 1] = 0xffffffffu;
            if (empty && k + 2 < n_live) issue(k + 2);      // every read of the stage is done
        }
        if (empty) {
            if (has_per && tid == 0) { tma_store_3d(&out_map, Z, tx0, ty0, b * K + plane_of(k)); tma_commit(); }
            continue;
        }
        // ---- pass B: AND census of the non-zero bytes (binary plane?) + bit packing, 16 pixels per shared load ----
        {
            uint32_t a = 0xffffffffu;
            uint16_t *b16 = reinterpret_cast<uint16_t *>(bits);
            for (int idx = tid; idx < n16; idx += B_THREADS) {
                const int r = idx / WS16, c = idx - r * WS16;
                const uint4 x = c4[idx];
                const uint32_t t0 = nonzero_bytes01(x.x), t1 = nonzero_bytes01(x.y), t2 = nonzero_bytes01(x.z), t3 = nonzero_bytes01(x.w);
                a &= (x.x | ~(t0 * 255u)) & (x.y | ~(t1 * 255u)) & (x.z | ~(t2 * 255u)) & (x.w | ~(t3 * 255u));
                const uint32_t pk = ((t0 * 0x01020408u) >> 24) | (((t1 * 0x01020408u) >> 24) << 4) |
                                    (((t2 * 0x01020408u) >> 24) << 8) | (((t3 * 0x01020408u) >> 24) << 12);
                b16[r * (2 * RW32) + c] = (uint16_t)pk;
            }
            a = __reduce_and_sync(0xffffffffu, a);
            if (lane == 0 && a != 0xffffffffu) atomicAnd(&red[slot + 1], a);
        }
        __syncthreads();                                    // S1b: census + bit rows complete, `cur` is dead on the binary route
        const uint32_t ac = red[slot + 1];
        const uint32_t va = (ac & (ac >> 8) & (ac >> 16) & (ac >> 24)) & 0xffu;         // AND of the non-zero bytes
        const bool binary = v == va;
        if (tid == 0 && binary && k + 2 < n_live) issue(k + 2);
        uint32_t bdw[8], mw[8];
        if (binary) {
            // ---- horizontal erosion on bit rows: window [x-d, x+d] for the four central 32-pixel words ----
            for (int it = tid; it < R * 4; it += B_THREADS) {
                const uint32_t *row32 = bits + (it >> 2) * RW32;
                const int s0 = hx + 32 * (it & 3) - d;
                uint32_t e = 0xffffffffu;
                for (int t = 0; t < w; ++t) {
                    const int s = s0 + t;
                    e &= __funnelshift_r(row32[s >> 5], row32[(s >> 5) + 1], s & 31);
                }
                Hb[it] = e;
            }
            if (has_per && tid == 0) tma_wait_read<1>();     // the store that last used this O slot has read it
            __syncthreads();                                 // S2
            uint32_t e = 0xffffffffu;
            for (int t = 0; t < w; ++t) e &= Hb[(row + t) * 4 + jw];
            const uint32_t *mrow = bits + (row + d) * RW32 + (hx >> 5) + jw;
            const uint32_t m = __funnelshift_r(mrow[0], mrow[1], hx & 31);
            const uint32_t bd = m & ~e;                      // mask - eroded, one bit per pixel
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                bdw[q] = nibble_to_bytes01((bd >> (4 * q)) & 15u) * v;
                mw[q] = nibble_to_bytes01((m >> (4 * q)) & 15u) * v;
            }
        } else {
            // ---- exact uint8 min filter (window doubling on packed bytes) ----
            {
                const uint4 *src4 = reinterpret_cast<const uint4 *>(cur + (size_t)(row + d) * WS + hw + 8 * jw);
                const uint4 a0 = src4[0], a1 = src4[1];
                mw[0] = a0.x; mw[1] = a0.y; mw[2] = a0.z; mw[3] = a0.w; mw[4] = a1.x; mw[5] = a1.y; mw[6] = a1.z; mw[7] = a1.w;
            }
            __syncthreads();                                 // every thread holds its originals before rows are overwritten
            for (int r = warp; r < R; r += B_THREADS / 32) {  // horizontal pass, in place, one warp per row
                uint32_t *rowp = cur + (size_t)r * WS;
                const int i0 = lane, i1 = lane + 32;
                for (int p = 1; p < p_top; p <<= 1) {
                    const int sw = p >> 2, sb = p & 3;
                    uint32_t n0, n1 = 0;
                    {
                        const int j0 = i0 + sw;
                        const uint32_t lo = j0 < WS ? rowp[j0] : 0u, hi = j0 + 1 < WS ? rowp[j0 + 1] : 0u;
                        n0 = __vminu4(rowp[i0], sb ? shift_bytes(lo, hi, sb) : lo);
                    }
                    if (i1 < WS) {
                        const int j0 = i1 + sw;
                        const uint32_t lo = j0 < WS ? rowp[j0] : 0u, hi = j0 + 1 < WS ? rowp[j0 + 1] : 0u;
                        n1 = __vminu4(rowp[i1], sb ? shift_bytes(lo, hi, sb) : lo);
                    }
                    __syncwarp();
                    rowp[i0] = n0;
                    if (i1 < WS) rowp[i1] = n1;
                    __syncwarp();
                }
                const int a1 = 4 * lane + lead, a2 = a1 + w - p_top;
                const int w1 = a1 >> 2, s1 = a1 & 3, w2 = a2 >> 2, s2 = a2 & 3;
                const uint32_t v1 = shift_bytes(rowp[w1], w1 + 1 < WS ? rowp[w1 + 1] : 0u, s1);
                const uint32_t v2 = shift_bytes(rowp[w2], w2 + 1 < WS ? rowp[w2 + 1] : 0u, s2);
                V1[r * B_TWW + lane] = __vminu4(v1, v2);
            }
            if (has_per && tid == 0) tma_wait_read<1>();
            __syncthreads();
            uint32_t *src = V1, *dst = cur;                  // vertical pass: ping-pong V1 <-> (dead) staging buffer
            for (int p = 1; p < p_top; p <<= 1) {
                for (int idx = tid; idx < R * B_TWW; idx += B_THREADS) {
                    const int r = idx >> 5;
                    const uint32_t other = (r + p < R) ? src[idx + p * B_TWW] : 0u;
                    dst[idx] = __vminu4(src[idx], other);
                }
                __syncthreads();
                uint32_t *t = src; src = dst; dst = t;
            }
            const uint32_t *e0 = src + row * B_TWW + 8 * jw, *e1 = src + (row + w - p_top) * B_TWW + 8 * jw;
#pragma unroll
            for (int q = 0; q < 8; ++q) bdw[q] = __vsub4(mw[q], __vminu4(e0[q], e1[q]));
        }
        // ---- common tail: wrapping uint8 sums for the union maps, per-instance tile out through TMA ----
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            uni_acc[q] = add_bytes(uni_acc[q], bdw[q]);
            fg_acc[q] = add_bytes(fg_acc[q], mw[q]);
        }
        if (has_per) {
            uint4 *o4 = reinterpret_cast<uint4 *>(O0 + (size_t)o_slot * B_TH * B_TWW + row * B_TWW + 8 * jw);
            o4[0] = make_uint4(bdw[0], bdw[1], bdw[2], bdw[3]);
            o4[1] = make_uint4(bdw[4], bdw[5], bdw[6], bdw[7]);
            fence_proxy_async();
        }
        __syncthreads();                                     // S3: O tile complete; bits / Hb / V1 / cur are free again
        if (tid == 0) {
            if (has_per) {
                tma_store_3d(&out_map, O0 + (size_t)o_slot * B_TH * B_TWW, tx0, ty0, b * K + plane_of(k));
                tma_commit();
            }
            if (!binary && k + 2 < n_live) issue(k + 2);
        }
        o_slot ^= 1;
    }
    if (!s_abort) {
        const int gy = ty0 + row, gx = tx0 + 32 * jw;
        if (gy < H && gx < W) {                              // W % 16 == 0: each 16-pixel half is all in or all out
            const size_t o = (size_t)b * HW + (size_t)gy * W + gx;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (gx + 16 * h >= W) break;
                if (boundary)
                    st_stream_u4(boundary + o + 16 * h, make_uint4(nonzero_bytes01(uni_acc[4 * h]), nonzero_bytes01(uni_acc[4 * h + 1]),
                                                                   nonzero_bytes01(uni_acc[4 * h + 2]), nonzero_bytes01(uni_acc[4 * h + 3])));
                if (fg)
                    st_stream_u4(fg + o + 16 * h, make_uint4(nonzero_bytes01(fg_acc[4 * h]), nonzero_bytes01(fg_acc[4 * h + 1]),
                                                             nonzero_bytes01(fg_acc[4 * h + 2]), nonzero_bytes01(fg_acc[4 * h + 3])));
            }
        }
    }
    if (tid == 0 || tid == 32) tma_wait_all();     // shared memory must outlive the bulk stores that read it
}

// ------------------------------------------------------------------------------------------------
// The same per-instance boundaries from BIT planes (one bit per pixel: quber_host_pack_masks / quber_rle_decode — what the
// host-buffer routes hold anyway).  No byte planes are ever materialised: a plane's (tile + halo) box is R x 5..6 words
// fetched with plain loads one live plane ahead, so there is no byte census, no packing pass and no TMA load (hence
// nothing to time out); the erosion on bit rows and the byte expansion of the 128 x 64 result are those of the binary
// route above.  Set pixels carry `v` in the outputs (the byte value the planes were packed from).  W % 32 == 0.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(B_THREADS)
boundary_bits_kernel(const uint32_t *__restrict__ planes, uint32_t v, const __grid_constant__ CUtensorMap out_map, int K, int H, int W,
                     int d, int has_per, uint8_t *__restrict__ boundary, uint8_t *__restrict__ fg,
                     const uint8_t *__restrict__ occ, int occ_ty, int occ_tx) {
    extern __shared__ __align__(128) unsigned char sm_raw[];
    __shared__ uint16_t s_live[B_MAXK_OCC];
    __shared__ int s_nlive;
    const BGeom g(d);
    const int hx = g.hx, R = g.R, RW32 = g.RW32, BW = g.SW >> 5;     // BW: box words per row (RW32 = BW + the zero pad word)
    uint32_t *O0 = reinterpret_cast<uint32_t *>(sm_raw);                                    // [2][64][32]
    uint32_t *Z = O0 + 2 * B_TH * B_TWW;                                                   // zero tile
    uint32_t *bits = Z + B_TH * B_TWW;                                                     // [R][RW32]
    uint32_t *Hb = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(bits) + g.bits_bytes);   // [R][4]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.z;
    const int ty0 = blockIdx.y * B_TH, tx0 = blockIdx.x * B_TW;
    const size_t HW = (size_t)H * W;
    const int WB = W >> 5;
    const int w = 2 * d + 1;
    for (int i = tid; i < B_TH * B_TWW; i += B_THREADS) Z[i] = 0u;
    for (int i = tid; i < R * RW32; i += B_THREADS) bits[i] = 0u;      // the pad word of every bit row stays 0
    const bool use_occ = occ != nullptr && K <= B_MAXK_OCC;
    if (warp == 0) {
        int n = 0;
        if (use_occ) {
            const int r0 = max(ty0 - d, 0) / QUBER_OCC_TILE_H, r1 = min((ty0 + B_TH - 1 + d) / QUBER_OCC_TILE_H, occ_ty - 1);
            const int c0 = max(tx0 - d, 0) / QUBER_OCC_TILE_W, c1 = min((tx0 + B_TW - 1 + d) / QUBER_OCC_TILE_W, occ_tx - 1);
            for (int k0 = 0; k0 < K; k0 += 32) {
                const int k = k0 + lane;
                int any = 0;
                if (k < K) {
                    const uint8_t *o = occ + ((size_t)b * K + k) * occ_ty * occ_tx;
                    for (int r = r0; r <= r1; ++r)
                        for (int c = c0; c <= c1; ++c) any |= __ldg(o + r * occ_tx + c);
                }
                const uint32_t m = __ballot_sync(0xffffffffu, any != 0);
                if (any) s_live[n + __popc(m & ((1u << lane) - 1u))] = (uint16_t)k;
                n += __popc(m);
            }
        }
        if (lane == 0) s_nlive = use_occ ? n : K;
    }
    fence_proxy_async();         // Z was written by the generic proxy, TMA stores read it
    __syncthreads();
    const int n_live = s_nlive;
    auto plane_of = [&](int j) { return use_occ ? (int)s_live[j] : j; };
    if (use_occ && has_per && tid == 32) {       // zero tiles of the planes that are not live
        int j = 0;
        for (int k = 0; k < K; ++k) {
            if (j < n_live && (int)s_live[j] == k) { ++j; continue; }
            tma_store_3d(&out_map, Z, tx0, ty0, b * K + k);
        }
        tma_commit();
    }
    // box word `it` (row r, word c) of a plane: pixels tx0 - hx + 32 c ... + 31 of image row ty0 - d + r, zero outside
    const int n_box = R * BW;
    auto fetch = [&](int j, int it) -> uint32_t {
        if (j >= n_live || it >= n_box) return 0u;
        const int r = it / BW, c = it - r * BW;
        const int gy = ty0 - d + r;
        if (gy < 0 || gy >= H) return 0u;
        const int x0 = tx0 - hx + 32 * c;
        const int wq = x0 >> 5, sh = x0 & 31;                          // arithmetic shift: floor for negative x0
        const uint32_t *row = planes + ((size_t)(b * K + plane_of(j)) * H + gy) * WB;
        const uint32_t lo = (wq >= 0 && wq < WB) ? __ldg(row + wq) : 0u;
        const uint32_t hi = (sh && wq + 1 >= 0 && wq + 1 < WB) ? __ldg(row + wq + 1) : 0u;
        return __funnelshift_r(lo, hi, sh);
    };
    constexpr int NF = 3;                       // box words per thread: R * BW <= (64 + 128) * 6 ... covered in NF passes for d <= 16
    const int passes = (n_box + B_THREADS - 1) / B_THREADS;
    uint32_t nxt[NF];
#pragma unroll
    for (int q = 0; q < NF; ++q) nxt[q] = fetch(0, tid + q * B_THREADS);
    const int row = tid >> 2, jw = tid & 3;     // this thread's 32 output pixels: tile row, 32-pixel column group
    uint32_t uni_acc[8], fg_acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) uni_acc[q] = fg_acc[q] = 0;
    int o_slot = 0;
    for (int k = 0; k < n_live; ++k) {
        uint32_t any = 0u;
#pragma unroll
        for (int q = 0; q < NF; ++q) {
            const int it = tid + q * B_THREADS;
            if (it < n_box) { bits[(it / BW) * RW32 + (it % BW)] = nxt[q]; any |= nxt[q]; }
        }
        for (int q = NF; q < passes; ++q) {      // large radii: the rest of the box, not prefetched
            const int it = tid + q * B_THREADS;
            if (it < n_box) { const uint32_t x = fetch(k, it); bits[(it / BW) * RW32 + (it % BW)] = x; any |= x; }
        }
#pragma unroll
        for (int q = 0; q < NF; ++q) nxt[q] = fetch(k + 1, tid + q * B_THREADS);
        if (!__syncthreads_or(any != 0u)) {       // S1: the box is empty after all (occupancy is per 128 x 16 tile)
            if (has_per && tid == 0) { tma_store_3d(&out_map, Z, tx0, ty0, b * K + plane_of(k)); tma_commit(); }
            continue;
        }
        // ---- horizontal erosion on bit rows: window [x-d, x+d] for the four central 32-pixel words ----
        for (int it = tid; it < R * 4; it += B_THREADS) {
            const uint32_t *row32 = bits + (it >> 2) * RW32;
            const int s0 = hx + 32 * (it & 3) - d;
            uint32_t e = 0xffffffffu;
            for (int t = 0; t < w; ++t) {
                const int s = s0 + t;
                e &= __funnelshift_r(row32[s >> 5], row32[(s >> 5) + 1], s & 31);
            }
            Hb[it] = e;
        }
        if (has_per && tid == 0) tma_wait_read<1>();     // the store that last used this O slot has read it
        __syncthreads();                                 // S2
        uint32_t e = 0xffffffffu;
        for (int t = 0; t < w; ++t) e &= Hb[(row + t) * 4 + jw];
        const uint32_t *mrow = bits + (row + d) * RW32 + (hx >> 5) + jw;
        const uint32_t m = __funnelshift_r(mrow[0], mrow[1], hx & 31);
        const uint32_t bd = m & ~e;                      // mask - eroded, one bit per pixel
        uint32_t bdw[8], mw[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            bdw[q] = nibble_to_bytes01((bd >> (4 * q)) & 15u) * v;
            mw[q] = nibble_to_bytes01((m >> (4 * q)) & 15u) * v;
            uni_acc[q] = add_bytes(uni_acc[q], bdw[q]);
            fg_acc[q] = add_bytes(fg_acc[q], mw[q]);
        }
        if (has_per) {
            uint4 *o4 = reinterpret_cast<uint4 *>(O0 + (size_t)o_slot * B_TH * B_TWW + row * B_TWW + 8 * jw);
            o4[0] = make_uint4(bdw[0], bdw[1], bdw[2], bdw[3]);
            o4[1] = make_uint4(bdw[4], bdw[5], bdw[6], bdw[7]);
            fence_proxy_async();
        }
        __syncthreads();                                 // S3: O tile complete; bits / Hb are free again
        if (has_per && tid == 0) {
            tma_store_3d(&out_map, O0 + (size_t)o_slot * B_TH * B_TWW, tx0, ty0, b * K + plane_of(k));
            tma_commit();
        }
        o_slot ^= 1;
    }
    const int gy = ty0 + row, gx = tx0 + 32 * jw;
    if (gy < H && gx < W) {
        const size_t o = (size_t)b * HW + (size_t)gy * W + gx;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (boundary)
                st_stream_u4(boundary + o + 16 * h, make_uint4(nonzero_bytes01(uni_acc[4 * h]), nonzero_bytes01(uni_acc[4 * h + 1]),
                                                               nonzero_bytes01(uni_acc[4 * h + 2]), nonzero_bytes01(uni_acc[4 * h + 3])));
            if (fg)
                st_stream_u4(fg + o + 16 * h, make_uint4(nonzero_bytes01(fg_acc[4 * h]), nonzero_bytes01(fg_acc[4 * h + 1]),
                                                         nonzero_bytes01(fg_acc[4 * h + 2]), nonzero_bytes01(fg_acc[4 * h + 3])));
        }
    }
    if (tid == 0 || tid == 32) tma_wait_all();     // shared memory must outlive the bulk stores that read it
}

__global__ void __launch_bounds__(256)
quadruple_kernel(const uint8_t *__restrict__ gt, const uint8_t *__restrict__ pred, size_t n, uint8_t *__restrict__ out) {
    const int b = blockIdx.y;
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (i >= n) return;
    const uint8_t *g = gt + (size_t)b * n + i, *p = pred + (size_t)b * n + i;
    uint8_t *o = out + (size_t)b * 4 * n + i;
    if (i + 16 <= n && ((n & 15) == 0)) {
        const uint4 gv = ld_stream_u4(g), pv = ld_stream_u4(p);
        uint4 tp, tn, fp, fn;
#define QUAD(c)                                                   \
        {                                                         \
            const uint32_t G = __vcmpne4(gv.c, 0u), P = __vcmpne4(pv.c, 0u); \
            tp.c = (G & P) & 0x01010101u; tn.c = (~G & ~P) & 0x01010101u;     \
            fp.c = (~G & P) & 0x01010101u; fn.c = (G & ~P) & 0x01010101u;     \
        }
        QUAD(x) QUAD(y) QUAD(z) QUAD(w)
#undef QUAD
        st_stream_u4(o, tp); st_stream_u4(o + n, tn); st_stream_u4(o + 2 * n, fp); st_stream_u4(o + 3 * n, fn);
    } else {
        for (int j = 0; j < 16 && i + j < n; ++j) {
            const bool G = g[j] != 0, P = p[j] != 0;
            o[j] = G && P; o[n + j] = !G && !P; o[2 * n + j] = !G && P; o[3 * n + j] = G && !P;
        }
    }
}

}  // namespace quber

extern "C" int quber_boundary_dilation(int H, int W, double ratio) {
    // util.py:80-84: int(round(ratio * sqrt(h^2 + w^2))) (Python round: half to even), at least 1
    const double diag = sqrt((double)H * H + (double)W * W);
    const long dd = (long)nearbyint(ratio * diag);
    return dd < 1 ? 1 : (int)dd;
}

namespace quber {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
// uint8 tensor [planes, H, W] viewed through a (box_w x box_h x 1) window; out-of-range bytes read as 0
static bool make_map(CUtensorMap *m, const void *base, int planes, int H, int W, int box_w, int box_h) {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)W, (cuuint64_t)W * H};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
}  // namespace quber

extern "C" int quber_mask_boundary(const uint8_t *masks, int B, int K, int H, int W, int d, uint8_t *per_instance,
                                   uint8_t *boundary, uint8_t *fg, const uint8_t *occupancy, int32_t *status, void *stream) {
    using namespace quber;
    if (!masks || B <= 0 || K <= 0 || H <= 0 || W <= 0 || d < 1 || d > 64) return QUBER_E_INVALID;
    if (!per_instance && !boundary && !fg) return QUBER_E_INVALID;
    if (W % 16 != 0 || !aligned16(masks) || (per_instance && !aligned16(per_instance)) ||
        (boundary && !aligned16(boundary)) || (fg && !aligned16(fg)))
        return QUBER_E_ALIGN;
    if (B > 65535) return QUBER_E_UNSUPPORTED;
    dim3 grid(cdiv(W, B_TW), cdiv(H, B_TH), B);
    const BGeom g(d);
    CUtensorMap in_map, out_map;
    if (!make_map(&in_map, masks, B * K, H, W, g.SW, g.R) ||
        !make_map(&out_map, per_instance ? per_instance : masks, B * K, H, W, B_TW, B_TH)) {
        set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled");
        return QUBER_E_CUDA;
    }
    QUBER_CUDA_TRY(cudaFuncSetAttribute(boundary_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    if (status) QUBER_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int32_t), (cudaStream_t)stream));
    boundary_tma_kernel<<<grid, B_THREADS, g.smem, (cudaStream_t)stream>>>(in_map, out_map, K, H, W, d, per_instance != nullptr,
                                                                          boundary, fg, status, occupancy,
                                                                          cdiv(H, QUBER_OCC_TILE_H), cdiv(W, QUBER_OCC_TILE_W));
    return check_launch("boundary_tma_kernel");
}

extern "C" int quber_mask_boundary_bits(const uint32_t *packed, int value, int B, int K, int H, int W, int d, uint8_t *per_instance,
                                        uint8_t *boundary, uint8_t *fg, const uint8_t *occupancy, void *stream) {
    using namespace quber;
    if (!packed || B <= 0 || K <= 0 || H <= 0 || W <= 0 || d < 1 || d > 64 || value < 1 || value > 255) return QUBER_E_INVALID;
    if (!per_instance && !boundary && !fg) return QUBER_E_INVALID;
    if (W % 32 != 0 || (reinterpret_cast<uintptr_t>(packed) & 3u) || (per_instance && !aligned16(per_instance)) ||
        (boundary && !aligned16(boundary)) || (fg && !aligned16(fg)))
        return QUBER_E_ALIGN;
    if (B > 65535) return QUBER_E_UNSUPPORTED;
    dim3 grid(cdiv(W, B_TW), cdiv(H, B_TH), B);
    const BGeom g(d);
    CUtensorMap out_map;
    if (!make_map(&out_map, per_instance ? (const void *)per_instance : (const void *)packed, B * K, H, W, B_TW, B_TH)) {
        set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled");
        return QUBER_E_CUDA;
    }
    const size_t smem = 3 * (size_t)B_TH * B_TWW * 4 + g.bits_bytes + g.hb_bytes + 128;
    QUBER_CUDA_TRY(cudaFuncSetAttribute(boundary_bits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    boundary_bits_kernel<<<grid, B_THREADS, smem, (cudaStream_t)stream>>>(packed, (uint32_t)value, out_map, K, H, W, d,
                                                                         per_instance != nullptr, boundary, fg, occupancy,
                                                                         cdiv(H, QUBER_OCC_TILE_H), cdiv(W, QUBER_OCC_TILE_W));
    return check_launch("boundary_bits_kernel");
}

extern "C" int quber_tma_timeouts(void) {
    unsigned int n = 0;
    if (cudaMemcpyFromSymbol(&n, quber::g_tma_timeouts, sizeof(n)) != cudaSuccess) return -1;
    return (int)n;
}

extern "C" int quber_quadruple_maps(const uint8_t *gt, const uint8_t *pred, int B, size_t n, uint8_t *out, void *stream) {
    using namespace quber;
    if (!gt || !pred || !out || B <= 0 || n == 0) return QUBER_E_INVALID;
    if ((n % 16 == 0) && (!aligned16(gt) || !aligned16(pred) || !aligned16(out))) return QUBER_E_ALIGN;
    dim3 grid((unsigned)((n + 16 * 256 - 1) / (16 * 256)), B);
    quadruple_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(gt, pred, n, out);
    return check_launch("quadruple_kernel");
}
