// Stand-alone probe of the TMA path quber_mask_boundary uses: one 3-D uint8 tiled load (with
// out-of-image start coordinates) into shared memory, copied back through a TMA store and by plain
// stores, compared with the host.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu
// Run:   ./tma_probe W H planes box_w box_h x y z [nowait]
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e), __FILE__, __LINE__, #x); return 2; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map, int box_bytes,
                      int x, int y, int z, uint8_t *plain_out, int *dbg, int do_store) {
    extern __shared__ __align__(128) unsigned char buf[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(box_bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(smem_u32(buf)), "l"(&in_map), "r"(smem_u32(&bar)), "r"(x), "r"(y), "r"(z) : "memory");
    }
    uint32_t done = 0;
    int spin = 0;
    for (; !done && spin < (1 << 22); ++spin)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    if (threadIdx.x == 0) { dbg[0] = done; dbg[1] = spin; }
    if (!done) return;
    for (int i = threadIdx.x; i < box_bytes; i += blockDim.x) plain_out[i] = buf[i];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0 && do_store) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                     ::"l"(&out_map), "r"(smem_u32(buf)), "r"(x), "r"(y), "r"(z) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        dbg[2] = 1;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char **argv) {
    if (argc < 9) { printf("usage\n"); return 1; }
    const int W = atoi(argv[1]), H = atoi(argv[2]), P = atoi(argv[3]), bw = atoi(argv[4]), bh = atoi(argv[5]);
    const int x = atoi(argv[6]), y = atoi(argv[7]), z = atoi(argv[8]);
    const int promo = argc > 9 ? atoi(argv[9]) : 1;
    const int do_store = argc > 10 ? atoi(argv[10]) : 1;
    void *fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess) { printf("no entry point\n"); return 2; }
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    const size_t n = (size_t)W * H * P;
    std::vector<uint8_t> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = (uint8_t)(1 + (i * 2654435761u >> 13) % 255);
    uint8_t *d_in, *d_out, *d_plain;
    int *d_dbg;
    CK(cudaMalloc(&d_in, n)); CK(cudaMalloc(&d_out, n)); CK(cudaMalloc(&d_plain, bw * bh)); CK(cudaMalloc(&d_dbg, 16));
    CK(cudaMemcpy(d_in, h.data(), n, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_out, 0, n)); CK(cudaMemset(d_dbg, 0, 16)); CK(cudaMemset(d_plain, 0xee, bw * bh));
    CUtensorMap im, om;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)P};
    const cuuint64_t strides[2] = {(cuuint64_t)W, (cuuint64_t)W * H};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    CUresult r1 = enc(&im, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_in, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, promo ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = enc(&om, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, promo ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d %d\n", (int)r1, (int)r2);
    if (r1 || r2) return 3;
    const int bytes = bw * bh;
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes + 128));
    probe<<<1, 256, bytes + 128>>>(im, om, bytes, x, y, z, d_plain, d_dbg, do_store);
    cudaError_t e = cudaDeviceSynchronize();
    printf("sync: %s\n", cudaGetErrorName(e));
    if (e != cudaSuccess) return 4;
    int dbg[4];
    CK(cudaMemcpy(dbg, d_dbg, 16, cudaMemcpyDeviceToHost));
    printf("done=%d spins=%d stored=%d\n", dbg[0], dbg[1], dbg[2]);
    std::vector<uint8_t> pl(bytes), out(n);
    CK(cudaMemcpy(pl.data(), d_plain, bytes, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out.data(), d_out, n, cudaMemcpyDeviceToHost));
    long bad_l = 0, bad_s = 0;
    for (int j = 0; j < bh; ++j)
        for (int i = 0; i < bw; ++i) {
            const int gx = x + i, gy = y + j;
            const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
            const uint8_t want = in ? h[((size_t)z * H + gy) * W + gx] : 0;
            bad_l += pl[j * bw + i] != want;
            if (in && do_store) bad_s += out[((size_t)z * H + gy) * W + gx] != want;
        }
    printf("load mismatches=%ld store mismatches=%ld\n", bad_l, bad_s);
    return (bad_l || bad_s) ? 5 : 0;
}
