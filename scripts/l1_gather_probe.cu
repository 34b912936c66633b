// Probe: what does the B200 L1 deliver for the MSDA gather pattern with NO set-up arithmetic at all?
// Same data layout and lane mapping as msda_fused_p4_kernel at C2 (value [B,S,M,32] fp32; 8 lanes x 16 B per (query, head);
// 12 taps x 4 corners = 48 gathered 128-byte lines per (query, head); 256 threads = 32 consecutive queries of one head),
// tap positions = the query's own pixel in every level + a fixed per-head / per-point integer offset (the module's initial
// offset grid), corner addresses by integer arithmetic only, corner values summed (4 FADD per corner instead of FMA + weights).
// The time of this kernel is the practical ceiling of any kernel that gathers these lines this way.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/l1_gather_probe scripts/l1_gather_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int M = 8, D = 32, L = 3, P = 4;
__constant__ int c_h[L], c_w[L], c_start[L];

template <int VARIANT, int MINB>
__global__ void __launch_bounds__(256, MINB) probe(const float *__restrict__ value, int S, int Q, float *__restrict__ out) {
    const int tid = threadIdx.x, lane = tid & 7, grp = tid >> 3;
    const int m = blockIdx.y, b = blockIdx.z;
    const int q = min(blockIdx.x * 32 + grp, Q - 1);
    // the query's pixel: level and (x, y)
    int lq = 0;
#pragma unroll
    for (int l = 1; l < L; ++l) if (q >= c_start[l]) lq = l;
    const int r = q - c_start[lq];
    const int qy = r / c_w[lq], qx = r - qy * c_w[lq];
    const float fx = (qx + 0.5f) / c_w[lq], fy = (qy + 0.5f) / c_h[lq];
    const int dxm = (m == 0 || m == 1 || m == 7) ? 1 : ((m >= 3 && m <= 5) ? -1 : 0);
    const int dym = (m >= 1 && m <= 3) ? 1 : ((m >= 5) ? -1 : 0);
    const float4 *v4 = reinterpret_cast<const float4 *>(value + ((size_t)b * S * M + m) * D) + lane;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int l = 0; l < L; ++l) {
        const int h = c_h[l], w = c_w[l];
        const int bx = (int)(fx * w - 0.5f), by = (int)(fy * h - 0.5f);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int x0 = min(max(bx + dxm * (p + 1), 0), w - 2), y0 = min(max(by + dym * (p + 1), 0), h - 2);
            const uint32_t a = (uint32_t)(c_start[l] + y0 * w + x0) * (M * D / 4);
            const uint32_t dx = M * D / 4, dy = (uint32_t)w * (M * D / 4);
            if (VARIANT == 0) {
                const float4 r1 = __ldg(v4 + a), r2 = __ldg(v4 + a + dx), r3 = __ldg(v4 + a + dy), r4 = __ldg(v4 + a + dy + dx);
                acc.x += (r1.x + r2.x) + (r3.x + r4.x); acc.y += (r1.y + r2.y) + (r3.y + r4.y);
                acc.z += (r1.z + r2.z) + (r3.z + r4.z); acc.w += (r1.w + r2.w) + (r3.w + r4.w);
            } else {                      // two of the four corners only: how much of the time is the gather itself
                const float4 r1 = __ldg(v4 + a), r3 = __ldg(v4 + a + dy);
                acc.x += r1.x + r3.x; acc.y += r1.y + r3.y; acc.z += r1.z + r3.z; acc.w += r1.w + r3.w;
            }
        }
    }
    if (blockIdx.x * 32 + grp < Q)
        reinterpret_cast<float4 *>(out + (((size_t)b * Q + q) * M + m) * D)[lane] = acc;
}

int main() {
    const int B = 32, hs[L] = {60, 30, 15}, ws[L] = {80, 40, 20};
    int st[L], S = 0;
    for (int l = 0; l < L; ++l) { st[l] = S; S += hs[l] * ws[l]; }
    const int Q = S;
    cudaMemcpyToSymbol(c_h, hs, sizeof(hs)); cudaMemcpyToSymbol(c_w, ws, sizeof(ws)); cudaMemcpyToSymbol(c_start, st, sizeof(st));
    const size_t n = (size_t)B * S * M * D;
    float *val[2], *out;
    for (int i = 0; i < 2; ++i) { cudaMalloc(&val[i], n * 4); cudaMemset(val[i], 0, n * 4); }
    cudaMalloc(&out, n * 4);
    dim3 grid((Q + 31) / 32, M, B);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto run = [&](int variant, int minb, int smem, int i) {
#define P_(V, MB) probe<V, MB><<<grid, 256, smem>>>(val[i & 1], S, Q, out)
        if (variant == 0) { if (minb == 5) P_(0, 5); else if (minb == 6) P_(0, 6); else P_(0, 8); }
        else { if (minb == 5) P_(1, 5); else if (minb == 6) P_(1, 6); else P_(1, 8); }
#undef P_
    };
    // resident CTAs per SM: asked of the compiler (registers) and limited from above by dynamic shared memory
    const int cfg[][2] = {{5, 72 * 1024}, {5, 44 * 1024}, {5, 0}, {6, 0}, {8, 0}};       // -> 3, 5, 5 (by registers: up to 8), 6.., 8
    for (auto &c : cfg) {
        cudaFuncSetAttribute(probe<0, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
        cudaFuncSetAttribute(probe<1, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
        for (int variant = 0; variant < 2; ++variant) {
            for (int i = 0; i < 4; ++i) run(variant, c[0], c[1], i);
            cudaDeviceSynchronize();
            const int N = 20;
            cudaEventRecord(e0);
            for (int i = 0; i < N; ++i) run(variant, c[0], c[1], i);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            int nb = 0;
            if (c[0] == 5) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, probe<0, 5>, 256, c[1]);
            else if (c[0] == 6) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, probe<0, 6>, 256, c[1]);
            else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, probe<0, 8>, 256, c[1]);
            const double lines = (double)B * Q * M * (variant == 0 ? 48 : 24);
            printf("minb %d smem %d KB (%d CTAs/SM) variant %d (%d corner lines per (query, head)): %.1f us/launch, %.2f TB/s through L1, err=%d\n",
                   c[0], c[1] / 1024, nb, variant, variant == 0 ? 48 : 24, ms / N * 1e3, lines * 128 / (ms / N * 1e-3) / 1e12, (int)cudaGetLastError());
        }
    }
    return 0;
}
