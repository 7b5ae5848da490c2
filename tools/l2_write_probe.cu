// Probe: how fast can one SM push a 32 KB shared-memory block (128 rows x 64 floats) into an L2-resident gradient buffer?
// Answers the question behind engine 2's W1-gradient flush: per-thread stores / reductions versus
// the TMA engine's bulk store / bulk reduction, with the destination shared by all CTAs (a reduction) or private.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/_variants/l2_write_probe tools/l2_write_probe.cu
//   tools/_variants/l2_write_probe            (on the GPU box)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ROWS = 128, COLS = 64, ROW_STRIDE = 600;  // the critic's W1 gradient: [128][600] floats, 64-column blocks
constexpr int BLOCKS = 10, ITERS = 8;                   // one "tile" = 10 blocks; 8 tiles per CTA

__device__ __forceinline__ void red_v4(float* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_store(void* g, const void* s, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(s32(s)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_red(void* g, const void* s, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(g), "r"(s32(s)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// mode: 0 STG.128 rows (256 B segments, row stride 2400 B)   1 STG.128 contiguous 32 KB
//       2 RED.v4 rows                                          3 RED.v4 contiguous
//       4 TMA bulk store, one 32 KB copy                       5 TMA bulk store, 128 row copies of 256 B
//       6 TMA bulk reduction, one 32 KB op                     7 TMA bulk reduction, 128 row ops of 256 B
// shared_dst: all CTAs write the same buffer (as a gradient reduction does) or each its own
__global__ void __launch_bounds__(256, 1) probe(float* base, int mode, int shared_dst, long long* cycles) {
    extern __shared__ __align__(128) float4 blk[];  // [128][16] float4
    for (int i = threadIdx.x; i < ROWS * COLS / 4; i += 256) blk[i] = make_float4(1.f, 1.f, 1.f, 1.f);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    float* dst0 = base + (shared_dst ? 0 : (size_t)blockIdx.x * ROWS * 640);
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
        for (int b0 = 0; b0 < BLOCKS; ++b0) {
            const int b = (b0 + blockIdx.x) % BLOCKS;
            const bool rows = mode == 0 || mode == 2 || mode == 5 || mode == 7;
            float* dst = rows ? dst0 + b * COLS : dst0 + (size_t)b * ROWS * COLS;
            if (mode <= 3) {
                for (int F = threadIdx.x; F < ROWS * 16; F += 256) {
                    const int r = F >> 4, c4 = F & 15;
                    const float4 v = blk[r * 16 + (c4 ^ (r & 7))];
                    float* p = rows ? dst + (size_t)r * ROW_STRIDE + 4 * c4 : dst + 4 * F;
                    if (mode <= 1) *reinterpret_cast<float4*>(p) = v;
                    else red_v4(p, v);
                }
            } else if (mode == 4 || mode == 6) {
                if (threadIdx.x == 0) {
                    if (mode == 4) bulk_store(dst, blk, ROWS * COLS * 4);
                    else bulk_red(dst, blk, ROWS * COLS * 4);
                    bulk_commit();
                }
            } else {
                if (threadIdx.x < ROWS) {
                    const int r = threadIdx.x;
                    if (mode == 5) bulk_store(dst + (size_t)r * ROW_STRIDE, blk + r * 16, COLS * 4);
                    else bulk_red(dst + (size_t)r * ROW_STRIDE, blk + r * 16, COLS * 4);
                    bulk_commit();
                }
            }
        }
    }
    if (mode >= 4) bulk_wait_all();
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
    float* buf;
    long long* cyc;
    const size_t per_cta = (size_t)ROWS * 640 * sizeof(float);
    cudaMalloc(&buf, per_cta * 148);
    cudaMalloc(&cyc, 148 * sizeof(long long));
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    const char* names[8] = {"STG.128 rows", "STG.128 contiguous", "RED.v4 rows", "RED.v4 contiguous", "TMA store 32 KB", "TMA store 128 x 256 B",
                            "TMA reduce 32 KB", "TMA reduce 128 x 256 B"};
    for (int grid : {1, 51, 97, 148}) {
        for (int shared_dst = 0; shared_dst < 2; ++shared_dst) {
            for (int mode = 0; mode < 8; ++mode) {
                cudaMemset(buf, 0, per_cta * 148);
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                probe<<<grid, 256, 32768>>>(buf, mode, shared_dst, cyc);  // warm
                cudaEventRecord(e0);
                probe<<<grid, 256, 32768>>>(buf, mode, shared_dst, cyc);
                cudaEventRecord(e1);
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("mode %d failed: %s\n", mode, cudaGetErrorString(cudaGetLastError())); return 1; }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                long long h[148];
                cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
                long long mx = 0;
                for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
                const double bytes = (double)ITERS * BLOCKS * ROWS * COLS * 4;
                printf("grid %3d  %-7s  %-24s  %8lld cycles  %6.1f B/clk/SM  %7.1f us  (%.0f cycles per 32 KB block)\n", grid, shared_dst ? "shared" : "private",
                       names[mode], mx, bytes / mx, ms * 1e3, (double)mx / (ITERS * BLOCKS));
            }
        }
    }
    return 0;
}
