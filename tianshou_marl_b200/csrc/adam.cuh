// adam.cuh -- block-level optimizer step shared by adam_kernel (mlp.cu) and the tcgen05 minibatch
// kernel's fused tail (mlp_tc.cu): global grad-norm clip (torch.nn.utils.clip_grad_norm_,
// reference algorithm_base.py:485-498) + torch.optim.Adam (optim.py:89-110); zeroes the gradients
// and, for the tcgen05 engine, refreshes the fp16 operand image of the parameters.
#pragma once
#include <cooperative_groups.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "ppo_math.cuh"

namespace tsm {
namespace opt {

// ---- fp16 operand image of a in(<=31) -> 64 -> 64 -> out(<=8) net (tcgen05 engine) ----------
// halves, chunked [k/8][n][k%8] per matrix (tc.cuh); biases ride on input feature 31 (a column of
// ones): b1 inside W1, b2 / b3 as 16-deep K blocks multiplied with X features 16..31.
constexpr int IMG_W1 = 0;                      // [4][64][8]
constexpr int IMG_W2 = IMG_W1 + 4 * 64 * 8;    // [8][64][8]
constexpr int IMG_W3 = IMG_W2 + 8 * 64 * 8;    // [8][16][8]
constexpr int IMG_B2 = IMG_W3 + 8 * 16 * 8;    // [2][64][8], only k = 15 used
constexpr int IMG_B3 = IMG_B2 + 2 * 64 * 8;    // [2][16][8]
constexpr int IMG_HALVES = IMG_B3 + 2 * 16 * 8;
constexpr int IMG_BYTES = IMG_HALVES * 2;      // 16 896
// IMG_COPIES identical copies (CTA b reads copy b % IMG_COPIES) would spread the start-up reads of a launch over L2;
// measured on B200: no gain for the readers and 2x the optimizer kernel's time at 16 copies, so one copy.
constexpr int IMG_COPIES = 1;
constexpr int IMG_ALLOC = IMG_BYTES * IMG_COPIES;

__host__ __device__ inline int image_index(int64_t i, const ppo::Offsets& o, int in_dim) {
    if (i < o.b1) {
        const int out = (int)(i / in_dim), in = (int)(i - (int64_t)out * in_dim);
        return IMG_W1 + ((in >> 3) * 64 + out) * 8 + (in & 7);
    }
    if (i < o.w2) return IMG_W1 + (3 * 64 + (int)(i - o.b1)) * 8 + 7;
    if (i < o.b2) {
        const int rel = (int)(i - o.w2), out = rel >> 6, in = rel & 63;
        return IMG_W2 + ((in >> 3) * 64 + out) * 8 + (in & 7);
    }
    if (i < o.w3) return IMG_B2 + (64 + (int)(i - o.b2)) * 8 + 7;
    if (i < o.b3) {
        const int rel = (int)(i - o.w3), out = rel >> 6, in = rel & 63;
        return IMG_W3 + ((in >> 3) * 16 + out) * 8 + (in & 7);
    }
    return IMG_B3 + (16 + (int)(i - o.b3)) * 8 + 7;
}

struct AdamArgs {
    float* p;
    float* g;
    float* m;
    float* v;
    int64_t* step;
    int64_t n;
    float lr, beta1, beta2, eps, max_norm, grad_scale;
    float* gnorm_out;
    __half* image;  // NULL unless engine 1
    int in_dim, out_dim;
    // data-parallel ranks on one node (world == 1: none of this is used).  Every rank PUSHES its local gradient g into
    // slot [seq & 1][rank] of every rank's receive area (peer_recv[q]: device pointers mapped over NVLink; plain stores, no
    // round trip), signals, waits for all ranks, and sums the world slots of its own receive area in rank order -- the
    // same order everywhere, so parameters stay bit-identical across ranks.
    int world, rank;
    float* peer_recv[8];      // [2][world][n_stride] floats on rank q
    uint32_t* peer_flag[8];   // peer_flag[q][r]: written by rank r into rank q's flag pad
    uint4* peer_ll[8];        // [2][world][n_stride / 2] x {v0, step, v1, step}: flag-in-data receive area (persistent epoch kernel)
    int64_t n_stride;         // n rounded up to a multiple of 4
    uint32_t seq;             // step number used as the flag value
    int* status;              // set to 1 if the cross-rank wait times out
    int debug;                // experiments only (TSM_DIST_DEBUG): 1 skip the barrier, 2 local gradients only
};
constexpr int ADAM_MAX_WORLD = 8;

// beta^t for integer t by binary exponentiation (a serial double pow() costs microseconds on one thread)
__device__ __forceinline__ double ipow(double b, int64_t t) {
    double r = 1.0;
    while (t > 0) {
        if (t & 1) r *= b;
        b *= b;
        t >>= 1;
    }
    return r;
}

// One Adam update of element i (torch.optim.Adam, no weight decay / amsgrad)
__device__ __forceinline__ float adam_elem(float p, float g, float& m, float& v, float coef, float step_size, float bc2_sqrt,
                                           float beta1, float beta2, float eps) {
    const float gi = g * coef;
    m = m + (gi - m) * (1.f - beta1);               // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(gi * gi, 1.f - beta2, v * beta2);      // mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = sqrtf(v) / bc2_sqrt + eps;
    return p - step_size * (m / denom);
}

// Same update with hardware approximations (sqrt.approx, rcp.approx: ~1 ulp each, ~15 instructions instead of ~70):
// used by the tcgen05 engine's persistent epoch kernel, where every CTA steps all parameters.
__device__ __forceinline__ float adam_elem_fast(float p, float g, float& m, float& v, float coef, float step_size, float inv_bc2_sqrt,
                                                float beta1, float beta2, float eps) {
#ifdef TSM_TC_EXACT_ADAM  // instrumented builds only (tools/tc_error_budget.py)
    return adam_elem(p, g, m, v, coef, step_size, 1.f / inv_bc2_sqrt, beta1, beta2, eps);
#endif
    const float gi = g * coef;
    m = m + (gi - m) * (1.f - beta1);
    v = fmaf(gi * gi, 1.f - beta2, v * beta2);
    float sq, r;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(v));
    asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(fmaf(sq, inv_bc2_sqrt, eps)));
    return fmaf(-step_size * m, r, p);
}

// One cluster of ADAM_CTAS CTAs: thread-owned float4 slices, so the gradient is read once, kept in registers for
// both the norm and the update, and zeroed by its owner; the squared-norm partials are exchanged through
// distributed shared memory (every CTA sums the same partials in the same order: identical clip coefficient).
constexpr int ADAM_CTAS = 8;
constexpr int ADAM_THREADS = 192;
constexpr int ADAM_SLOTS = 4;  // float4 per thread held in registers: covers n <= 8 * 192 * 4 * 4 = 24 576 parameters

__device__ __forceinline__ void adam_store_image(const AdamArgs& a, const ppo::Offsets& o, int64_t i, float pn) {
    const int idx = image_index(i, o, a.in_dim);
    const __half h = __float2half_rn(pn);
#pragma unroll
    for (int c = 0; c < IMG_COPIES; ++c) a.image[c * IMG_HALVES + idx] = h;
}

// my gradient slice -> slot [seq & 1][rank] of every rank's receive area
__device__ __forceinline__ int64_t adam_slot(const AdamArgs& a, int q) { return ((int64_t)(a.seq & 1u) * a.world + q) * a.n_stride; }
__device__ __forceinline__ void adam_push4(const AdamArgs& a, int64_t j, float4 v) {
    for (int q = 0; q < a.world; ++q)
        if (q != a.rank) reinterpret_cast<float4*>(a.peer_recv[q] + adam_slot(a, a.rank))[j] = v;
}
__device__ __forceinline__ void adam_push1(const AdamArgs& a, int64_t i, float v) {
    for (int q = 0; q < a.world; ++q)
        if (q != a.rank) (a.peer_recv[q] + adam_slot(a, a.rank))[i] = v;
}
// sum over ranks in rank order; `mine` is this rank's own contribution (already in registers)
__device__ __forceinline__ float4 adam_sum4(const AdamArgs& a, int64_t j, float4 mine) {
    if (a.debug & 2) return mine;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int q = 0; q < a.world; ++q) {
        const float4 x = q == a.rank ? mine : __ldcv(reinterpret_cast<const float4*>(a.peer_recv[a.rank] + adam_slot(a, q)) + j);
        s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
    }
    return s;
}
__device__ __forceinline__ float adam_sum1(const AdamArgs& a, int64_t i, float mine) {
    if (a.debug & 2) return mine;
    float s = 0.f;
    for (int q = 0; q < a.world; ++q) s += q == a.rank ? mine : __ldcv(a.peer_recv[a.rank] + adam_slot(a, q) + i);
    return s;
}

// Cross-rank barrier of step `seq`, executed by one thread: publish "my gradient buffer of this step is complete" to
// every peer's flag pad, then wait for every peer's flag.  The wait is bounded (about 2 s): a missing rank sets
// *status instead of hanging the device.  Receive areas are double buffered by step parity, so one barrier per step
// suffices: a rank overwrites parity b of a peer only after that peer has signalled the step after it last read b.
__device__ __forceinline__ void adam_rank_barrier(const AdamArgs& a) {
    __threadfence_system();
    for (int q = 0; q < a.world; ++q)
        if (q != a.rank) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.peer_flag[q] + a.rank), "r"(a.seq) : "memory");
    const long long t0 = clock64();
    for (int q = 0; q < a.world; ++q) {
        if (q == a.rank) continue;
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.peer_flag[a.rank] + q) : "memory");
            if ((int32_t)(v - a.seq) >= 0) break;
            if (clock64() - t0 > 4000000000LL) {
                if (a.status) *a.status = 1;
                break;
            }
        } while (true);
    }
}

__device__ __forceinline__ void adam_cluster(const AdamArgs& a) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ float sh[ADAM_THREADS / 32];
    __shared__ float s_part;
    __shared__ float s_c[3];
    const int tid = threadIdx.x;
    const int64_t gtid = (int64_t)cluster.block_rank() * ADAM_THREADS + tid;
    constexpr int64_t GT = (int64_t)ADAM_CTAS * ADAM_THREADS;
    float* __restrict__ P = a.p;
    float* __restrict__ G = a.g;
    float* __restrict__ M = a.m;
    float* __restrict__ V = a.v;
    uintptr_t al = reinterpret_cast<uintptr_t>(P) | reinterpret_cast<uintptr_t>(G) | reinterpret_cast<uintptr_t>(M) |
                   reinterpret_cast<uintptr_t>(V);
    for (int q = 0; q < a.world && a.world > 1; ++q) al |= reinterpret_cast<uintptr_t>(a.peer_recv[q]);
    const bool vec = (al & 15) == 0;
    const int64_t n4 = vec ? a.n / 4 : 0;  // float4 body, scalar tail
    const bool in_regs = n4 <= GT * ADAM_SLOTS;
    const bool dist = a.world > 1;
    float4 g[ADAM_SLOTS], p[ADAM_SLOTS], m[ADAM_SLOTS], v[ADAM_SLOTS];
    float sq = 0.f;
    const float gs = a.grad_scale;
    // ---- local gradient (and, across ranks, its push to every peer) ----
    if (in_regs) {
#pragma unroll
        for (int u = 0; u < ADAM_SLOTS; ++u) {
            const int64_t j = gtid + u * GT;
            if (j < n4) {
                g[u] = __ldcg(reinterpret_cast<const float4*>(G) + j);
                if (dist) adam_push4(a, j, g[u]);
                p[u] = reinterpret_cast<const float4*>(P)[j];
                m[u] = reinterpret_cast<const float4*>(M)[j];
                v[u] = reinterpret_cast<const float4*>(V)[j];
            }
        }
    } else if (dist) {
        for (int64_t j = gtid; j < n4; j += GT) adam_push4(a, j, __ldcg(reinterpret_cast<const float4*>(G) + j));
    }
    if (dist) {
        for (int64_t i = 4 * n4 + gtid; i < a.n; i += GT) adam_push1(a, i, __ldcg(G + i));
        cluster.sync();  // all pushes of this cluster happen-before the system-scope fence + flag stores of thread 0 below
        if (gtid == 0 && !(a.debug & 1)) adam_rank_barrier(a);
        cluster.sync();
    }
    // ---- total gradient and its squared norm ----
    if (in_regs) {
#pragma unroll
        for (int u = 0; u < ADAM_SLOTS; ++u)
            if (gtid + u * GT < n4) {
                if (dist) g[u] = adam_sum4(a, gtid + u * GT, g[u]);
                sq = fmaf(g[u].x * gs, g[u].x * gs, sq); sq = fmaf(g[u].y * gs, g[u].y * gs, sq);
                sq = fmaf(g[u].z * gs, g[u].z * gs, sq); sq = fmaf(g[u].w * gs, g[u].w * gs, sq);
            }
    } else {
        for (int64_t j = gtid; j < n4; j += GT) {
            float4 x = __ldcg(reinterpret_cast<const float4*>(G) + j);
            if (dist) x = adam_sum4(a, j, x);
            sq = fmaf(x.x * gs, x.x * gs, sq); sq = fmaf(x.y * gs, x.y * gs, sq);
            sq = fmaf(x.z * gs, x.z * gs, sq); sq = fmaf(x.w * gs, x.w * gs, sq);
        }
    }
    for (int64_t i = 4 * n4 + gtid; i < a.n; i += GT) {
        float x = __ldcg(G + i);
        if (dist) x = adam_sum1(a, i, x);
        x *= gs;
        sq = fmaf(x, x, sq);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sq += __shfl_down_sync(0xffffffffu, sq, d);
    if ((tid & 31) == 0) sh[tid >> 5] = sq;
    __syncthreads();
    if (tid == 0) {
        float t = 0.f;
        for (int w = 0; w < ADAM_THREADS / 32; ++w) t += sh[w];
        s_part = t;
    }
    cluster.sync();
    if (tid == 0) {
        float t = 0.f;
        for (unsigned c = 0; c < (unsigned)ADAM_CTAS; ++c) t += *cluster.map_shared_rank(&s_part, c);
        const float norm = sqrtf(t);
        float c = 1.f;
        if (a.max_norm > 0.f) c = fminf(a.max_norm / (norm + 1e-6f), 1.f);
        const int64_t step = *a.step + 1;
        s_c[0] = c * gs;
        s_c[1] = a.lr / (float)(1.0 - ipow((double)a.beta1, step));
        s_c[2] = (float)sqrt(1.0 - ipow((double)a.beta2, step));
        if (cluster.block_rank() == 0 && a.gnorm_out) *a.gnorm_out = norm;
    }
    cluster.sync();  // also: every CTA has read *a.step and its peers' partials before anything below changes them
    if (gtid == 0) *a.step += 1;
    const float coef = s_c[0], step_size = s_c[1], bc2_sqrt = s_c[2];
    const ppo::Offsets o = ppo::offsets(a.in_dim, 64, a.out_dim);
    if (in_regs) {
#pragma unroll
        for (int u = 0; u < ADAM_SLOTS; ++u) {
            const int64_t j = gtid + u * GT;
            if (j < n4) {
                float4 q;
                q.x = adam_elem(p[u].x, g[u].x, m[u].x, v[u].x, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
                q.y = adam_elem(p[u].y, g[u].y, m[u].y, v[u].y, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
                q.z = adam_elem(p[u].z, g[u].z, m[u].z, v[u].z, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
                q.w = adam_elem(p[u].w, g[u].w, m[u].w, v[u].w, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
                reinterpret_cast<float4*>(P)[j] = q;
                reinterpret_cast<float4*>(M)[j] = m[u];
                reinterpret_cast<float4*>(V)[j] = v[u];
                reinterpret_cast<float4*>(G)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (a.image) {
                    adam_store_image(a, o, 4 * j + 0, q.x); adam_store_image(a, o, 4 * j + 1, q.y);
                    adam_store_image(a, o, 4 * j + 2, q.z); adam_store_image(a, o, 4 * j + 3, q.w);
                }
            }
        }
    } else {
        for (int64_t i = gtid; i < 4 * n4; i += GT) {
            float mm = M[i], vv = V[i], gi = __ldcg(G + i);
            if (dist) gi = adam_sum1(a, i, gi);
            const float pn = adam_elem(P[i], gi, mm, vv, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
            P[i] = pn; M[i] = mm; V[i] = vv; G[i] = 0.f;
            if (a.image) adam_store_image(a, o, i, pn);
        }
    }
    for (int64_t i = 4 * n4 + gtid; i < a.n; i += GT) {
        float mm = M[i], vv = V[i], gi = __ldcg(G + i);
        if (dist) gi = adam_sum1(a, i, gi);
        const float pn = adam_elem(P[i], gi, mm, vv, coef, step_size, bc2_sqrt, a.beta1, a.beta2, a.eps);
        P[i] = pn; M[i] = mm; V[i] = vv; G[i] = 0.f;
        if (a.image) adam_store_image(a, o, i, pn);
    }
}

}  // namespace opt
}  // namespace tsm
