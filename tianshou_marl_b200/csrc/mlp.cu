// mlp.cu -- K4/K5/K6 (fp32 CUDA-core path): actor-critic MLP forward, fused PPO
// forward+loss+backward over a minibatch, fused grad-norm-clip + Adam (sm_100a).
//
// Replaces, for nets of the shape  in -> H -> H -> out  with ReLU (H in {64,128}),
//   tianshou/utils/net/common.py:90-181,286-369 (MLP / Net forward),
//   tianshou/utils/net/discrete.py:67-89,116-124 (DiscreteActor / DiscreteCritic heads),
//   tianshou/algorithm/modelfree/reinforce.py:167-192 (ProbabilisticActorPolicy.forward: dist + sample),
//   tianshou/algorithm/modelfree/ppo.py:157-161 (logp_old), :179-216 (clipped-surrogate minibatch loss),
//   tianshou/algorithm/algorithm_base.py:485-498 (Optimizer.step: zero_grad, backward, clip_grad_norm_, Adam),
//   tianshou/algorithm/optim.py:89-110 (Adam hyper-parameters).
//
// This file is the exact-fp32 path (parity mode, and the checker for the tcgen05 path in
// mlp_tc.cu).  A CTA owns a tile of TM rows; activations live in shared memory TRANSPOSED
// ([feature][row], leading dimension TM+4) so that both the layer GEMMs (k = feature) and the
// weight-gradient GEMMs (k = row) read 16-byte vectors without bank conflicts.  Weight
// gradients accumulate in registers across all tiles of a (persistent) CTA and are added to
// the flat gradient buffer once per CTA.
//
// Flat parameter layout (float32, torch.nn.Linear convention weight[out][in]):
//   W1[H][IN] b1[H] W2[H][H] b2[H] W3[OUT][H] b3[OUT]
// For the shared actor-critic OUT = n_act + 1: rows 0..n_act-1 of W3 are the
// actor head (DiscreteActor.last), row n_act is the critic head (DiscreteCritic.last).
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "philox.cuh"
#include "feistel.cuh"
#include "ppo_math.cuh"
#include "adam.cuh"

namespace tsm {
namespace mlp {

using namespace tsm::ppo;

constexpr int THREADS = 256;

template <int H>
struct Cfg {
    static constexpr int TM = (H == 64) ? 128 : 64;  // rows per tile
    static constexpr int LD = TM + 4;                // leading dim of transposed activation tiles
    static constexpr int LDW = H + 4;                // leading dim of W2 / W3 (torch layout) in smem
    static constexpr int RC = TM / 64;               // row chunks of 4 per thread in the layer GEMMs
    static constexpr int CC = H / 64;                // col chunks of 4 per thread
};

template <int H>
struct Smem {
    // all offsets in floats
    int xrows;  // 16 * ceil(in/16): rows of sXT (zero beyond in_dim)
    int w1t, w2, w3, b1, b2, b3, xt, h1t, h2t, dot, slot, red, total;
    __host__ __device__ Smem(int in_dim) {
        using C = Cfg<H>;
        xrows = (in_dim + 15) / 16 * 16;
        int p = 0;
        w1t = p; p += xrows * H;
        w2 = p;  p += H * C::LDW;
        w3 = p;  p += OUTP * C::LDW;
        b1 = p;  p += H;
        b2 = p;  p += H;
        b3 = p;  p += OUTP;
        xt = p;  p += xrows * C::LD;
        h1t = p; p += H * C::LD;
        h2t = p; p += H * C::LD;
        dot = p; p += OUTP * C::LD;      // head outputs [TM][OUTP] (forward) / dOut^T [OUTP][LD] (backward)
        slot = p; p += 2 * C::TM;        // int64 storage slot per tile row
        red = p; p += 64;
        total = p;
    }
};

// ---------------------------------------------------------------------------------------------
template <int H>
__device__ __forceinline__ void load_weights(float* smem, const Smem<H>& L, const float* __restrict__ params, int in_dim,
                                             int out_dim) {
    using C = Cfg<H>;
    const Offsets o = offsets(in_dim, H, out_dim);
    for (int i = threadIdx.x; i < L.xrows * H; i += THREADS) {  // sW1T[k][j] = W1[j][k]
        const int k = i / H, j = i - k * H;
        smem[L.w1t + i] = k < in_dim ? params[o.w1 + (int64_t)j * in_dim + k] : 0.f;
    }
    for (int i = threadIdx.x; i < H * H; i += THREADS) {
        const int r = i / H, c = i - r * H;
        smem[L.w2 + r * C::LDW + c] = params[o.w2 + i];
    }
    for (int i = threadIdx.x; i < OUTP * H; i += THREADS) {
        const int r = i / H, c = i - r * H;
        smem[L.w3 + r * C::LDW + c] = r < out_dim ? params[o.w3 + (int64_t)r * H + c] : 0.f;
    }
    for (int i = threadIdx.x; i < H; i += THREADS) {
        smem[L.b1 + i] = params[o.b1 + i];
        smem[L.b2 + i] = params[o.b2 + i];
    }
    if (threadIdx.x < OUTP) smem[L.b3 + threadIdx.x] = threadIdx.x < out_dim ? params[o.b3 + threadIdx.x] : 0.f;
}

// gather TM rows of x (row-major [.][in_dim]) into sXT[k][row]; rows >= n_valid are zero
template <int H>
__device__ __forceinline__ void load_x_tile(float* smem, const Smem<H>& L, const float* __restrict__ x, int in_dim,
                                            int n_valid) {
    using C = Cfg<H>;
    const int64_t* sSlot = reinterpret_cast<const int64_t*>(smem + L.slot);
    float* sXT = smem + L.xt;
    if ((in_dim & 1) == 0 && (reinterpret_cast<uintptr_t>(x) & 7) == 0) {
        const int nv = in_dim >> 1;
        for (int i = threadIdx.x; i < C::TM * nv; i += THREADS) {
            const int r = i / nv, v = i - r * nv;
            float2 val = make_float2(0.f, 0.f);
            if (r < n_valid) val = *reinterpret_cast<const float2*>(x + sSlot[r] * in_dim + 2 * v);
            sXT[(2 * v) * C::LD + r] = val.x;
            sXT[(2 * v + 1) * C::LD + r] = val.y;
        }
    } else {
        for (int i = threadIdx.x; i < C::TM * in_dim; i += THREADS) {
            const int r = i / in_dim, k = i - r * in_dim;
            sXT[k * C::LD + r] = r < n_valid ? x[sSlot[r] * in_dim + k] : 0.f;
        }
    }
}

// acc[rc][cc][i][j] = sum_k A^T[k][row] * W(k, col);  row = rc*64 + rg*4 + i, col = cc*64 + cg*4 + j
// WT: W stored [k][col] (ld = ldw, 16-byte vector loads); else W stored [col][k] (torch layout, scalar loads)
template <int H, bool WT>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ sAT, const float* __restrict__ sW, int ldw, int K,
                                          float (&acc)[Cfg<H>::RC][Cfg<H>::CC][4][4]) {
    using C = Cfg<H>;
    const int rg = threadIdx.x & 15, cg = threadIdx.x >> 4;
#pragma unroll
    for (int rc = 0; rc < C::RC; ++rc)
#pragma unroll
        for (int cc = 0; cc < C::CC; ++cc)
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[rc][cc][i][j] = 0.f;
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
        float a[C::RC][4], w[C::CC][4];
#pragma unroll
        for (int rc = 0; rc < C::RC; ++rc) {
            const float4 t = *reinterpret_cast<const float4*>(sAT + k * C::LD + rc * 64 + rg * 4);
            a[rc][0] = t.x; a[rc][1] = t.y; a[rc][2] = t.z; a[rc][3] = t.w;
        }
#pragma unroll
        for (int cc = 0; cc < C::CC; ++cc) {
            if constexpr (WT) {
                const float4 t = *reinterpret_cast<const float4*>(sW + k * ldw + cc * 64 + cg * 4);
                w[cc][0] = t.x; w[cc][1] = t.y; w[cc][2] = t.z; w[cc][3] = t.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) w[cc][j] = sW[(cc * 64 + cg * 4 + j) * ldw + k];
            }
        }
#pragma unroll
        for (int rc = 0; rc < C::RC; ++rc)
#pragma unroll
            for (int cc = 0; cc < C::CC; ++cc)
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[rc][cc][i][j] = fmaf(a[rc][i], w[cc][j], acc[rc][cc][i][j]);
    }
}

// out^T[col][row] = relu(acc + bias[col])
template <int H>
__device__ __forceinline__ void store_relu(const float (&acc)[Cfg<H>::RC][Cfg<H>::CC][4][4], const float* __restrict__ sBias,
                                           float* __restrict__ sOT) {
    using C = Cfg<H>;
    const int rg = threadIdx.x & 15, cg = threadIdx.x >> 4;
#pragma unroll
    for (int rc = 0; rc < C::RC; ++rc)
#pragma unroll
        for (int cc = 0; cc < C::CC; ++cc)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = cc * 64 + cg * 4 + j;
                const float b = sBias[col];
                float4 v;
                v.x = fmaxf(acc[rc][cc][0][j] + b, 0.f);
                v.y = fmaxf(acc[rc][cc][1][j] + b, 0.f);
                v.z = fmaxf(acc[rc][cc][2][j] + b, 0.f);
                v.w = fmaxf(acc[rc][cc][3][j] + b, 0.f);
                *reinterpret_cast<float4*>(sOT + col * C::LD + rc * 64 + rg * 4) = v;
            }
}

// in place: sHT[col][row] = (sHT[col][row] > 0) ? acc : 0     (ReLU backward on the stored activation)
template <int H>
__device__ __forceinline__ void store_masked(const float (&acc)[Cfg<H>::RC][Cfg<H>::CC][4][4], float* __restrict__ sHT) {
    using C = Cfg<H>;
    const int rg = threadIdx.x & 15, cg = threadIdx.x >> 4;
#pragma unroll
    for (int rc = 0; rc < C::RC; ++rc)
#pragma unroll
        for (int cc = 0; cc < C::CC; ++cc)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4* p = reinterpret_cast<float4*>(sHT + (cc * 64 + cg * 4 + j) * C::LD + rc * 64 + rg * 4);
                const float4 h = *p;
                float4 v;
                v.x = h.x > 0.f ? acc[rc][cc][0][j] : 0.f;
                v.y = h.y > 0.f ? acc[rc][cc][1][j] : 0.f;
                v.z = h.z > 0.f ? acc[rc][cc][2][j] : 0.f;
                v.w = h.w > 0.f ? acc[rc][cc][3][j] : 0.f;
                *p = v;
            }
}

// head: sOut[row][o] = b3[o] + sum_k H2^T[k][row] * W3[o][k]
template <int H>
__device__ __forceinline__ void head_forward(const float* __restrict__ smem, const Smem<H>& L, float* __restrict__ sOut) {
    using C = Cfg<H>;
    constexpr int OPT = OUTP * C::TM / THREADS;  // outputs per thread (4 at TM=128, 2 at TM=64)
    const int row = threadIdx.x % C::TM, og = threadIdx.x / C::TM;
    float o[OPT];
#pragma unroll
    for (int q = 0; q < OPT; ++q) o[q] = smem[L.b3 + og * OPT + q];
    const float* sH2T = smem + L.h2t;
    const float* sW3 = smem + L.w3;
#pragma unroll 4
    for (int k = 0; k < H; ++k) {
        const float a = sH2T[k * C::LD + row];
#pragma unroll
        for (int q = 0; q < OPT; ++q) o[q] = fmaf(a, sW3[(og * OPT + q) * C::LDW + k], o[q]);
    }
#pragma unroll
    for (int q = 0; q < OPT; ++q) sOut[row * OUTP + og * OPT + q] = o[q];
}

// the two hidden layers + head for the tile currently in sXT
template <int H>
__device__ __forceinline__ void forward_tile(float* smem, const Smem<H>& L, int in_dim) {
    using C = Cfg<H>;
    float acc[C::RC][C::CC][4][4];
    tile_gemm<H, true>(smem + L.xt, smem + L.w1t, H, in_dim, acc);
    store_relu<H>(acc, smem + L.b1, smem + L.h1t);
    __syncthreads();
    tile_gemm<H, false>(smem + L.h1t, smem + L.w2, C::LDW, H, acc);
    store_relu<H>(acc, smem + L.b2, smem + L.h2t);
    __syncthreads();
    head_forward<H>(smem, L, smem + L.dot);
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// forward kernel.  mode: 0 raw outputs [n][out_dim]; 1 sample action (Philox inverse-CDF);
// 2 arg-max action (dist.mode, deterministic_eval); 3 evaluate given actions (log_prob).
struct FwdArgs {
    const float* params;
    const float* x;          // [.][in_dim] rows addressed through `rows` (NULL = identity)
    const int64_t* rows;     // optional gather list of x rows (also the scatter target when scatter != 0)
    int64_t n;
    int in_dim, out_dim, n_act, value_index, mode, scatter, probs;
    uint64_t seed, counter;
    int64_t rng_row0;        // added to the row index in the sampling key
    float* out;              // mode 0: [n][out_dim]
    int32_t* act;            // modes 1,2: written; mode 3: read
    float* logp;             // [n]
    float* value;            // [n] (if value_index >= 0)
    float* entropy;          // [n] optional
};

template <int H>
__global__ void __launch_bounds__(THREADS) forward_kernel(FwdArgs a) {
    using C = Cfg<H>;
    extern __shared__ __align__(16) float smem[];
    const Smem<H> L(a.in_dim);
    load_weights<H>(smem, L, a.params, a.in_dim, a.out_dim);
    int64_t* sSlot = reinterpret_cast<int64_t*>(smem + L.slot);
    const int64_t n_tiles = (a.n + C::TM - 1) / C::TM;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t base = tile * C::TM;
        const int n_valid = (int)min((int64_t)C::TM, a.n - base);
        __syncthreads();  // previous tile fully consumed (and weights visible on the first pass)
        if (threadIdx.x < C::TM) sSlot[threadIdx.x] = threadIdx.x < n_valid ? (a.rows ? a.rows[base + threadIdx.x] : base + threadIdx.x) : 0;
        __syncthreads();
        load_x_tile<H>(smem, L, a.x, a.in_dim, n_valid);
        __syncthreads();
        forward_tile<H>(smem, L, a.in_dim);
        if (threadIdx.x < n_valid) {
            const int r = threadIdx.x;
            const int64_t dst = (a.scatter && a.rows) ? sSlot[r] : base + r;
            float z[OUTP];
            const float4 z0 = *reinterpret_cast<const float4*>(smem + L.dot + r * OUTP);
            const float4 z1 = *reinterpret_cast<const float4*>(smem + L.dot + r * OUTP + 4);
            z[0] = z0.x; z[1] = z0.y; z[2] = z0.z; z[3] = z0.w; z[4] = z1.x; z[5] = z1.y; z[6] = z1.z; z[7] = z1.w;
            if (a.value && a.value_index >= 0) {
                float v = 0.f;
#pragma unroll
                for (int j = 0; j < OUTP; ++j)
                    if (j == a.value_index) v = z[j];
                a.value[dst] = v;
            }
            if (a.mode == 0) {
                if (a.out) {
#pragma unroll
                    for (int j = 0; j < OUTP; ++j)
                        if (j < a.out_dim) a.out[dst * a.out_dim + j] = z[j];
                }
            } else {
                const Soft s = softmax_row(z, a.n_act, a.probs != 0);
                int act = 0;
                if (a.mode == 1) {
                    const int64_t key = dst + a.rng_row0;
                    const Philox4 rnd = philox4x32((uint32_t)key, (uint32_t)((uint64_t)key >> 32), (uint32_t)a.counter,
                                                   (uint32_t)(a.counter >> 32), (uint32_t)a.seed, (uint32_t)(a.seed >> 32) ^ STREAM_ACT);
                    const float u = u01(rnd.x);
                    float c = 0.f;
                    act = a.n_act - 1;
                    bool found = false;
#pragma unroll
                    for (int j = 0; j < OUTP; ++j) {
                        if (j < a.n_act) {
                            c += s.p[j];
                            if (!found && u < c) { act = j; found = true; }
                        }
                    }
                } else if (a.mode == 2) {
                    float best = -INFINITY;
#pragma unroll
                    for (int j = 0; j < OUTP; ++j)
                        if (j < a.n_act && z[j] > best) { best = z[j]; act = j; }
                } else {
                    act = a.act[dst];
                }
                float lp = 0.f;
#pragma unroll
                for (int j = 0; j < OUTP; ++j)
                    if (j == act) lp = s.logp[j];
                if (a.mode != 3 && a.act) a.act[dst] = act;
                if (a.logp) a.logp[dst] = lp;
                if (a.entropy) a.entropy[dst] = s.ent;
                if (a.out) {
#pragma unroll
                    for (int j = 0; j < OUTP; ++j)
                        if (j < a.n_act) a.out[dst * a.n_act + j] = z[j];
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// G[i][j] += sum_k P^T[i][k] * Q^T[j][k] over the TM rows of the tile; i = ig + 16*ii, j = jg + 16*jj
template <int H, int NI, int NJ>
__device__ __forceinline__ void wgrad(const float* __restrict__ sPT, const float* __restrict__ sQT, float (&g)[NI][NJ]) {
    using C = Cfg<H>;
    const int ig = threadIdx.x & 15, jg = threadIdx.x >> 4;
#pragma unroll 2
    for (int k = 0; k < C::TM; k += 4) {
        float4 p[NI], q[NJ];
#pragma unroll
        for (int ii = 0; ii < NI; ++ii) p[ii] = *reinterpret_cast<const float4*>(sPT + (ig + 16 * ii) * C::LD + k);
#pragma unroll
        for (int jj = 0; jj < NJ; ++jj) q[jj] = *reinterpret_cast<const float4*>(sQT + (jg + 16 * jj) * C::LD + k);
#pragma unroll
        for (int ii = 0; ii < NI; ++ii)
#pragma unroll
            for (int jj = 0; jj < NJ; ++jj) {
                float s = g[ii][jj];
                s = fmaf(p[ii].x, q[jj].x, s);
                s = fmaf(p[ii].y, q[jj].y, s);
                s = fmaf(p[ii].z, q[jj].z, s);
                s = fmaf(p[ii].w, q[jj].w, s);
                g[ii][jj] = s;
            }
    }
}

struct PpoArgs {
    const float* params;
    float* grads;           // flat, accumulated with atomics
    const float* obs;       // time-major [size][n_rows][in_dim]
    const int32_t* act;     // [size][n_rows]
    const float* act_f;     // Gaussian nets: [size][n_rows][n_act]
    int n_sigma;            // > 0: Gaussian policy, log sigma at params[offsets.total ...]
    float mu_bound;         // Gaussian policy: > 0 = mean is mu_bound * tanh(output)
    int cond_sigma;         // Gaussian policy with per-row log sigma in head outputs [n_act, 2 n_act)
    const float* adv;
    const float* ret;
    const float* logp_old;
    const float* v_old;     // only for value_clip
    BatchView view;
    int64_t mb_start, mb_len;  // batch positions [mb_start, mb_start + mb_len)
    const float* mb_stats;     // {mean, 1/(std+eps)} of adv over this minibatch (device) or NULL
    PpoCfg cfg;
    int in_dim, out_dim, n_act, value_index;
    double* loss_sums;         // [4]: sum clip_loss_i, sum vf_i, sum ent_i, count  (device, accumulated)
};

template <int H, int XJ>
__global__ void __launch_bounds__(THREADS) ppo_minibatch_kernel(PpoArgs a) {
    using C = Cfg<H>;
    extern __shared__ __align__(16) float smem[];
    const Smem<H> L(a.in_dim);
    load_weights<H>(smem, L, a.params, a.in_dim, a.out_dim);
    int64_t* sSlot = reinterpret_cast<int64_t*>(smem + L.slot);
    float* sDOT = smem + L.dot;
    for (int i = threadIdx.x; i < (L.xrows - a.in_dim) * C::LD; i += THREADS) smem[L.xt + a.in_dim * C::LD + i] = 0.f;
    constexpr int NI = H / 16;
    float g2[NI][NI], g1[NI][XJ];
    float g3[H / 32];
    float gb1 = 0.f, gb2 = 0.f, gb3 = 0.f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
#pragma unroll
        for (int j = 0; j < NI; ++j) g2[i][j] = 0.f;
#pragma unroll
        for (int j = 0; j < XJ; ++j) g1[i][j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < H / 32; ++j) g3[j] = 0.f;
    double l_clip = 0.0, l_vf = 0.0, l_ent = 0.0;
    const float inv_n = 1.0f / (float)a.mb_len;
    float mean = 0.f, rstd = 1.f;
    if ((a.cfg.adv_norm || a.cfg.plain_pg == 2) && a.mb_stats) { mean = a.mb_stats[0]; rstd = a.mb_stats[1]; }
    const int64_t sigma_off = offsets(a.in_dim, H, a.out_dim).total;  // Gaussian nets: log sigma sits behind b3
    float dls[OUTP];  // Gaussian nets: this thread's share of d(mean loss) / d log sigma
#pragma unroll
    for (int j = 0; j < OUTP; ++j) dls[j] = 0.f;

    const int64_t n_tiles = (a.mb_len + C::TM - 1) / C::TM;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t base = a.mb_start + tile * C::TM;
        const int n_valid = (int)min((int64_t)C::TM, a.mb_start + a.mb_len - base);
        __syncthreads();
        if (threadIdx.x < C::TM) sSlot[threadIdx.x] = threadIdx.x < n_valid ? slot_of(a.view, base + threadIdx.x) : 0;
        __syncthreads();
        load_x_tile<H>(smem, L, a.obs, a.in_dim, n_valid);
        __syncthreads();
        forward_tile<H>(smem, L, a.in_dim);
        // ---- per-row loss and d(loss)/d(head outputs) (ppo.py:181-211) ----
        float dz[OUTP];
#pragma unroll
        for (int j = 0; j < OUTP; ++j) dz[j] = 0.f;
        if (threadIdx.x < n_valid) {
            const int r = threadIdx.x;
            const int64_t s = sSlot[r];
            float z[OUTP];
            const float4 z0 = *reinterpret_cast<const float4*>(sDOT + r * OUTP);
            const float4 z1 = *reinterpret_cast<const float4*>(sDOT + r * OUTP + 4);
            z[0] = z0.x; z[1] = z0.y; z[2] = z0.z; z[3] = z0.w; z[4] = z1.x; z[5] = z1.y; z[6] = z1.z; z[7] = z1.w;
            // the loss itself is shared with the tcgen05 engine (ppo_math.cuh); IEEE exp / log / division here
            const int act = (a.n_act > 0 && !a.n_sigma && !a.cond_sigma) ? a.act[s] : 0;
            const float adv_s = a.n_act > 0 ? a.adv[s] : 0.f, lpo = a.n_act > 0 ? a.logp_old[s] : 0.f;
            const float ret_s = a.value_index >= 0 ? a.ret[s] : 0.f;
            const float vo = (a.value_index >= 0 && a.cfg.value_clip) ? a.v_old[s] : 0.f;
            const RowLoss rl = (a.n_sigma || a.cond_sigma) ? row_loss_gauss(z, a.n_act, a.value_index, a.act_f + s * a.n_act, a.params + sigma_off, adv_s, ret_s, lpo, vo,
                                                          mean, rstd, a.cfg, inv_n, dz, dls, a.mu_bound, a.cond_sigma != 0)
                                         : row_loss<false>(z, a.n_act, a.value_index, act, adv_s, ret_s, lpo, vo, mean, rstd, a.cfg, inv_n, dz);
            l_clip += (double)rl.clip;
            l_ent += (double)rl.ent;
            l_vf += (double)rl.vf;
        }
        __syncthreads();  // every thread has read its head outputs before the region is reused as dOut^T
        if (threadIdx.x < C::TM) {
#pragma unroll
            for (int j = 0; j < OUTP; ++j) sDOT[j * C::LD + threadIdx.x] = dz[j];
        }
        __syncthreads();
        // ---- head gradients: dW3[o][j] += sum_row dOut[row][o] * H2[row][j];  db3[o] ----
        {
            const int o = threadIdx.x & 7, jq = threadIdx.x >> 3;
            float b3 = 0.f;
#pragma unroll 2
            for (int k = 0; k < C::TM; k += 4) {
                const float4 p = *reinterpret_cast<const float4*>(sDOT + o * C::LD + k);
#pragma unroll
                for (int jj = 0; jj < H / 32; ++jj) {
                    const float4 q = *reinterpret_cast<const float4*>(smem + L.h2t + (jq + 32 * jj) * C::LD + k);
                    g3[jj] = fmaf(p.x, q.x, fmaf(p.y, q.y, fmaf(p.z, q.z, fmaf(p.w, q.w, g3[jj]))));
                }
                b3 += (p.x + p.y) + (p.z + p.w);
            }
            if (jq == 0) gb3 += b3;
        }
        __syncthreads();
        // ---- D2 = relu'(H2) * (dOut . W3), in place over H2 ----
        float acc[C::RC][C::CC][4][4];
        tile_gemm<H, true>(sDOT, smem + L.w3, C::LDW, OUTP, acc);
        store_masked<H>(acc, smem + L.h2t);
        __syncthreads();
        // ---- dW2[i][j] += sum_row D2[row][i] * H1[row][j];  db2[i] ----
        wgrad<H, NI, NI>(smem + L.h2t, smem + L.h1t, g2);
        if (threadIdx.x < H) {
            float s = 0.f;
            for (int k = 0; k < C::TM; k += 4) {
                const float4 p = *reinterpret_cast<const float4*>(smem + L.h2t + threadIdx.x * C::LD + k);
                s += (p.x + p.y) + (p.z + p.w);
            }
            gb2 += s;
        }
        __syncthreads();
        // ---- D1 = relu'(H1) * (D2 . W2), in place over H1 ----
        tile_gemm<H, true>(smem + L.h2t, smem + L.w2, C::LDW, H, acc);
        store_masked<H>(acc, smem + L.h1t);
        __syncthreads();
        // ---- dW1[i][j] += sum_row D1[row][i] * X[row][j];  db1[i] ----
        wgrad<H, NI, XJ>(smem + L.h1t, smem + L.xt, g1);
        if (threadIdx.x < H) {
            float s = 0.f;
            for (int k = 0; k < C::TM; k += 4) {
                const float4 p = *reinterpret_cast<const float4*>(smem + L.h1t + threadIdx.x * C::LD + k);
                s += (p.x + p.y) + (p.z + p.w);
            }
            gb1 += s;
        }
    }
    // ---- flush this CTA's gradient partials ----
    const Offsets o = offsets(a.in_dim, H, a.out_dim);
    {
        const int ig = threadIdx.x & 15, jg = threadIdx.x >> 4;
#pragma unroll
        for (int ii = 0; ii < NI; ++ii) {
            const int i = ig + 16 * ii;
#pragma unroll
            for (int jj = 0; jj < NI; ++jj) atomicAdd(a.grads + o.w2 + (int64_t)i * H + jg + 16 * jj, g2[ii][jj]);
#pragma unroll
            for (int jj = 0; jj < XJ; ++jj) {
                const int j = jg + 16 * jj;
                if (j < a.in_dim) atomicAdd(a.grads + o.w1 + (int64_t)i * a.in_dim + j, g1[ii][jj]);
            }
        }
        const int oo = threadIdx.x & 7, jq = threadIdx.x >> 3;
        if (oo < a.out_dim) {
#pragma unroll
            for (int jj = 0; jj < H / 32; ++jj) atomicAdd(a.grads + o.w3 + (int64_t)oo * H + jq + 32 * jj, g3[jj]);
            if (jq == 0) atomicAdd(a.grads + o.b3 + oo, gb3);
        }
        if (threadIdx.x < H) {
            atomicAdd(a.grads + o.b1 + threadIdx.x, gb1);
            atomicAdd(a.grads + o.b2 + threadIdx.x, gb2);
        }
    }
    if (a.n_sigma) {  // log-sigma gradients: warp shuffle, one atomic per warp and action dimension
#pragma unroll
        for (int j = 0; j < OUTP; ++j) {
            float v = dls[j];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if ((threadIdx.x & 31) == 0 && j < a.n_sigma && v != 0.f) atomicAdd(a.grads + sigma_off + j, v);
        }
    }
    // ---- loss sums: warp shuffle -> shared -> one atomic per CTA ----
    if (a.loss_sums) {
        __syncthreads();
        double* sRed = reinterpret_cast<double*>(smem + L.slot);  // reuse (TM int64 >= 8*3 doubles)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            l_clip += __shfl_down_sync(0xffffffffu, l_clip, d);
            l_vf += __shfl_down_sync(0xffffffffu, l_vf, d);
            l_ent += __shfl_down_sync(0xffffffffu, l_ent, d);
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) { sRed[warp * 3] = l_clip; sRed[warp * 3 + 1] = l_vf; sRed[warp * 3 + 2] = l_ent; }
        __syncthreads();
        if (threadIdx.x < 3) {
            double s = 0.0;
            for (int w = 0; w < THREADS / 32; ++w) s += sRed[w * 3 + threadIdx.x];
            atomicAdd(a.loss_sums + threadIdx.x, s);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// mean and 1/(std_unbiased + eps) of adv over each minibatch (ppo.py:184-186), one CTA each.  The reads are a random
// gather (DRAM + page-walk latency of microseconds): 1024 threads x 8 independent loads in flight per thread.
constexpr int STATS_THREADS = 1024;
constexpr int STATS_UNROLL = 8;
__global__ void __launch_bounds__(STATS_THREADS) adv_stats_kernel(const float* __restrict__ adv, BatchView view,
                                                                 const int64_t* __restrict__ mb_bounds, float eps,
                                                                 float* __restrict__ stats, int32_t* __restrict__ slots) {
    const int64_t s0 = mb_bounds[blockIdx.x], s1 = mb_bounds[blockIdx.x + 1];
    double sum = 0.0, sq = 0.0;
    for (int64_t p0 = s0 + threadIdx.x; p0 < s1; p0 += (int64_t)STATS_THREADS * STATS_UNROLL) {
        float v[STATS_UNROLL];
#pragma unroll
        for (int u = 0; u < STATS_UNROLL; ++u) {
            const int64_t p = p0 + (int64_t)u * STATS_THREADS;
            v[u] = 0.f;
            if (p < s1) {
                const int64_t slot = slot_of(view, p);
                v[u] = __ldg(adv + slot);
                if (slots) slots[p] = (int32_t)slot;  // the slot table as a by-product (tsm_ppo_batch.slots)
            }
        }
#pragma unroll
        for (int u = 0; u < STATS_UNROLL; ++u) {
            const double x = (double)v[u];
            sum += x;
            sq += x * x;
        }
    }
    __shared__ double sh[2][STATS_THREADS / 32];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum += __shfl_down_sync(0xffffffffu, sum, d);
        sq += __shfl_down_sync(0xffffffffu, sq, d);
    }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = sum; sh[1][threadIdx.x >> 5] = sq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < STATS_THREADS / 32; ++w) { a += sh[0][w]; b += sh[1][w]; }
        const double n = (double)(s1 - s0);
        const double mean = n > 0 ? a / n : 0.0;
        double var = n > 1 ? (b - n * mean * mean) / (n - 1.0) : NAN;  // torch.std of one element is nan
        if (var < 0.0) var = 0.0;
        stats[2 * blockIdx.x] = (float)mean;
        stats[2 * blockIdx.x + 1] = (float)(1.0 / (sqrt(var) + (double)eps));
    }
}

// The same statistics for a keyed (Feistel) permutation over a lock-step view, with the cycle walk kept busy.  The network maps
// positions into [0, 2^bits) and is re-applied while the value is >= n (75 % of the domain at C2's n = 25.2 M): in the plain loop
// above a warp repeats each element until its slowest lane is inside the domain (3.4 passes on average where 1.33 are needed).
// Here every thread walks a private list of WALK_N positions: a lane that lands inside the domain stores the result and feeds its
// NEXT position into the following pass, so a pass does useful work in every lane (27 passes per 16 elements instead of 54).
// Results are identical (same permutation, same accumulation order); the gather of adv then runs over the stored slots with all
// loads in flight.
constexpr int WALK_N = 16;
__global__ void __launch_bounds__(STATS_THREADS) adv_stats_walk_kernel(const float* __restrict__ adv, BatchView view,
                                                                      const int64_t* __restrict__ mb_bounds, float eps,
                                                                      float* __restrict__ stats, int32_t* __restrict__ slots) {
    extern __shared__ uint32_t s_q[];  // [WALK_N][STATS_THREADS] permuted indices of the thread's current batch
    const int64_t s0 = mb_bounds[blockIdx.x], s1 = mb_bounds[blockIdx.x + 1];
    const FeistelPre pre = view.pre;
    const uint32_t n = (uint32_t)view.perm_n, len = (uint32_t)view.len;
    double sum = 0.0, sq = 0.0;
    for (int64_t p0 = s0 + threadIdx.x; p0 < s1; p0 += (int64_t)STATS_THREADS * WALK_N) {
        const int64_t left = (s1 - p0 + STATS_THREADS - 1) / STATS_THREADS;  // positions p0, p0 + THREADS, ... below s1
        const int n_el = left < WALK_N ? (int)left : WALK_N;
        int cur = 0;
        uint32_t x = (uint32_t)p0;
        while (cur < n_el) {
            uint32_t lo = x & pre.mask_a, hi = (x >> pre.a) & pre.mask_b;
#pragma unroll
            for (int r = 0; r < FEISTEL_ROUNDS; ++r) {
                if (r & 1) lo ^= mix32(hi ^ pre.rk[r]) & pre.mask_a;
                else hi ^= mix32(lo ^ pre.rk[r]) & pre.mask_b;
            }
            x = (hi << pre.a) | lo;
            if (x < n) {
                s_q[cur * STATS_THREADS + threadIdx.x] = x;
                ++cur;
                x = (uint32_t)(p0 + (int64_t)cur * STATS_THREADS);
            }
        }
        float v[WALK_N];
#pragma unroll
        for (int u = 0; u < WALK_N; ++u) {
            v[u] = 0.f;
            if (u < n_el) {
                const uint32_t q = s_q[u * STATS_THREADS + threadIdx.x];
                const uint32_t row = q / len, j = q - row * len;
                int64_t t = view.start + j;
                if (t >= view.size) t -= view.size;
                const int64_t slot = t * view.n_rows + row;
                v[u] = __ldg(adv + slot);
                if (slots) slots[p0 + (int64_t)u * STATS_THREADS] = (int32_t)slot;
            }
        }
#pragma unroll
        for (int u = 0; u < WALK_N; ++u) {
            const double xv = (double)v[u];
            sum += xv;
            sq += xv * xv;
        }
    }
    __shared__ double sh[2][STATS_THREADS / 32];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum += __shfl_down_sync(0xffffffffu, sum, d);
        sq += __shfl_down_sync(0xffffffffu, sq, d);
    }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = sum; sh[1][threadIdx.x >> 5] = sq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < STATS_THREADS / 32; ++w) { a += sh[0][w]; b += sh[1][w]; }
        const double nn = (double)(s1 - s0);
        const double mean = nn > 0 ? a / nn : 0.0;
        double var = nn > 1 ? (b - nn * mean * mean) / (nn - 1.0) : NAN;  // torch.std of one element is nan
        if (var < 0.0) var = 0.0;
        stats[2 * blockIdx.x] = (float)mean;
        stats[2 * blockIdx.x + 1] = (float)(1.0 / (sqrt(var) + (double)eps));
    }
}

__global__ void __launch_bounds__(256) slot_table_kernel(BatchView view, int64_t p0, int64_t n, int32_t* __restrict__ slots) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        slots[p0 + i] = (int32_t)slot_of(view, p0 + i);
}

// ---------------------------------------------------------------------------------------------
// one CTA: global grad-norm clip (torch.nn.utils.clip_grad_norm_) + torch.optim.Adam step; zeroes grads (adam.cuh)
// (one 8-CTA cluster, see adam.cuh).  PDL: lets the next minibatch kernel start its parameter-independent prologue.
__global__ void __cluster_dims__(opt::ADAM_CTAS, 1, 1) __launch_bounds__(opt::ADAM_THREADS) adam_kernel(opt::AdamArgs a) {
    pdl_launch_dependents();
    pdl_wait();
    opt::adam_cluster(a);
}
constexpr int ADAM_THREADS = 512;  // pack_image_kernel

// fp32 parameters -> fp16 operand image of the tcgen05 engine (one CTA)
__global__ void __launch_bounds__(ADAM_THREADS) pack_image_kernel(const float* __restrict__ p, int64_t n, int in_dim, int out_dim,
                                                                 __half* __restrict__ image) {
    for (int i = threadIdx.x; i < opt::IMG_HALVES * opt::IMG_COPIES; i += ADAM_THREADS) image[i] = __float2half_rn(0.f);
    __syncthreads();
    const Offsets o = offsets(in_dim, 64, out_dim);
    for (int64_t i = threadIdx.x; i < n; i += ADAM_THREADS) {
        const int idx = opt::image_index(i, o, in_dim);
        const __half h = __float2half_rn(p[i]);
        for (int c = 0; c < opt::IMG_COPIES; ++c) image[c * opt::IMG_HALVES + idx] = h;
    }
}

}  // namespace mlp
}  // namespace tsm

namespace tsm {
int tcw_net_pack(const tsm_net* net, cudaStream_t s);  // engine 2 (mlp_tcw.cu)
int tcw_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter, uint64_t seed, uint64_t counter,
                    float* out, int32_t* act, float* logp, float* value, float* entropy, int64_t rng_row0, cudaStream_t s);
int tc_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter, uint64_t seed,
                   uint64_t counter, float* out, int32_t* act, float* logp, float* value, float* entropy, int64_t rng_row0, cudaStream_t s);
int tc_ppo_minibatch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t mb_start, int64_t mb_len,
                     const float* mb_stats, double* loss_sums, int pdl, cudaStream_t s);
int tc_ppo_epoch_persistent(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t start, int64_t mb_size,
                            int64_t n_mb, int64_t total, const float* mb_stats, double* loss_sums, const opt::AdamArgs& adam,
                            cudaStream_t s);
}  // namespace tsm

using namespace tsm;
using namespace tsm::mlp;

static int check_net(const tsm_net* n, const char* fn) {
    if (!n) return fail(TSM_ERR_INVALID, "%s: net is NULL", fn);
    if (n->engine != 0 && n->engine != 1) return fail(TSM_ERR_INVALID, "%s: engine must be 0 (fp32) or 1 (tcgen05)", fn);
    if (n->hidden != 64 && n->hidden != 128) return fail(TSM_ERR_UNSUPPORTED, "%s: hidden must be 64 or 128", fn);
    if (n->in_dim < 1 || n->in_dim > 64) return fail(TSM_ERR_UNSUPPORTED, "%s: in_dim must be in 1..64 on the fp32 path", fn);
    if (n->out_dim < 1 || n->out_dim > OUTP) return fail(TSM_ERR_UNSUPPORTED, "%s: out_dim must be in 1..8", fn);
    if (n->n_act < 0 || n->n_act > n->out_dim) return fail(TSM_ERR_INVALID, "%s: 0 <= n_act <= out_dim", fn);
    if (n->value_index >= n->out_dim) return fail(TSM_ERR_INVALID, "%s: value_index < out_dim", fn);
    if (n->value_index >= 0 && n->value_index < n->n_act) return fail(TSM_ERR_INVALID, "%s: value_index must not alias a logit", fn);
    if (!n->params) return fail(TSM_ERR_INVALID, "%s: params is NULL", fn);
    if (n->engine == 1 && !n->tc_image) return fail(TSM_ERR_INVALID, "%s: engine 1 needs tc_image", fn);
    if (n->n_sigma != 0 && (n->n_sigma != n->n_act || n->engine != 0 || n->probs_param))
        return fail(TSM_ERR_INVALID, "%s: n_sigma must be 0 or n_act, on engine 0, without probs_param", fn);
    if (n->cond_sigma != 0 && (n->cond_sigma != 1 || n->n_sigma != 0 || n->engine != 0 || n->probs_param || n->n_act < 1 ||
                               n->out_dim < 2 * n->n_act || (n->value_index >= 0 && n->value_index < 2 * n->n_act)))
        return fail(TSM_ERR_INVALID, "%s: cond_sigma needs engine 0, n_sigma == 0, no probs_param, out_dim >= 2 n_act and the value head behind the sigma outputs", fn);
    if (!(n->mu_bound >= 0.f) || (n->mu_bound > 0.f && n->n_sigma == 0 && n->cond_sigma == 0))
        return fail(TSM_ERR_INVALID, "%s: mu_bound must be >= 0, and 0 unless the net is Gaussian (n_sigma > 0 or cond_sigma)", fn);
    if (n->n_params != offsets(n->in_dim, n->hidden, n->out_dim).total + n->n_sigma)
        return fail(TSM_ERR_INVALID, "%s: n_params does not match the layout (want %lld)", fn,
                    (long long)offsets(n->in_dim, n->hidden, n->out_dim).total + n->n_sigma);
    return TSM_OK;
}

template <int H>
static size_t smem_bytes(int in_dim) {
    return (size_t)Smem<H>(in_dim).total * sizeof(float);
}

template <typename K>
static int prep_kernel(K kernel, size_t smem, int* ctas_per_sm) {
    TSM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    TSM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, THREADS, smem));
    if (occ < 1) return fail(TSM_ERR_UNSUPPORTED, "%s: kernel does not fit on an SM", __func__);
    *ctas_per_sm = occ;
    return TSM_OK;
}

extern "C" int64_t tsm_net_param_count(int in_dim, int hidden, int out_dim) { return offsets(in_dim, hidden, out_dim).total; }

static int net_forward_impl(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter,
                            uint64_t seed, uint64_t counter, float* out, int32_t* act, float* logp, float* value,
                            float* entropy, int64_t rng_row0, tsm_stream_t stream) {
    if (net && net->engine == 2)
        return tcw_net_forward(net, x, rows, n, mode, scatter, seed, counter, out, act, logp, value, entropy, rng_row0, as_stream(stream));
    int rc = check_net(net, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n >= 0 && mode >= 0 && mode <= 3, "n >= 0, 0 <= mode <= 3");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(x != nullptr, "x != NULL");
    TSM_REQUIRE(mode == 0 || net->n_act > 0, "sampling / evaluation needs n_act > 0");
    TSM_REQUIRE(mode != 3 || act != nullptr, "mode 3 (evaluate) needs act");
    TSM_REQUIRE((net->n_sigma == 0 && net->cond_sigma == 0) || mode == 0, "Gaussian nets (n_sigma > 0 or cond_sigma): mode 0 only");
    if (net->engine == 1)
        return tc_net_forward(net, x, rows, n, mode, scatter, seed, counter, out, act, logp, value, entropy, rng_row0, as_stream(stream));
    FwdArgs a{};
    a.params = net->params; a.x = x; a.rows = rows; a.n = n;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim; a.n_act = net->n_act; a.value_index = net->value_index;
    a.mode = mode; a.scatter = scatter; a.seed = seed; a.counter = counter; a.rng_row0 = rng_row0; a.probs = net->probs_param;
    a.out = out; a.act = act; a.logp = logp; a.value = value; a.entropy = entropy;
    cudaStream_t s = as_stream(stream);
    int occ = 1;
    if (net->hidden == 64) {
        const size_t sm = smem_bytes<64>(net->in_dim);
        rc = prep_kernel(forward_kernel<64>, sm, &occ);
        if (rc) return rc;
        const int64_t tiles = ceil_div(n, Cfg<64>::TM);
        const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)sm_count() * occ);
        forward_kernel<64><<<grid, THREADS, sm, s>>>(a);
    } else {
        const size_t sm = smem_bytes<128>(net->in_dim);
        rc = prep_kernel(forward_kernel<128>, sm, &occ);
        if (rc) return rc;
        const int64_t tiles = ceil_div(n, Cfg<128>::TM);
        const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)sm_count() * occ);
        forward_kernel<128><<<grid, THREADS, sm, s>>>(a);
    }
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter,
                               uint64_t seed, uint64_t counter, float* out, int32_t* act, float* logp, float* value,
                               float* entropy, tsm_stream_t stream) {
    return net_forward_impl(net, x, rows, n, mode, scatter, seed, counter, out, act, logp, value, entropy, 0, stream);
}

extern "C" int tsm_net_forward_range(const tsm_net* net, const float* x, int64_t row0, int64_t n, int mode, uint64_t seed,
                                     uint64_t counter, int32_t* act, float* logp, float* value, tsm_stream_t stream) {
    TSM_REQUIRE(net != nullptr && row0 >= 0, "net != NULL, row0 >= 0");
    TSM_REQUIRE(mode == 1 || mode == 2, "mode: 1 sample, 2 argmax");
    return net_forward_impl(net, x ? x + row0 * net->in_dim : nullptr, nullptr, n, mode, 0, seed, counter, nullptr, act ? act + row0 : nullptr,
                            logp ? logp + row0 : nullptr, value ? value + row0 : nullptr, nullptr, row0, stream);
}

static BatchView make_view(const tsm_ppo_batch* b) {
    return make_batch_view(b->indices, b->size, b->n_rows, b->start, b->len, b->perm_key, b->perm_n);
}

static int check_batch(const tsm_ppo_batch* b, const char* fn) {
    if (!b) return fail(TSM_ERR_INVALID, "%s: batch is NULL", fn);
    if (b->size <= 0 || b->n_rows <= 0) return fail(TSM_ERR_INVALID, "%s: size and n_rows must be > 0", fn);
    if (!b->indices && (b->len <= 0 || b->len > b->size || b->start < 0 || b->start >= b->size))
        return fail(TSM_ERR_INVALID, "%s: need 0 < len <= size and 0 <= start < size", fn);
    if (b->perm_n < 0 || (!b->indices && b->perm_n > 0 && b->perm_n != b->len * b->n_rows))
        return fail(TSM_ERR_INVALID, "%s: perm_n must be 0 or len * n_rows", fn);
    return TSM_OK;
}

extern "C" int tsm_ppo_adv_stats(const tsm_ppo_batch* batch, const int64_t* mb_bounds, int64_t n_minibatches, float eps,
                                 float* stats_out, tsm_stream_t stream) {
    int rc = check_batch(batch, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n_minibatches >= 0, "n_minibatches >= 0");
    if (n_minibatches == 0) return TSM_OK;
    TSM_REQUIRE(batch->adv && mb_bounds && stats_out, "adv / mb_bounds / stats_out != NULL");
    // keyed permutation over a view whose indices fit 32 bits: the walk kernel (identical results)
    static const bool no_walk = getenv("TSM_ADV_STATS_PLAIN") != nullptr;
    if (!batch->indices && batch->perm_n > 0 && batch->perm_n < ((int64_t)1 << 31) && batch->len < ((int64_t)1 << 31) && !no_walk) {
        constexpr size_t smem = (size_t)WALK_N * STATS_THREADS * sizeof(uint32_t);
        static bool attr = false;
        if (!attr) {
            TSM_CUDA(cudaFuncSetAttribute(adv_stats_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr = true;
        }
        adv_stats_walk_kernel<<<(unsigned)n_minibatches, STATS_THREADS, smem, as_stream(stream)>>>(batch->adv, make_view(batch), mb_bounds, eps,
                                                                                                   stats_out, slots_of(batch));
    } else {
        adv_stats_kernel<<<(unsigned)n_minibatches, STATS_THREADS, 0, as_stream(stream)>>>(batch->adv, make_view(batch), mb_bounds, eps,
                                                                                    stats_out, slots_of(batch));
    }
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_ppo_slot_table(const tsm_ppo_batch* batch, int64_t p0, int64_t n, tsm_stream_t stream) {
    int rc = check_batch(batch, __func__);
    if (rc) return rc;
    TSM_REQUIRE(batch->slots != nullptr, "batch.slots != NULL");
    TSM_REQUIRE(batch->size * batch->n_rows < (int64_t)1 << 31, "size * n_rows < 2^31");
    TSM_REQUIRE(p0 >= 0 && n >= 0, "p0, n >= 0");
    if (n == 0) return TSM_OK;
    const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n, 256), (int64_t)sm_count() * 8);
    slot_table_kernel<<<grid, 256, 0, as_stream(stream)>>>(make_view(batch), p0, n, batch->slots);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

template <int H>
static int launch_ppo(const PpoArgs& a, cudaStream_t s) {
    const size_t sm = smem_bytes<H>(a.in_dim);
    const int xj = (a.in_dim + 15) / 16;
    int occ = 1, rc = 0;
    const int64_t tiles = ceil_div(a.mb_len, Cfg<H>::TM);
#define TSM_PPO_CASE(XJ)                                                                   \
    case XJ: {                                                                             \
        rc = prep_kernel(ppo_minibatch_kernel<H, XJ>, sm, &occ);                            \
        if (rc) return rc;                                                                 \
        const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)sm_count() * occ); \
        ppo_minibatch_kernel<H, XJ><<<grid, THREADS, sm, s>>>(a);                           \
        break;                                                                             \
    }
    switch (xj) {
        TSM_PPO_CASE(1)
        TSM_PPO_CASE(2)
        TSM_PPO_CASE(3)
        TSM_PPO_CASE(4)
        default: return fail(TSM_ERR_UNSUPPORTED, "%s: in_dim > 64", __func__);
    }
#undef TSM_PPO_CASE
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_ppo_minibatch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t mb_start,
                                 int64_t mb_len, const float* mb_stats, double* loss_sums, tsm_stream_t stream) {
    int rc = check_net(net, __func__);
    if (rc) return rc;
    rc = check_batch(batch, __func__);
    if (rc) return rc;
    TSM_REQUIRE(cfg != nullptr, "cfg != NULL");
    TSM_REQUIRE(net->grads != nullptr, "net.grads != NULL");
    TSM_REQUIRE(mb_start >= 0 && mb_len >= 0, "mb_start, mb_len >= 0");
    if (mb_len == 0) return TSM_OK;
    TSM_REQUIRE(!(cfg->adv_norm || cfg->plain_pg == 2) || mb_stats, "advantage normalisation / plain_pg 2 need mb_stats");
    if (net->engine == 1) return tc_ppo_minibatch(net, batch, cfg, mb_start, mb_len, mb_stats, loss_sums, 0, as_stream(stream));
    TSM_REQUIRE(batch->obs != nullptr, "batch.obs != NULL");
    TSM_REQUIRE(net->n_act == 0 || (((net->n_sigma || net->cond_sigma) ? (const void*)batch->act_f : (const void*)batch->act) && batch->adv && batch->logp_old),
                "policy loss needs act (act_f for Gaussian nets) / adv / logp_old");
    TSM_REQUIRE(net->value_index < 0 || batch->ret, "value loss needs ret");
    TSM_REQUIRE(!cfg->value_clip || batch->v_old, "value_clip needs v_old");
    PpoArgs a{};
    a.params = net->params; a.grads = net->grads;
    a.obs = batch->obs; a.act = batch->act; a.adv = batch->adv; a.ret = batch->ret; a.logp_old = batch->logp_old; a.v_old = batch->v_old;
    a.act_f = batch->act_f; a.n_sigma = net->n_sigma; a.mu_bound = net->mu_bound; a.cond_sigma = net->cond_sigma;
    a.view = make_view(batch);
    a.mb_start = mb_start; a.mb_len = mb_len; a.mb_stats = mb_stats;
    a.cfg.eps_clip = cfg->eps_clip; a.cfg.dual_clip = cfg->dual_clip; a.cfg.vf_coef = cfg->vf_coef; a.cfg.ent_coef = cfg->ent_coef;
    a.cfg.value_clip = cfg->value_clip; a.cfg.adv_norm = cfg->adv_norm;
    a.cfg.plain_pg = cfg->plain_pg; a.cfg.value_heads = cfg->value_heads; a.cfg.probs_param = net->probs_param;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim; a.n_act = net->n_act; a.value_index = net->value_index;
    a.loss_sums = loss_sums;
    return net->hidden == 64 ? launch_ppo<64>(a, as_stream(stream)) : launch_ppo<128>(a, as_stream(stream));
}

static opt::AdamArgs adam_args(const tsm_net* net, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                               float grad_scale, float* grad_norm_out) {
    opt::AdamArgs a{};
    a.p = net->params; a.g = net->grads; a.m = net->adam_m; a.v = net->adam_v; a.step = net->adam_step; a.n = net->n_params;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.max_norm = max_grad_norm; a.grad_scale = grad_scale;
    a.gnorm_out = grad_norm_out;
    a.image = net->engine == 1 ? reinterpret_cast<__half*>(net->tc_image) : nullptr;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim;
    a.world = 1; a.rank = 0; a.seq = 0; a.status = nullptr; a.n_stride = 0; a.debug = 0;
    return a;
}

static int check_peers(const tsm_peers* p, const char* fn) {
    if (!p) return fail(TSM_ERR_INVALID, "%s: peers is NULL", fn);
    if (p->world < 1 || p->world > TSM_MAX_RANKS || p->rank < 0 || p->rank >= p->world)
        return fail(TSM_ERR_INVALID, "%s: need 1 <= world <= %d and 0 <= rank < world", fn, TSM_MAX_RANKS);
    for (int q = 0; q < p->world; ++q)
        if (!p->recv[q] || !p->flags[q]) return fail(TSM_ERR_INVALID, "%s: peers.recv / flags of rank %d is NULL", fn, q);
    return TSM_OK;
}

static opt::AdamArgs adam_args_dist(const tsm_net* net, const tsm_peers* peers, uint32_t seq, float lr, float beta1, float beta2,
                                    float eps, float max_grad_norm, float* grad_norm_out) {
    opt::AdamArgs a = adam_args(net, lr, beta1, beta2, eps, max_grad_norm, 1.0f / (float)peers->world, grad_norm_out);
    a.world = peers->world; a.rank = peers->rank; a.seq = seq; a.status = peers->status;
    a.n_stride = (net->n_params + 3) / 4 * 4;
    static const int dbg = getenv("TSM_DIST_DEBUG") ? atoi(getenv("TSM_DIST_DEBUG")) : 0;
    a.debug = dbg;
    for (int q = 0; q < peers->world; ++q) {
        a.peer_recv[q] = peers->recv[q];
        a.peer_flag[q] = peers->flags[q];
        a.peer_ll[q] = reinterpret_cast<uint4*>(peers->ll[q]);
    }
    return a;
}

static int launch_adam(const opt::AdamArgs& a, int pdl, cudaStream_t s) {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(opt::ADAM_CTAS); lc.blockDim = dim3(opt::ADAM_THREADS); lc.dynamicSmemBytes = 0; lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl ? 1 : 0;
    TSM_CUDA(cudaLaunchKernelEx(&lc, adam_kernel, a));
    return TSM_OK;
}

extern "C" int tsm_adam_step(const tsm_net* net, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                             float grad_scale, float* grad_norm_out, tsm_stream_t stream) {
    TSM_REQUIRE(net != nullptr, "net != NULL");
    TSM_REQUIRE(net->params && net->grads && net->adam_m && net->adam_v && net->adam_step, "params/grads/adam state != NULL");
    TSM_REQUIRE(net->n_params > 0, "n_params > 0");
    TSM_REQUIRE(net->engine != 1 || net->tc_image, "engine 1 needs tc_image");
    return launch_adam(adam_args(net, lr, beta1, beta2, eps, max_grad_norm, grad_scale, grad_norm_out), 0, as_stream(stream));
}

extern "C" int tsm_net_pack(const tsm_net* net, tsm_stream_t stream) {
    if (net && net->engine == 2) return tcw_net_pack(net, as_stream(stream));
    int rc = check_net(net, __func__);
    if (rc) return rc;
    if (net->engine != 1) return TSM_OK;
    pack_image_kernel<<<1, ADAM_THREADS, 0, as_stream(stream)>>>(net->params, net->n_params, net->in_dim, net->out_dim,
                                                                 reinterpret_cast<__half*>(net->tc_image));
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_ppo_epoch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, const int64_t* h_mb_bounds,
                             int64_t n_minibatches, const float* mb_stats, float lr, float beta1, float beta2, float eps,
                             float max_grad_norm, double* loss_sums, tsm_stream_t stream) {
    TSM_REQUIRE(h_mb_bounds != nullptr && n_minibatches >= 0, "h_mb_bounds != NULL");
    if (net && net->engine == 1 && n_minibatches >= 1 && !getenv("TSM_NO_PERSISTENT")) {
        // uniform minibatches (Batch.split with merge_last: all of size bs, the last one possibly longer): one persistent
        // cooperative launch runs the whole epoch, optimizer steps included (mlp_tc_kernels.cuh: tc_ppo_epoch_kernel)
        const int64_t bs = h_mb_bounds[1] - h_mb_bounds[0];
        bool uniform = bs > 0;
        for (int64_t m = 1; m + 1 < n_minibatches && uniform; ++m) uniform = h_mb_bounds[m + 1] - h_mb_bounds[m] == bs;
        if (uniform && n_minibatches > 1) uniform = h_mb_bounds[n_minibatches] - h_mb_bounds[n_minibatches - 1] >= bs;
        if (uniform) {
            int rc = check_net(net, __func__);
            if (rc) return rc;
            rc = check_batch(batch, __func__);
            if (rc) return rc;
            TSM_REQUIRE(cfg && net->grads && net->adam_m && net->adam_v && net->adam_step, "cfg / grads / adam state != NULL");
            TSM_REQUIRE(!(cfg->adv_norm || cfg->plain_pg == 2) || mb_stats, "advantage normalisation / plain_pg 2 need mb_stats");
            rc = tc_ppo_epoch_persistent(net, batch, cfg, h_mb_bounds[0], bs, n_minibatches, h_mb_bounds[n_minibatches] - h_mb_bounds[0],
                                         mb_stats, loss_sums, adam_args(net, lr, beta1, beta2, eps, max_grad_norm, 1.0f, nullptr),
                                         as_stream(stream));
            if (rc <= 0) return rc;  // done, or a real error; rc == 1: not launchable here, fall through
        }
    }
    for (int64_t m = 0; m < n_minibatches; ++m) {
        const int64_t s0 = h_mb_bounds[m], len = h_mb_bounds[m + 1] - h_mb_bounds[m];
        const float* st = mb_stats ? mb_stats + 2 * m : nullptr;
        double* ls = loss_sums ? loss_sums + 4 * m : nullptr;
        if (net && net->engine == 1 && len > 0) {
            // programmatic dependent launches: minibatch(m) -> adam(m) -> minibatch(m+1); each dependent starts its prologue
            // (TMEM allocation, first batch gather) while its predecessor drains.  The first minibatch launches normally.
            int rc = check_net(net, __func__);
            if (rc) return rc;
            rc = check_batch(batch, __func__);
            if (rc) return rc;
            TSM_REQUIRE(cfg && net->grads && net->adam_m && net->adam_v && net->adam_step, "cfg / grads / adam state != NULL");
            TSM_REQUIRE(!(cfg->adv_norm || cfg->plain_pg == 2) || st, "advantage normalisation / plain_pg 2 need mb_stats");
            static const int no_pdl = getenv("TSM_NO_PDL") != nullptr;
            rc = tc_ppo_minibatch(net, batch, cfg, s0, len, st, ls, m > 0 && !no_pdl, as_stream(stream));
            if (rc) return rc;
            rc = launch_adam(adam_args(net, lr, beta1, beta2, eps, max_grad_norm, 1.0f, nullptr), !no_pdl, as_stream(stream));
            if (rc) return rc;
            continue;
        }
        int rc = tsm_ppo_minibatch(net, batch, cfg, s0, len, st, ls, stream);
        if (rc) return rc;
        rc = tsm_adam_step(net, lr, beta1, beta2, eps, max_grad_norm, 1.0f, nullptr, stream);
        if (rc) return rc;
    }
    return TSM_OK;
}

extern "C" int tsm_adam_step_dist(const tsm_net* net, const tsm_peers* peers, uint32_t seq, float lr, float beta1, float beta2, float eps,
                                  float max_grad_norm, float* grad_norm_out, tsm_stream_t stream) {
    TSM_REQUIRE(net != nullptr, "net != NULL");
    int rc = check_peers(peers, __func__);
    if (rc) return rc;
    TSM_REQUIRE(net->params && net->grads && net->adam_m && net->adam_v && net->adam_step && net->n_params > 0, "params / grads / adam state != NULL");
    TSM_REQUIRE(net->engine != 1 || net->tc_image, "engine 1 needs tc_image");
    TSM_REQUIRE(seq >= 1, "seq >= 1");
    return launch_adam(adam_args_dist(net, peers, seq, lr, beta1, beta2, eps, max_grad_norm, grad_norm_out), 0, as_stream(stream));
}

extern "C" int tsm_ppo_epoch_dist(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, const int64_t* h_mb_bounds,
                                  int64_t n_minibatches, const float* mb_stats, float lr, float beta1, float beta2, float eps,
                                  float max_grad_norm, double* loss_sums, const tsm_peers* peers, uint32_t seq0, tsm_stream_t stream) {
    TSM_REQUIRE(h_mb_bounds != nullptr && n_minibatches >= 0 && net != nullptr, "h_mb_bounds / net != NULL");
    int rc = check_peers(peers, __func__);
    if (rc) return rc;
    TSM_REQUIRE(seq0 >= 1, "seq0 >= 1");
    bool have_ll = true;
    for (int q = 0; q < peers->world; ++q) have_ll = have_ll && peers->ll[q] != nullptr;
    if (net->engine == 1 && n_minibatches >= 1 && have_ll && !getenv("TSM_NO_PERSISTENT")) {
        // as tsm_ppo_epoch: one persistent cooperative launch per epoch; the gradient exchange runs inside it
        const int64_t bs = h_mb_bounds[1] - h_mb_bounds[0];
        bool uniform = bs > 0;
        for (int64_t m = 1; m + 1 < n_minibatches && uniform; ++m) uniform = h_mb_bounds[m + 1] - h_mb_bounds[m] == bs;
        if (uniform && n_minibatches > 1) uniform = h_mb_bounds[n_minibatches] - h_mb_bounds[n_minibatches - 1] >= bs;
        if (uniform) {
            rc = check_net(net, __func__);
            if (rc) return rc;
            rc = check_batch(batch, __func__);
            if (rc) return rc;
            TSM_REQUIRE(cfg && net->grads && net->adam_m && net->adam_v && net->adam_step, "cfg / grads / adam state != NULL");
            TSM_REQUIRE(!(cfg->adv_norm || cfg->plain_pg == 2) || mb_stats, "advantage normalisation / plain_pg 2 need mb_stats");
            rc = tc_ppo_epoch_persistent(net, batch, cfg, h_mb_bounds[0], bs, n_minibatches, h_mb_bounds[n_minibatches] - h_mb_bounds[0],
                                         mb_stats, loss_sums, adam_args_dist(net, peers, seq0, lr, beta1, beta2, eps, max_grad_norm, nullptr),
                                         as_stream(stream));
            if (rc <= 0) return rc;
        }
    }
    for (int64_t m = 0; m < n_minibatches; ++m) {
        const uint32_t seq = seq0 + (uint32_t)m;
        const int64_t s0 = h_mb_bounds[m], len = h_mb_bounds[m + 1] - h_mb_bounds[m];
        const tsm_net& local = *net;
        const float* st = mb_stats ? mb_stats + 2 * m : nullptr;
        double* ls = loss_sums ? loss_sums + 4 * m : nullptr;
        if (local.engine == 1 && len > 0) {
            rc = check_net(&local, __func__);
            if (rc) return rc;
            rc = check_batch(batch, __func__);
            if (rc) return rc;
            TSM_REQUIRE(cfg && local.adam_m && local.adam_v && local.adam_step, "cfg / adam state != NULL");
            TSM_REQUIRE(!(cfg->adv_norm || cfg->plain_pg == 2) || st, "advantage normalisation / plain_pg 2 need mb_stats");
            rc = tc_ppo_minibatch(&local, batch, cfg, s0, len, st, ls, m > 0, as_stream(stream));
        } else {
            rc = tsm_ppo_minibatch(&local, batch, cfg, s0, len, st, ls, stream);
        }
        if (rc) return rc;
        rc = launch_adam(adam_args_dist(&local, peers, seq, lr, beta1, beta2, eps, max_grad_norm, nullptr), local.engine == 1,
                         as_stream(stream));
        if (rc) return rc;
    }
    return TSM_OK;
}
