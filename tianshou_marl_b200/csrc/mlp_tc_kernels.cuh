// mlp_tc_kernels.cuh -- device code of the tcgen05 actor-critic path (included by mlp_tc.cu only).
//
// Net: in(<=31) -> 64 -> 64 -> out(<=8).  One CTA = 128 threads = one tile of 128 rows; thread t
// owns row t (TMEM lane t).  All GEMM operands are fp16 in the chunked layout of tc.cuh with 8
// features per 16-byte chunk:  act[f/8][row 0..127][f%8],  weight[in/8][out][in%8].
//
//   forward      H1 = relu(X W1^T + b1)        A = X   (K-major, K = 32)   B = W1 (K-major)  N = 64
//                H2 = relu(H1 W2^T + b2)       A = H1  (K-major, K = 64)   B = W2 (K-major)  N = 64
//                OUT = H2 W3^T + b3            A = H2  (K-major, K = 64)   B = W3 (K-major)  N = 16
//   data grads   DH2 = G W3                    A = G   (K-major, K = 16)   B = W3 (MN-major) N = 64
//                DH1 = D2 W2                   A = D2  (K-major, K = 64)   B = W2 (MN-major) N = 64
//   weight grads (K = the 128 rows, both operands MN-major views of the activation images,
//                 M = 128 made of two adjacent 64-feature images):
//                [dW3^T; .] += [H2; H1]^T G          N = 16   (rows 0..63 used)
//                [dW2 ; .] += [D2; H1]^T H1          N = 64   (rows 0..63 used)
//                [dW1|db1 ; .|db2] += [D1; D2]^T X   N = 32   (X feature 31 is a column of ones)
// G = S * d(loss)/d(OUT) with S a power of two (fp16 range), D2 / D1 = ReLU-masked DH2 / DH1.
// Weight-gradient accumulators stay in TMEM across all tiles of the CTA and are flushed once.
#pragma once

constexpr int TC_THREADS = 128;
constexpr int TC_ROWS = 128;
constexpr int TC_H = 64;
constexpr int TC_XF = 32;    // padded input features
constexpr int TC_OF = 16;    // padded head outputs (UMMA N >= 16 at M = 128)
constexpr int CHUNK_B = TC_ROWS * 16;  // bytes of one 8-feature chunk of an activation image (2048)

// shared-memory map (bytes).  [D1][H2/D2][H1][X] must stay adjacent and in this order.
constexpr int SM_D1 = 0;
constexpr int SM_H2 = SM_D1 + 8 * CHUNK_B;   // 16 KB each
constexpr int SM_H1 = SM_H2 + 8 * CHUNK_B;
constexpr int SM_X = SM_H1 + 8 * CHUNK_B;    // 4 chunks = 8 KB
constexpr int SM_G = SM_X + 4 * CHUNK_B;     // 2 chunks = 4 KB (chunk 1 stays zero)
constexpr int SM_W1 = SM_G + 2 * CHUNK_B;    // [32/8][64][8] fp16 = 4 KB
constexpr int SM_W2 = SM_W1 + 4 * 64 * 16;   // [64/8][64][8] = 8 KB
constexpr int SM_W3 = SM_W2 + 8 * 64 * 16;   // [64/8][16][8] = 2 KB
constexpr int SM_B1 = SM_W3 + 8 * 16 * 16;   // fp32 biases
constexpr int SM_B2 = SM_B1 + 64 * 4;
constexpr int SM_B3 = SM_B2 + 64 * 4;
constexpr int SM_RED = SM_B3 + 16 * 4;       // 8 warps x 4 doubles scratch
constexpr int SM_BAR = SM_RED + 256;         // mbarrier + tmem base
constexpr int SM_TOTAL = SM_BAR + 64;
constexpr int SM_FWD_TOTAL = SM_TOTAL;       // forward kernel uses the same map (D1 / G unused)

// TMEM columns (fp32 accumulators, 128 lanes)
constexpr uint32_t TM_SCR = 0;     // 64: H1pre / H2pre / DH2 / DH1 in turn
constexpr uint32_t TM_OUT = 64;    // 16
constexpr uint32_t TM_GW2 = 96;    // 64
constexpr uint32_t TM_GW1 = 160;   // 32
constexpr uint32_t TM_GW3 = 192;   // 16
constexpr uint32_t TM_COLS_BWD = 256;
constexpr uint32_t TM_COLS_FWD = 128;

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// weights: flat fp32 torch layout -> fp16 images [in/8][out][in%8] (zero padded)
__device__ __forceinline__ void tc_load_weights(unsigned char* smem, const float* __restrict__ params, int in_dim, int out_dim) {
    const Offsets o = offsets(in_dim, TC_H, out_dim);
    __half* w1 = reinterpret_cast<__half*>(smem + SM_W1);
    __half* w2 = reinterpret_cast<__half*>(smem + SM_W2);
    __half* w3 = reinterpret_cast<__half*>(smem + SM_W3);
    for (int i = threadIdx.x; i < TC_XF * TC_H; i += TC_THREADS) {  // i = out * 32 + in
        const int out = i / TC_XF, in = i - out * TC_XF;
        const float v = in < in_dim ? params[o.w1 + (int64_t)out * in_dim + in] : 0.f;
        w1[((in >> 3) * TC_H + out) * 8 + (in & 7)] = __float2half_rn(v);
    }
    for (int i = threadIdx.x; i < TC_H * TC_H; i += TC_THREADS) {
        const int out = i / TC_H, in = i - out * TC_H;
        w2[((in >> 3) * TC_H + out) * 8 + (in & 7)] = __float2half_rn(params[o.w2 + i]);
    }
    for (int i = threadIdx.x; i < TC_OF * TC_H; i += TC_THREADS) {
        const int out = i / TC_H, in = i - out * TC_H;
        const float v = out < out_dim ? params[o.w3 + (int64_t)out * TC_H + in] : 0.f;
        w3[((in >> 3) * TC_OF + out) * 8 + (in & 7)] = __float2half_rn(v);
    }
    float* b1 = reinterpret_cast<float*>(smem + SM_B1);
    float* b2 = reinterpret_cast<float*>(smem + SM_B2);
    float* b3 = reinterpret_cast<float*>(smem + SM_B3);
    if (threadIdx.x < TC_H) {
        b1[threadIdx.x] = params[o.b1 + threadIdx.x];
        b2[threadIdx.x] = params[o.b2 + threadIdx.x];
    }
    if (threadIdx.x < TC_OF) b3[threadIdx.x] = threadIdx.x < out_dim ? params[o.b3 + threadIdx.x] : 0.f;
}

// this thread's input row (fp32 registers, feature 31 = `one`) -> X image
__device__ __forceinline__ void tc_store_x(unsigned char* smem, const float (&x)[TC_XF]) {
#pragma unroll
    for (int c = 0; c < TC_XF / 8; ++c) {
        uint4 v;
        v.x = pack_h2(x[8 * c + 0], x[8 * c + 1]);
        v.y = pack_h2(x[8 * c + 2], x[8 * c + 3]);
        v.z = pack_h2(x[8 * c + 4], x[8 * c + 5]);
        v.w = pack_h2(x[8 * c + 6], x[8 * c + 7]);
        *reinterpret_cast<uint4*>(smem + SM_X + c * CHUNK_B + threadIdx.x * 16) = v;
    }
}

// gather this thread's row of x (row-major [.][in_dim] fp32) into registers; zero when !valid
__device__ __forceinline__ void tc_load_row(const float* __restrict__ x, int64_t row, int in_dim, bool valid, float one, float (&r)[TC_XF]) {
#pragma unroll
    for (int k = 0; k < TC_XF; ++k) r[k] = 0.f;
    if (valid) {
        const float* p = x + row * in_dim;
        if ((in_dim & 1) == 0 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) {
#pragma unroll
            for (int k = 0; k < TC_XF / 2; ++k)
                if (2 * k < in_dim) {
                    const float2 v = *reinterpret_cast<const float2*>(p + 2 * k);
                    r[2 * k] = v.x;
                    r[2 * k + 1] = v.y;
                }
        } else {
#pragma unroll
            for (int k = 0; k < TC_XF; ++k)
                if (k < in_dim) r[k] = p[k];
        }
        r[TC_XF - 1] = one;
    }
}

// SCR accumulator (64 cols of this thread's lane) + bias, ReLU -> fp16 image at smem + dst
__device__ __forceinline__ void tc_epilogue_relu(unsigned char* smem, uint32_t tmem_lane_base, int bias_off, int dst) {
    const float* bias = reinterpret_cast<const float*>(smem + bias_off);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float v[16];
        tc::tmem_ld16(tmem_lane_base + TM_SCR + 16 * q, v);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = fmaxf(v[8 * h + i] + bias[16 * q + 8 * h + i], 0.f);
            uint4 w;
            w.x = pack_h2(f[0], f[1]); w.y = pack_h2(f[2], f[3]); w.z = pack_h2(f[4], f[5]); w.w = pack_h2(f[6], f[7]);
            *reinterpret_cast<uint4*>(smem + dst + (2 * q + h) * CHUNK_B + threadIdx.x * 16) = w;
        }
    }
}

// SCR accumulator masked by (act image at `mask_src` > 0) -> fp16 image at `dst` (may equal mask_src)
__device__ __forceinline__ void tc_epilogue_mask(unsigned char* smem, uint32_t tmem_lane_base, int mask_src, int dst) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float v[16];
        tc::tmem_ld16(tmem_lane_base + TM_SCR + 16 * q, v);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint4 m = *reinterpret_cast<const uint4*>(smem + mask_src + (2 * q + h) * CHUNK_B + threadIdx.x * 16);
            const __half2* mh = reinterpret_cast<const __half2*>(&m);
            float f[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 a = __half22float2(mh[i]);
                f[2 * i] = a.x > 0.f ? v[8 * h + 2 * i] : 0.f;
                f[2 * i + 1] = a.y > 0.f ? v[8 * h + 2 * i + 1] : 0.f;
            }
            uint4 w;
            w.x = pack_h2(f[0], f[1]); w.y = pack_h2(f[2], f[3]); w.z = pack_h2(f[4], f[5]); w.w = pack_h2(f[6], f[7]);
            *reinterpret_cast<uint4*>(smem + dst + (2 * q + h) * CHUNK_B + threadIdx.x * 16) = w;
        }
    }
}

// descriptor helpers for the four operand views (fp16, 8 features per chunk)
__device__ __forceinline__ uint64_t desc_act_k(uint32_t base, int kstep) {  // rows x features, K = features
    return tc::make_desc(base + kstep * 2 * CHUNK_B, CHUNK_B, 128);
}
__device__ __forceinline__ uint64_t desc_act_mn(uint32_t base, int kstep) {  // features x rows, K = rows
    return tc::make_desc(base + kstep * 256, 128, CHUNK_B);
}
__device__ __forceinline__ uint64_t desc_w_k(uint32_t base, int n_out, int kstep) {  // N = out, K = in
    return tc::make_desc(base + kstep * 2 * n_out * 16, n_out * 16, 128);
}
__device__ __forceinline__ uint64_t desc_w_mn(uint32_t base, int n_out, int kstep) {  // N = in, K = out
    return tc::make_desc(base + kstep * 256, 128, n_out * 16);
}

// publish generic-proxy smem writes + TMEM reads to the MMA-issuing thread
__device__ __forceinline__ void tc_sync_for_mma() {
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
}

// ---------------------------------------------------------------------------------------------
struct TcFwdArgs {
    const float* params;
    const float* x;
    const int64_t* rows;
    int64_t n;
    int in_dim, out_dim, n_act, value_index, mode, scatter;
    uint64_t seed, counter;
    float* out;
    int32_t* act;
    float* logp;
    float* value;
    float* entropy;
};

struct TcSetup {
    uint64_t* bar;
    uint32_t tmem, lane_base, sbase, phase;
};

__device__ __forceinline__ TcSetup tc_setup(unsigned char* smem, const float* params, int in_dim, int out_dim, uint32_t tmem_cols) {
    TcSetup s;
    s.bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    uint32_t* tmem_s = reinterpret_cast<uint32_t*>(smem + SM_BAR + 8);
    const int warp = threadIdx.x >> 5;
    tc_load_weights(smem, params, in_dim, out_dim);
    for (int i = threadIdx.x; i < 2 * CHUNK_B / 16; i += TC_THREADS)  // G image (chunk 1 stays zero forever)
        *reinterpret_cast<uint4*>(smem + SM_G + i * 16) = make_uint4(0, 0, 0, 0);
    if (warp == 0) tc::tmem_alloc(tmem_s, tmem_cols);
    if (threadIdx.x == 0) {
        tc::mbar_init(s.bar, 1);
        tc::fence_barrier_init();
    }
    tc_sync_for_mma();
    s.tmem = *tmem_s;
    s.lane_base = s.tmem + ((uint32_t)(warp * 32) << 16);
    s.sbase = tc::smem_u32(smem);
    s.phase = 0;
    return s;
}

__device__ __forceinline__ void tc_wait(TcSetup& s) {
    tc::mbar_wait(s.bar, s.phase);
    s.phase ^= 1;
    tc::fence_after_sync();
}

// X -> H1 -> H2 -> OUT for the tile currently in the X image; leaves OUT in TMEM (waited)
__device__ __forceinline__ void tc_forward_tile(unsigned char* smem, TcSetup& s) {
    constexpr uint32_t ID64 = tc::make_idesc_f16(128, 64, 0, 0, 0, 0);
    constexpr uint32_t ID16 = tc::make_idesc_f16(128, 16, 0, 0, 0, 0);
    tc_sync_for_mma();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_XF / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase + SM_X, ks), desc_w_k(s.sbase + SM_W1, TC_H, ks), ID64, ks > 0);
        tc::commit(s.bar);
    }
    tc_wait(s);
    tc_epilogue_relu(smem, s.lane_base, SM_B1, SM_H1);
    tc_sync_for_mma();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase + SM_H1, ks), desc_w_k(s.sbase + SM_W2, TC_H, ks), ID64, ks > 0);
        tc::commit(s.bar);
    }
    tc_wait(s);
    tc_epilogue_relu(smem, s.lane_base, SM_B2, SM_H2);
    tc_sync_for_mma();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_OUT, desc_act_k(s.sbase + SM_H2, ks), desc_w_k(s.sbase + SM_W3, TC_OF, ks), ID16, ks > 0);
        tc::commit(s.bar);
    }
    tc_wait(s);
}

__device__ __forceinline__ void tc_read_out(unsigned char* smem, const TcSetup& s, float (&z)[OUTP]) {
    float o[16];
    tc::tmem_ld16(s.lane_base + TM_OUT, o);
    const float* b3 = reinterpret_cast<const float*>(smem + SM_B3);
#pragma unroll
    for (int j = 0; j < OUTP; ++j) z[j] = o[j] + b3[j];
}

__global__ void __launch_bounds__(TC_THREADS) tc_forward_kernel(TcFwdArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    TcSetup s = tc_setup(smem, a.params, a.in_dim, a.out_dim, TM_COLS_FWD);
    const int64_t n_tiles = (a.n + TC_ROWS - 1) / TC_ROWS;
    const int r = threadIdx.x;
    float xr[TC_XF];
    int64_t tile = blockIdx.x;
    int64_t nsrc = 0;
    bool nvalid = false;
    if (tile < n_tiles) {
        nvalid = tile * TC_ROWS + r < a.n;
        nsrc = nvalid ? (a.rows ? a.rows[tile * TC_ROWS + r] : tile * TC_ROWS + r) : 0;
        tc_load_row(a.x, nsrc, a.in_dim, nvalid, 0.f, xr);
    }
    for (; tile < n_tiles; tile += gridDim.x) {
        const bool valid = nvalid;
        const int64_t src = nsrc;
        const int64_t dst = (a.scatter && a.rows) ? src : tile * TC_ROWS + r;
        tc_store_x(smem, xr);
        const int64_t nt = tile + gridDim.x;
        if (nt < n_tiles) {  // prefetch the next tile's row while this one computes
            nvalid = nt * TC_ROWS + r < a.n;
            nsrc = nvalid ? (a.rows ? a.rows[nt * TC_ROWS + r] : nt * TC_ROWS + r) : 0;
            tc_load_row(a.x, nsrc, a.in_dim, nvalid, 0.f, xr);
        }
        tc_forward_tile(smem, s);
        float z[OUTP];
        tc_read_out(smem, s, z);
        if (valid) {
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
                const Soft sm = softmax_row(z, a.n_act);
                int act = 0;
                if (a.mode == 1) {
                    const Philox4 rnd = philox4x32((uint32_t)dst, (uint32_t)((uint64_t)dst >> 32), (uint32_t)a.counter,
                                                   (uint32_t)(a.counter >> 32), (uint32_t)a.seed, (uint32_t)(a.seed >> 32) ^ STREAM_ACT);
                    const float u = u01(rnd.x);
                    float c = 0.f;
                    act = a.n_act - 1;
                    bool found = false;
#pragma unroll
                    for (int j = 0; j < OUTP; ++j) {
                        if (j < a.n_act) {
                            c += sm.p[j];
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
                    if (j == act) lp = sm.logp[j];
                if (a.mode != 3 && a.act) a.act[dst] = act;
                if (a.logp) a.logp[dst] = lp;
                if (a.entropy) a.entropy[dst] = sm.ent;
                if (a.out) {
#pragma unroll
                    for (int j = 0; j < OUTP; ++j)
                        if (j < a.n_act) a.out[dst * a.n_act + j] = z[j];
                }
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) tc::tmem_dealloc(s.tmem, TM_COLS_FWD);
}

// ---------------------------------------------------------------------------------------------
struct TcPpoArgs {
    const float* params;
    float* grads;
    const float* obs;
    const int32_t* act;
    const float* adv;
    const float* ret;
    const float* logp_old;
    const float* v_old;
    BatchView view;
    int64_t mb_start, mb_len;
    const float* mb_stats;
    PpoCfg cfg;
    int in_dim, out_dim, n_act, value_index;
    float gscale, inv_gscale;  // S and 1/S (power of two)
    double* loss_sums;
};

struct RowData {
    int act;
    float adv, ret, logp_old, v_old;
};

__device__ __forceinline__ void tc_prefetch_row(const TcPpoArgs& a, int64_t tile, int r, bool& valid, float (&xr)[TC_XF], RowData& d) {
    const int64_t p = a.mb_start + tile * TC_ROWS + r;
    valid = tile * TC_ROWS + r < a.mb_len;
    d = RowData{0, 0.f, 0.f, 0.f, 0.f};
    int64_t slot = 0;
    if (valid) {
        slot = slot_of(a.view, p);
        if (a.n_act > 0) {
            d.act = a.act[slot];
            d.adv = a.adv[slot];
            d.logp_old = a.logp_old[slot];
        }
        if (a.value_index >= 0) d.ret = a.ret[slot];
        if (a.cfg.value_clip) d.v_old = a.v_old[slot];
    }
    tc_load_row(a.obs, slot, a.in_dim, valid, 1.0f, xr);
}

__global__ void __launch_bounds__(TC_THREADS) tc_ppo_minibatch_kernel(TcPpoArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    TcSetup s = tc_setup(smem, a.params, a.in_dim, a.out_dim, TM_COLS_BWD);
    constexpr uint32_t ID_DH = tc::make_idesc_f16(128, 64, 0, 0, 0, 1);    // K-major A, MN-major B
    constexpr uint32_t ID_G64 = tc::make_idesc_f16(128, 64, 0, 0, 1, 1);   // both MN-major (K = rows)
    constexpr uint32_t ID_G32 = tc::make_idesc_f16(128, 32, 0, 0, 1, 1);
    constexpr uint32_t ID_G16 = tc::make_idesc_f16(128, 16, 0, 0, 1, 1);
    const int r = threadIdx.x;
    const int64_t n_tiles = (a.mb_len + TC_ROWS - 1) / TC_ROWS;
    const float inv_n = 1.0f / (float)a.mb_len;
    float mean = 0.f, rstd = 1.f;
    if (a.cfg.adv_norm && a.mb_stats) { mean = a.mb_stats[0]; rstd = a.mb_stats[1]; }
    float gb3[OUTP];
#pragma unroll
    for (int j = 0; j < OUTP; ++j) gb3[j] = 0.f;
    double l_clip = 0.0, l_vf = 0.0, l_ent = 0.0;

    float xr[TC_XF];
    RowData nd;
    bool nvalid = false;
    int64_t tile = blockIdx.x;
    if (tile < n_tiles) tc_prefetch_row(a, tile, r, nvalid, xr, nd);
    uint32_t it = 0;
    for (; tile < n_tiles; tile += gridDim.x, ++it) {
        const bool valid = nvalid;
        const RowData d = nd;
        tc_store_x(smem, xr);
        if (tile + gridDim.x < n_tiles) tc_prefetch_row(a, tile + gridDim.x, r, nvalid, xr, nd);
        tc_forward_tile(smem, s);
        // ---- loss and G = S * d(loss)/d(OUT) for this thread's row ----
        float z[OUTP], dz[OUTP];
        tc_read_out(smem, s, z);
        const RowLoss rl = row_loss(z, a.n_act, a.value_index, d.act, d.adv, d.ret, d.logp_old, d.v_old, mean, rstd, a.cfg, inv_n, dz);
        if (valid) {
            l_clip += (double)rl.clip;
            l_vf += (double)rl.vf;
            l_ent += (double)rl.ent;
        }
        {
            float g[OUTP];
#pragma unroll
            for (int j = 0; j < OUTP; ++j) {
                if (!valid) dz[j] = 0.f;
                gb3[j] += dz[j];
                g[j] = fminf(fmaxf(dz[j] * a.gscale, -65000.f), 65000.f);
            }
            uint4 w;
            w.x = pack_h2(g[0], g[1]); w.y = pack_h2(g[2], g[3]); w.z = pack_h2(g[4], g[5]); w.w = pack_h2(g[6], g[7]);
            *reinterpret_cast<uint4*>(smem + SM_G + r * 16) = w;
        }
        tc_sync_for_mma();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [dW3^T; .] += [H2; H1]^T G
                tc::mma_f16(s.tmem + TM_GW3, desc_act_mn(s.sbase + SM_H2, ks), desc_act_mn(s.sbase + SM_G, ks), ID_G16, (it | ks) > 0);
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase + SM_G, 0), desc_w_mn(s.sbase + SM_W3, TC_OF, 0), ID_DH, 0);  // DH2 = G W3
            tc::commit(s.bar);
        }
        tc_wait(s);
        tc_epilogue_mask(smem, s.lane_base, SM_H2, SM_H2);  // D2 over H2 (the dW3 GEMM has completed)
        tc_sync_for_mma();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int ks = 0; ks < TC_H / 16; ++ks)  // DH1 = D2 W2
                tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase + SM_H2, ks), desc_w_mn(s.sbase + SM_W2, TC_H, ks), ID_DH, ks > 0);
            tc::commit(s.bar);
        }
        tc_wait(s);
        tc_epilogue_mask(smem, s.lane_base, SM_H1, SM_D1);
        tc_sync_for_mma();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [dW2; .] += [D2; H1]^T H1
                tc::mma_f16(s.tmem + TM_GW2, desc_act_mn(s.sbase + SM_H2, ks), desc_act_mn(s.sbase + SM_H1, ks), ID_G64, (it | ks) > 0);
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [dW1 | db1; . | db2] += [D1; D2]^T X
                tc::mma_f16(s.tmem + TM_GW1, desc_act_mn(s.sbase + SM_D1, ks), desc_act_mn(s.sbase + SM_X, ks), ID_G32, (it | ks) > 0);
            tc::commit(s.bar);
        }
        tc_wait(s);  // X / H1 / H2 / D1 are free again
    }
    // ---- flush: TMEM gradient accumulators -> flat gradient buffer (scaled back by 1/S) ----
    const Offsets o = offsets(a.in_dim, TC_H, a.out_dim);
    const float inv_s = a.inv_gscale;
    if (it > 0) {
        if (r < TC_H) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {  // dW2[r][16q..]
                float v[16];
                tc::tmem_ld16(s.lane_base + TM_GW2 + 16 * q, v);
#pragma unroll
                for (int e = 0; e < 16; ++e) atomicAdd(a.grads + o.w2 + (int64_t)r * TC_H + 16 * q + e, v[e] * inv_s);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {  // dW1[r][k], db1[r] (column 31)
                float v[16];
                tc::tmem_ld16(s.lane_base + TM_GW1 + 16 * q, v);
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const int k = 16 * q + e;
                    if (k < a.in_dim) atomicAdd(a.grads + o.w1 + (int64_t)r * a.in_dim + k, v[e] * inv_s);
                    if (k == TC_XF - 1) atomicAdd(a.grads + o.b1 + r, v[e] * inv_s);
                }
            }
            {  // dW3^T[r][o]
                float v[16];
                tc::tmem_ld16(s.lane_base + TM_GW3, v);
#pragma unroll
                for (int e = 0; e < OUTP; ++e)
                    if (e < a.out_dim) atomicAdd(a.grads + o.w3 + (int64_t)e * TC_H + r, v[e] * inv_s);
            }
        } else {  // lanes 64..127 of the [D1; D2]^T X accumulator: column 31 = db2
            float v[16];
            tc::tmem_ld16(s.lane_base + TM_GW1 + 16, v);
            atomicAdd(a.grads + o.b2 + (r - TC_H), v[15] * inv_s);
        }
    }
    // db3 and the loss sums: warp shuffle, then one atomic per warp
#pragma unroll
    for (int dlt = 16; dlt > 0; dlt >>= 1) {
#pragma unroll
        for (int j = 0; j < OUTP; ++j) gb3[j] += __shfl_down_sync(0xffffffffu, gb3[j], dlt);
        l_clip += __shfl_down_sync(0xffffffffu, l_clip, dlt);
        l_vf += __shfl_down_sync(0xffffffffu, l_vf, dlt);
        l_ent += __shfl_down_sync(0xffffffffu, l_ent, dlt);
    }
    if ((threadIdx.x & 31) == 0 && it > 0) {
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j < a.out_dim) atomicAdd(a.grads + o.b3 + j, gb3[j]);
        if (a.loss_sums) {
            atomicAdd(a.loss_sums + 0, l_clip);
            atomicAdd(a.loss_sums + 1, l_vf);
            atomicAdd(a.loss_sums + 2, l_ent);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) tc::tmem_dealloc(s.tmem, TM_COLS_BWD);
}
