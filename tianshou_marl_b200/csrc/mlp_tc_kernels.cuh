// mlp_tc_kernels.cuh -- device code of the tcgen05 actor-critic path (included by mlp_tc.cu only).
//
// Net: in(<=31) -> 64 -> 64 -> out(<=8).  One CTA = 128 threads = one tile of 128 rows; thread t
// owns row t (TMEM lane t).  All GEMM operands are fp16 in the chunked layout of tc.cuh with 8
// features per 16-byte chunk:  act[f/8][row 0..127][f%8],  weight[in/8][out][in%8].  Input feature
// 31 is a column of ones; the biases ride on it (adam.cuh: b1 inside the W1 image, b2 / b3 as one
// extra K = 16 block against X features 16..31), so no epilogue adds a bias.
//
//   forward      H1 = relu([X|1] W1b^T)                 A = X   (K-major, K = 32)   B = W1b      N = 64
//                H2 = relu(H1 W2^T + [X|1]_hi B2^T)     A = H1  (K = 64) + X hi (K = 16)         N = 64
//                OUT = H2 W3^T + [X|1]_hi B3^T          A = H2  (K = 64) + X hi (K = 16)         N = 16
//   data grads   DH2 = G W3                             A = G   (K-major, K = 16)   B = W3 (MN-major) N = 64
//                DH1 = D2 W2                            A = D2  (K-major, K = 64)   B = W2 (MN-major) N = 64
//   weight grads (K = the 128 rows, both operands MN-major views of the activation images,
//                 M = 128 made of two adjacent 64-feature images, [H2][H1] or [D2][D1]):
//                [dW3^T; .]        += [H2; H1]^T G        N = 16   (lanes 0..63 used)
//                [dW2 ; .]         += [D2; D1]^T H1       N = 64   (lanes 0..63 used)
//                [.|db2; dW1|db1]  += [D2; D1]^T [X|1]    N = 32   (column 31 = bias gradients)
// G = S * d(loss)/d(OUT) with S a power of two (fp16 range); D2 / D1 = ReLU-masked DH2 / DH1.
// Only the five dependent GEMM stages (3 forward, 2 data-gradient) sit on a tile's critical path: the
// 24 weight-gradient MMAs are issued last and never waited for inside the tile (the next tile's first
// wait covers them; X is double buffered).  Weight-gradient accumulators stay in TMEM across all tiles
// of the CTA and are flushed once, through a shared-memory staging copy so that the global atomics
// are coalesced.
#pragma once

constexpr int TC_THREADS = 128;
constexpr int TC_ROWS = 128;
constexpr int TC_H = 64;
constexpr int TC_XF = 32;    // padded input features (31 = ones)
constexpr int TC_OF = 16;    // padded head outputs (UMMA N >= 16 at M = 128)
constexpr int CHUNK_B = TC_ROWS * 16;  // bytes of one 8-feature chunk of an activation image (2048)

// shared-memory map (bytes).  [H2][H1] and [D2][D1] must each stay adjacent and in this order.
constexpr int SM_H2 = 0;                       // 16 KB
constexpr int SM_H1 = SM_H2 + 8 * CHUNK_B;     // 16 KB
constexpr int SM_X = SM_H1 + 8 * CHUNK_B;      // 2 x 8 KB, double buffered
constexpr int SM_G = SM_X + 2 * 4 * CHUNK_B;   // 2 chunks = 4 KB (chunk 1 stays zero)
constexpr int SM_IMG = SM_G + 2 * CHUNK_B;     // weight image (adam.cuh)
constexpr int SM_W1 = SM_IMG + opt::IMG_W1 * 2;
constexpr int SM_W2 = SM_IMG + opt::IMG_W2 * 2;
constexpr int SM_W3 = SM_IMG + opt::IMG_W3 * 2;
constexpr int SM_B2I = SM_IMG + opt::IMG_B2 * 2;
constexpr int SM_B3I = SM_IMG + opt::IMG_B3 * 2;
constexpr int SM_RED = SM_IMG + opt::IMG_BYTES;  // 4 warps x 8 floats + 4 x 3 doubles
constexpr int SM_BAR = SM_RED + 384;             // 2 mbarriers + tmem base
constexpr int SM_FWD_TOTAL = SM_BAR + 64;        // the forward kernel stops here (~69 KB: three CTAs per SM)
constexpr int SM_D2 = SM_FWD_TOTAL;              // 16 KB  (backward only)
constexpr int SM_D1 = SM_D2 + 8 * CHUNK_B;       // 16 KB
constexpr int SM_TOTAL = SM_D1 + 8 * CHUNK_B;    // ~101 KB: two CTAs per SM
constexpr int SM_STAGE = SM_H2;                  // flush staging: fp32 copy of the flat gradient (<= 32 KB)

// TMEM columns (fp32 accumulators, 128 lanes)
constexpr uint32_t TM_SCR = 0;     // 64: H1pre / H2pre / DH2 / DH1 in turn
constexpr uint32_t TM_OUT = 64;    // 16
constexpr uint32_t TM_GW2 = 96;    // 64
constexpr uint32_t TM_GW1 = 160;   // 32
constexpr uint32_t TM_GW3 = 192;   // 16
constexpr uint32_t TM_COLS_BWD = 256;
constexpr uint32_t TM_COLS_FWD = 128;

__device__ __forceinline__ uint32_t h2_bits(__half2 h) { return *reinterpret_cast<const uint32_t*>(&h); }
__device__ __forceinline__ uint32_t pack_h2(float a, float b) { return h2_bits(__floats2half2_rn(a, b)); }
__device__ __forceinline__ uint32_t pack_relu_h2(float a, float b) {
    return h2_bits(__hmax2(__floats2half2_rn(a, b), __float2half2_rn(0.f)));
}

// this thread's input row (fp32 registers, feature 31 = 1) -> X image at smem + xoff
__device__ __forceinline__ void tc_store_x(unsigned char* smem, int xoff, const float (&x)[TC_XF]) {
#pragma unroll
    for (int c = 0; c < TC_XF / 8; ++c) {
        uint4 v;
        v.x = pack_h2(x[8 * c + 0], x[8 * c + 1]);
        v.y = pack_h2(x[8 * c + 2], x[8 * c + 3]);
        v.z = pack_h2(x[8 * c + 4], x[8 * c + 5]);
        v.w = pack_h2(x[8 * c + 6], x[8 * c + 7]);
        *reinterpret_cast<uint4*>(smem + xoff + c * CHUNK_B + threadIdx.x * 16) = v;
    }
}

// gather this thread's row of x (row-major [.][in_dim] fp32) into registers; zero when !valid
__device__ __forceinline__ void tc_load_row(const float* __restrict__ x, int64_t row, int in_dim, bool valid, float (&r)[TC_XF]) {
#pragma unroll
    for (int k = 0; k < TC_XF; ++k) r[k] = 0.f;
    if (valid) {
        const float* p = x + row * in_dim;
        if ((in_dim & 1) == 0 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) {
#pragma unroll
            for (int k = 0; k < TC_XF / 2; ++k)
                if (2 * k < in_dim) {
                    const float2 v = __ldg(reinterpret_cast<const float2*>(p + 2 * k));
                    r[2 * k] = v.x;
                    r[2 * k + 1] = v.y;
                }
        } else {
#pragma unroll
            for (int k = 0; k < TC_XF; ++k)
                if (k < in_dim) r[k] = __ldg(p + k);
        }
        r[TC_XF - 1] = 1.0f;
    }
}

// relu(SCR accumulator, 64 cols of this thread's lane) -> fp16 image at smem + dst
__device__ __forceinline__ void tc_epilogue_relu(unsigned char* smem, uint32_t tmem_lane_base, int dst) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(tmem_lane_base + TM_SCR + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            uint4 w;
            w.x = pack_relu_h2(v[8 * h + 0], v[8 * h + 1]);
            w.y = pack_relu_h2(v[8 * h + 2], v[8 * h + 3]);
            w.z = pack_relu_h2(v[8 * h + 4], v[8 * h + 5]);
            w.w = pack_relu_h2(v[8 * h + 6], v[8 * h + 7]);
            *reinterpret_cast<uint4*>(smem + dst + (4 * q + h) * CHUNK_B + threadIdx.x * 16) = w;
        }
    }
}

// SCR accumulator masked by (activation image at `img` > 0) -> fp16 image at `dst`
__device__ __forceinline__ void tc_epilogue_mask(unsigned char* smem, uint32_t tmem_lane_base, int img, int dst) {
    const __half2 zero = __float2half2_rn(0.f);
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(tmem_lane_base + TM_SCR + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int cell = (4 * q + h) * CHUNK_B + threadIdx.x * 16;
            const uint4 m = *reinterpret_cast<const uint4*>(smem + img + cell);
            const __half2* mh = reinterpret_cast<const __half2*>(&m);
            uint4 w;
            w.x = pack_h2(v[8 * h + 0], v[8 * h + 1]) & __hgt2_mask(mh[0], zero);
            w.y = pack_h2(v[8 * h + 2], v[8 * h + 3]) & __hgt2_mask(mh[1], zero);
            w.z = pack_h2(v[8 * h + 4], v[8 * h + 5]) & __hgt2_mask(mh[2], zero);
            w.w = pack_h2(v[8 * h + 6], v[8 * h + 7]) & __hgt2_mask(mh[3], zero);
            *reinterpret_cast<uint4*>(smem + dst + cell) = w;
        }
    }
}

// descriptor helpers for the four operand views (fp16, 8 features per chunk).  sb4 = (CTA smem base) >> 4; every
// offset is a compile-time constant, so each descriptor costs one integer add at run time.
__device__ __forceinline__ uint64_t desc_at(uint32_t sb4, uint32_t off_bytes, uint32_t lbo, uint32_t sbo) {
    const uint32_t lo = sb4 + ((off_bytes >> 4) + ((lbo >> 4) << 16));  // start address field cannot carry: smem < 256 KB
    const uint32_t hi = (sbo >> 4) | (1u << 14);                        // descriptor version bit 46
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t desc_act_k(uint32_t sb4, int off, int kstep) {  // rows x features, K = features
    return desc_at(sb4, off + kstep * 2 * CHUNK_B, CHUNK_B, 128);
}
__device__ __forceinline__ uint64_t desc_act_mn(uint32_t sb4, int off, int kstep) {  // features x rows, K = rows
    return desc_at(sb4, off + kstep * 256, 128, CHUNK_B);
}
__device__ __forceinline__ uint64_t desc_w_k(uint32_t sb4, int off, int n_out, int kstep) {  // N = out, K = in
    return desc_at(sb4, off + kstep * 2 * n_out * 16, n_out * 16, 128);
}
__device__ __forceinline__ uint64_t desc_w_mn(uint32_t sb4, int off, int n_out, int kstep) {  // N = in, K = out
    return desc_at(sb4, off + kstep * 256, 128, n_out * 16);
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
    const unsigned char* image;
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

#ifdef TSM_TC_TRACE
__device__ long long g_trace[4][64];
#define TRACE(row, slot) do { if (threadIdx.x == 0 && (row) >= 0) g_trace[row][slot] = clock64(); } while (0)
#else
#define TRACE(row, slot) do { } while (0)
#endif

struct TcSetup {
    uint64_t* bar;   // bar[0]: waited phases; bar[1]: the unwaited weight-gradient commits of the backward
    uint32_t tmem, lane_base, sbase, phase;  // sbase: CTA smem base >> 4
};

// part A: nothing here depends on the parameters (runs ahead of pdl_wait())
__device__ __forceinline__ void tc_setup_begin(unsigned char* smem, uint32_t tmem_cols) {
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    uint32_t* tmem_s = reinterpret_cast<uint32_t*>(smem + SM_BAR + 16);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 2 * CHUNK_B / 16; i += TC_THREADS)  // G image (chunk 1 stays zero forever)
        *reinterpret_cast<uint4*>(smem + SM_G + i * 16) = make_uint4(0, 0, 0, 0);
    if (warp == 0) tc::tmem_alloc(tmem_s, tmem_cols);
    if (threadIdx.x == 0) {
        tc::mbar_init(bar, 1);
        tc::mbar_init(bar + 1, 1);
        tc::fence_barrier_init();
    }
}

// part B: weight image (straight 16-byte copies; the image is rebuilt by the optimizer step, not here), then publish
__device__ __forceinline__ TcSetup tc_setup_finish(unsigned char* smem, const unsigned char* image) {
    TcSetup s;
    s.bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    uint32_t* tmem_s = reinterpret_cast<uint32_t*>(smem + SM_BAR + 16);
    const int warp = threadIdx.x >> 5;
    {
        const uint4* src = reinterpret_cast<const uint4*>(image + (blockIdx.x % opt::IMG_COPIES) * opt::IMG_BYTES);
        uint4* dst = reinterpret_cast<uint4*>(smem + SM_IMG);
        constexpr int N16 = opt::IMG_BYTES / 16, PER = (N16 + TC_THREADS - 1) / TC_THREADS;
        uint4 v[PER];  // all loads first (one L2 round trip), then the stores
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int i = threadIdx.x + u * TC_THREADS;
            if (i < N16) v[u] = __ldcg(src + i);
        }
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int i = threadIdx.x + u * TC_THREADS;
            if (i < N16) dst[i] = v[u];
        }
    }
    tc_sync_for_mma();
    s.tmem = *tmem_s;
    s.lane_base = s.tmem + ((uint32_t)(warp * 32) << 16);
    s.sbase = tc::smem_u32(smem) >> 4;  // descriptor units
    s.phase = 0;
    return s;
}

__device__ __forceinline__ void tc_wait(TcSetup& s) {
    tc::mbar_wait(s.bar, s.phase);
    s.phase ^= 1;
    tc::fence_after_sync();
}

// [X|1] -> H1 -> H2 -> OUT for the tile in the X image at xoff; leaves OUT in TMEM (waited)
__device__ __forceinline__ void tc_forward_tile(unsigned char* smem, TcSetup& s, int xoff, int trow = -1) {
    constexpr uint32_t ID64 = tc::make_idesc_f16(128, 64, 0, 0, 0, 0);
    constexpr uint32_t ID16 = tc::make_idesc_f16(128, 16, 0, 0, 0, 0);
    TRACE(trow, 1);
    tc_sync_for_mma();
    TRACE(trow, 2);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_XF / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, xoff, ks), desc_w_k(s.sbase, SM_W1, TC_H, ks), ID64, ks > 0);
        tc::commit(s.bar);
    }
    TRACE(trow, 3);
    tc_wait(s);
    TRACE(trow, 4);
    tc_epilogue_relu(smem, s.lane_base, SM_H1);
    TRACE(trow, 5);
    tc_sync_for_mma();
    TRACE(trow, 6);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_H1, ks), desc_w_k(s.sbase, SM_W2, TC_H, ks), ID64, ks > 0);
        tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, xoff, 1), desc_w_k(s.sbase, SM_B2I, TC_H, 0), ID64, 1);  // + b2
        tc::commit(s.bar);
    }
    TRACE(trow, 7);
    tc_wait(s);
    TRACE(trow, 8);
    tc_epilogue_relu(smem, s.lane_base, SM_H2);
    TRACE(trow, 9);
    tc_sync_for_mma();
    TRACE(trow, 10);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_OUT, desc_act_k(s.sbase, SM_H2, ks), desc_w_k(s.sbase, SM_W3, TC_OF, ks), ID16, ks > 0);
        tc::mma_f16(s.tmem + TM_OUT, desc_act_k(s.sbase, xoff, 1), desc_w_k(s.sbase, SM_B3I, TC_OF, 0), ID16, 1);  // + b3
        tc::commit(s.bar);
    }
    TRACE(trow, 11);
    tc_wait(s);
    TRACE(trow, 12);
}

__device__ __forceinline__ void tc_read_out(const TcSetup& s, float (&z)[OUTP]) {
    float o[16];
    tc::tmem_ld16(s.lane_base + TM_OUT, o);
#pragma unroll
    for (int j = 0; j < OUTP; ++j) z[j] = o[j];
}

__global__ void __launch_bounds__(TC_THREADS) tc_forward_kernel(TcFwdArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tc_setup_begin(smem, TM_COLS_FWD);
    TcSetup s = tc_setup_finish(smem, a.image);
    const int64_t n_tiles = (a.n + TC_ROWS - 1) / TC_ROWS;
    const int r = threadIdx.x;
    float xr[TC_XF];
    int64_t tile = blockIdx.x;
    int64_t nsrc = 0;
    bool nvalid = false;
    if (tile < n_tiles) {
        nvalid = tile * TC_ROWS + r < a.n;
        nsrc = nvalid ? (a.rows ? a.rows[tile * TC_ROWS + r] : tile * TC_ROWS + r) : 0;
        tc_load_row(a.x, nsrc, a.in_dim, nvalid, xr);
    }
    for (uint32_t it = 0; tile < n_tiles; tile += gridDim.x, ++it) {
        const bool valid = nvalid;
        const int64_t src = nsrc;
        const int64_t dst = (a.scatter && a.rows) ? src : tile * TC_ROWS + r;
        const int xoff = SM_X + (int)(it & 1) * 4 * CHUNK_B;
        tc_store_x(smem, xoff, xr);
        const int64_t nt = tile + gridDim.x;
        if (nt < n_tiles) {  // prefetch the next tile's row while this one computes
            nvalid = nt * TC_ROWS + r < a.n;
            nsrc = nvalid ? (a.rows ? a.rows[nt * TC_ROWS + r] : nt * TC_ROWS + r) : 0;
            tc_load_row(a.x, nsrc, a.in_dim, nvalid, xr);
        }
        tc_forward_tile(smem, s, xoff);
        float z[OUTP];
        tc_read_out(s, z);
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
                const Soft sm = softmax_row<true>(z, a.n_act);
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
    const unsigned char* image;
    float* grads;
    const unsigned char* records;  // [slot] 96-byte packed rows (tsm_ppo_pack_records)
    BatchView view;
    int64_t mb_start, mb_len;
    const float* mb_stats;
    PpoCfg cfg;
    int in_dim, out_dim, n_act, value_index;
    float gscale, inv_gscale;  // S and 1/S (power of two)
    double* loss_sums;
    int debug;  // experiments only (TSM_TC_DEBUG): 1 skip flush, 2 skip backward GEMM phases, 4 skip the loss math
};

constexpr int REC_BYTES = 96;  // halves x[32] | float adv, ret, logp_old | int32 act | float v_old | 12 bytes pad

// one minibatch row: six 16-byte loads from one record (prefetched a tile ahead, kept in registers)
struct RowRec {
    uint4 x[4];   // fp16 input features, already in X-image chunk form
    uint4 sc;     // adv, ret, logp_old, act
    float v_old;
};

__device__ __forceinline__ void tc_prefetch_row(const TcPpoArgs& a, int64_t tile, int r, bool& valid, RowRec& d) {
    const int64_t p = a.mb_start + tile * TC_ROWS + r;
    valid = tile * TC_ROWS + r < a.mb_len;
    if (valid) {
        const int64_t slot = (a.debug & 8) ? p : slot_of(a.view, p);
        const uint4* rec = reinterpret_cast<const uint4*>(a.records + slot * REC_BYTES);
#pragma unroll
        for (int c = 0; c < 4; ++c) d.x[c] = __ldg(rec + c);
        d.sc = __ldg(rec + 4);
        d.v_old = __ldg(reinterpret_cast<const float*>(rec + 5));
    } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) d.x[c] = make_uint4(0, 0, 0, 0);
        d.sc = make_uint4(0, 0, 0, 0);
        d.v_old = 0.f;
    }
}

__device__ __forceinline__ void tc_store_rec(unsigned char* smem, int xoff, const RowRec& d) {
#pragma unroll
    for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(smem + xoff + c * CHUNK_B + threadIdx.x * 16) = d.x[c];
}

// physical position of flat gradient element i inside the staging copy: the W2 block is XOR-swizzled
// so that both the per-lane row writes (stride 64 floats) and the linear read-out are bank-conflict free
__device__ __forceinline__ int stage_pos(int i, int w2_lo, int w2_hi) {
    if (i >= w2_lo && i < w2_hi) {
        const int rel = i - w2_lo, rr = rel >> 6, cc = rel & 63;
        return w2_lo + (rr << 6) + (cc ^ (rr & 31));
    }
    return i;
}

__global__ void __launch_bounds__(TC_THREADS) tc_ppo_minibatch_kernel(TcPpoArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    [[maybe_unused]] const int krow = blockIdx.x == 0 ? 0 : (blockIdx.x == gridDim.x - 1 ? 2 : -1);
    TRACE(krow, 30);
    pdl_launch_dependents();
    tc_setup_begin(smem, TM_COLS_BWD);
    TRACE(krow, 31);
    constexpr uint32_t ID_DH = tc::make_idesc_f16(128, 64, 0, 0, 0, 1);    // K-major A, MN-major B
    constexpr uint32_t ID_G64 = tc::make_idesc_f16(128, 64, 0, 0, 1, 1);   // both MN-major (K = rows)
    constexpr uint32_t ID_G32 = tc::make_idesc_f16(128, 32, 0, 0, 1, 1);
    constexpr uint32_t ID_G16 = tc::make_idesc_f16(128, 16, 0, 0, 1, 1);
    const int r = threadIdx.x;
    const int64_t n_tiles = (a.mb_len + TC_ROWS - 1) / TC_ROWS;
    const float inv_n = 1.0f / (float)a.mb_len;
    float gb3[OUTP];
#pragma unroll
    for (int j = 0; j < OUTP; ++j) gb3[j] = 0.f;
    float f_clip = 0.f, f_vf = 0.f, f_ent = 0.f;  // per-thread partial sums of up to 16 tiles, then folded into doubles
    double l_clip = 0.0, l_vf = 0.0, l_ent = 0.0;

    RowRec nd;
    bool nvalid = false;
    int64_t tile = blockIdx.x;
    if (tile < n_tiles) tc_prefetch_row(a, tile, r, nvalid, nd);  // batch records only: safe ahead of the dependency wait
    TRACE(krow, 32);
    pdl_wait();  // parameters (image), gradients and mb_stats of the preceding launches are now visible
    TcSetup s = tc_setup_finish(smem, a.image);
    TRACE(krow, 33);
    float mean = 0.f, rstd = 1.f;
    if ((a.cfg.adv_norm || a.cfg.plain_pg == 2) && a.mb_stats) { mean = __ldcg(a.mb_stats); rstd = __ldcg(a.mb_stats + 1); }
    uint32_t it = 0;
    for (; tile < n_tiles; tile += gridDim.x, ++it) {
        const bool valid = nvalid;
        const int xoff = SM_X + (int)(it & 1) * 4 * CHUNK_B;
        const int trow = (it < 2 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? (blockIdx.x != 0) * 2 + (int)it : -1;
        TRACE(trow, 0);
        tc_store_rec(smem, xoff, nd);
        const float d_adv = __uint_as_float(nd.sc.x), d_ret = __uint_as_float(nd.sc.y), d_logp_old = __uint_as_float(nd.sc.z);
        const int d_act = (int)nd.sc.w;
        const float d_v_old = nd.v_old;
        if (tile + gridDim.x < n_tiles) tc_prefetch_row(a, tile + gridDim.x, r, nvalid, nd);
        tc_forward_tile(smem, s, xoff, trow);  // its first wait also covers the previous tile's weight-gradient GEMMs
        // ---- loss and G = S * d(loss)/d(OUT) for this thread's row ----
        float z[OUTP], dz[OUTP];
        tc_read_out(s, z);
        RowLoss rl{0.f, 0.f, 0.f};
        if (a.debug & 4) {
#pragma unroll
            for (int j = 0; j < OUTP; ++j) dz[j] = z[j] * inv_n;
        } else {
            rl = row_loss<true>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz);
        }
        if (valid) {
            f_clip += rl.clip;
            f_vf += rl.vf;
            f_ent += rl.ent;
        }
        if ((it & 15) == 15) {
            l_clip += (double)f_clip; l_vf += (double)f_vf; l_ent += (double)f_ent;
            f_clip = f_vf = f_ent = 0.f;
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
        if (a.debug & 2) {
            tc_sync_for_mma();
            if (threadIdx.x == 0) tc::commit(s.bar + 1);
            continue;
        }
        TRACE(trow, 13);
        tc_sync_for_mma();
        TRACE(trow, 14);
        if (threadIdx.x == 0) {
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_G, 0), desc_w_mn(s.sbase, SM_W3, TC_OF, 0), ID_DH, 0);  // DH2 = G W3
            tc::commit(s.bar);
        }
        TRACE(trow, 15);
        tc_wait(s);
        TRACE(trow, 16);
        tc_epilogue_mask(smem, s.lane_base, SM_H2, SM_D2);
        TRACE(trow, 17);
        tc_sync_for_mma();
        TRACE(trow, 18);
        if (threadIdx.x == 0) {
#pragma unroll
            for (int ks = 0; ks < TC_H / 16; ++ks)  // DH1 = D2 W2
                tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_D2, ks), desc_w_mn(s.sbase, SM_W2, TC_H, ks), ID_DH, ks > 0);
            tc::commit(s.bar);
        }
        TRACE(trow, 19);
        tc_wait(s);
        TRACE(trow, 20);
        tc_epilogue_mask(smem, s.lane_base, SM_H1, SM_D1);
        TRACE(trow, 21);
        tc_sync_for_mma();
        TRACE(trow, 22);
        if (threadIdx.x == 0) {
            // the weight-gradient GEMMs: off the critical path, completion tracked by bar[1] only
            const uint32_t acc0 = it > 0;
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [dW3^T; .] += [H2; H1]^T G
                tc::mma_f16(s.tmem + TM_GW3, desc_act_mn(s.sbase, SM_H2, ks), desc_act_mn(s.sbase, SM_G, ks), ID_G16, ks > 0 ? 1u : acc0);
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [dW2; .] += [D2; D1]^T H1
                tc::mma_f16(s.tmem + TM_GW2, desc_act_mn(s.sbase, SM_D2, ks), desc_act_mn(s.sbase, SM_H1, ks), ID_G64, ks > 0 ? 1u : acc0);
#pragma unroll
            for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [. | db2; dW1 | db1] += [D2; D1]^T [X | 1]
                tc::mma_f16(s.tmem + TM_GW1, desc_act_mn(s.sbase, SM_D2, ks), desc_act_mn(s.sbase, xoff, ks), ID_G32, ks > 0 ? 1u : acc0);
            tc::commit(s.bar + 1);  // not waited here: the next tile's first wait covers it (in-order completion); X is double buffered
        }
        TRACE(trow, 23);
    }
    TRACE(krow, 34);
    const Offsets o = offsets(a.in_dim, TC_H, a.out_dim);
    const float inv_s = a.inv_gscale;
    if (it > 0 && !(a.debug & 1)) {
        // ---- flush: TMEM gradient accumulators -> smem staging (flat gradient layout) -> coalesced atomics ----
        tc::mbar_wait(s.bar + 1, (it - 1) & 1);
        tc::fence_after_sync();
        float* stage = reinterpret_cast<float*>(smem + SM_STAGE);
        const int w2_lo = (int)o.w2;
        if (r < TC_H) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {  // dW2[r][32q..]
                float v[32];
                tc::tmem_ld32(s.lane_base + TM_GW2 + 32 * q, v);
#pragma unroll
                for (int e = 0; e < 32; ++e) stage[w2_lo + (r << 6) + ((32 * q + e) ^ (r & 31))] = v[e];
            }
            float v[16];
            tc::tmem_ld16(s.lane_base + TM_GW3, v);  // dW3^T[r][o]
#pragma unroll
            for (int e = 0; e < OUTP; ++e)
                if (e < a.out_dim) stage[o.w3 + e * TC_H + r] = v[e];
            tc::tmem_ld16(s.lane_base + TM_GW1 + 16, v);  // D2^T [X|1], column 31 = db2[r]
            stage[o.b2 + r] = v[15];
        } else {
            float v[32];
            tc::tmem_ld32(s.lane_base + TM_GW1, v);  // dW1[r-64][k], db1[r-64] (column 31)
            const int rr = r - TC_H;
#pragma unroll
            for (int k = 0; k < 32; ++k)
                if (k < a.in_dim) stage[o.w1 + rr * a.in_dim + k] = v[k];
            stage[o.b1 + rr] = v[31];
        }
    }
    TRACE(krow, 35);
    // db3 and the loss sums: warp shuffle, then shared memory
    l_clip += (double)f_clip; l_vf += (double)f_vf; l_ent += (double)f_ent;
    float* red_f = reinterpret_cast<float*>(smem + SM_RED);
    double* red_d = reinterpret_cast<double*>(smem + SM_RED + 192);
#pragma unroll
    for (int dlt = 16; dlt > 0; dlt >>= 1) {
#pragma unroll
        for (int j = 0; j < OUTP; ++j) gb3[j] += __shfl_down_sync(0xffffffffu, gb3[j], dlt);
        l_clip += __shfl_down_sync(0xffffffffu, l_clip, dlt);
        l_vf += __shfl_down_sync(0xffffffffu, l_vf, dlt);
        l_ent += __shfl_down_sync(0xffffffffu, l_ent, dlt);
    }
    if ((r & 31) == 0) {
        const int w = r >> 5;
#pragma unroll
        for (int j = 0; j < OUTP; ++j) red_f[w * OUTP + j] = gb3[j];
        red_d[w * 3 + 0] = l_clip; red_d[w * 3 + 1] = l_vf; red_d[w * 3 + 2] = l_ent;
    }
    tc::fence_before_sync();
    __syncthreads();
    if ((r >> 5) == 0) tc::tmem_dealloc(s.tmem, TM_COLS_BWD);
    if (it > 0 && !(a.debug & 1)) {
        const float* stage = reinterpret_cast<const float*>(smem + SM_STAGE);
        const int w2_lo = (int)o.w2, w2_hi = (int)o.b2, n_stage = (int)o.b3;
        for (int i = r; i < n_stage; i += TC_THREADS) atomicAdd(a.grads + i, stage[stage_pos(i, w2_lo, w2_hi)] * inv_s);
        if (r < a.out_dim) atomicAdd(a.grads + o.b3 + r, red_f[r] + red_f[OUTP + r] + red_f[2 * OUTP + r] + red_f[3 * OUTP + r]);
        if (a.loss_sums && r < 3) atomicAdd(a.loss_sums + r, red_d[r] + red_d[3 + r] + red_d[6 + r] + red_d[9 + r]);
    }
    TRACE(krow, 38);
}
