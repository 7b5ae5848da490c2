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

// Instrumented builds only (tools/tc_error_budget.py separates the tcgen05 engine's error sources): -DTSM_TC_EXACT_LOSS keeps
// the IEEE loss math of the fp32 engine, -DTSM_TC_EXACT_ADAM the IEEE optimizer arithmetic.  The shipped library has neither.
#ifdef TSM_TC_EXACT_LOSS
constexpr bool TC_FAST_LOSS = false;
#else
constexpr bool TC_FAST_LOSS = true;
#endif

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
constexpr int SM_BAR = SM_RED + 384;             // mbarriers at +0 (MMA, waited), +8 (weight-gradient MMAs), +24 (records); TMEM base at +16
constexpr int SM_FWD_TOTAL = SM_BAR + 64;        // the forward kernel stops here (~69 KB: three CTAs per SM)
constexpr int SM_D2 = SM_FWD_TOTAL;              // 16 KB  (backward only)
constexpr int SM_D1 = SM_D2 + 8 * CHUNK_B;       // 16 KB
constexpr int REC_BYTES = 96;  // halves x[32] | float adv, ret, logp_old | int32 act | float v_old | 12 bytes pad
constexpr int REC_PITCH = 128; // distance between records in device memory: a record never straddles a 128-byte line (at a 96-byte
                               // pitch the gather moved 2.0x the record bytes through DRAM); the TMA box is the first 96 bytes
constexpr int SM_REC = (SM_D1 + 8 * CHUNK_B + 127) & ~127;  // 12 KB: the next tile's batch records, landed by TMA gather4 (128-byte aligned)
constexpr int SM_TOTAL = SM_REC + TC_ROWS * REC_BYTES;  // 115,712 B: still two CTAs per SM (2 x (this + 1 KB) = 228 KB exactly)
constexpr int SM_STAGE = SM_H2;                  // flush staging: fp32 copy of the flat gradient (<= 32 KB)

// TMEM columns (fp32 accumulators, 128 lanes)
constexpr uint32_t TM_SCR = 0;     // 64: H1pre / H2pre / DH2 / DH1 in turn
constexpr uint32_t TM_OUT = 64;    // 16
constexpr uint32_t TM_GW2 = 96;    // 64
constexpr uint32_t TM_GW1 = 160;   // 32
constexpr uint32_t TM_GW3 = 192;   // 16
constexpr uint32_t TM_COLS_BWD = 256;
constexpr uint32_t TM_COLS_FWD = 128;

// A "group" is one 128-thread warpgroup working on one tile.  The forward and minibatch kernels run one group per CTA;
// the persistent epoch kernel runs TC_GROUPS of them per CTA, each with its own shared-memory map, TMEM columns,
// mbarriers and named barrier (id 1 + group), so every helper below addresses by group-local thread index.
constexpr int TC_GROUPS = 2;
__device__ __forceinline__ int tc_tid() { return (int)(threadIdx.x & (TC_THREADS - 1)); }
__device__ __forceinline__ int tc_grp() { return (int)(threadIdx.x / TC_THREADS); }
// The thread that issues a group's MMAs: one elected lane of the group's first warp.  The warp index is a shuffle
// broadcast, so the branch is warp-uniform for the compiler, and TcSetup::tmem / sbase are broadcast values too: the
// descriptors are then built in uniform registers.  (With `if (tid == 0)` every operand went through R2UR inside an
// elect loop: ~15 dependent instructions = 85 cycles per MMA, 3.9k of a tile's 10.3k cycles.)
// HV ("halves"): threads per tile row.  1 everywhere except the persistent epoch kernel, which runs a group as 2 x 128
// threads: thread t of a group works on row t & 127 and on half (t >> 7) of that row's columns (warps 0-3 and 4-7 of a
// group see the same TMEM lane quarters), so every CUDA-core stage between two MMAs is half as long.
template <int HV = 1>
__device__ __forceinline__ int tc_warp_u() { return __shfl_sync(0xffffffffu, (int)((threadIdx.x >> 5) & (HV * TC_THREADS / 32 - 1)), 0); }
__device__ __forceinline__ bool tc_issuer(bool w0) { return w0 && tc::elect_one(); }  // w0 = TcSetup::w0 (warp-uniform, computed once)
template <int HV = 1>
__device__ __forceinline__ void tc_group_sync() {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + (int)(threadIdx.x / (HV * TC_THREADS))), "n"(HV * TC_THREADS) : "memory");
}

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
        *reinterpret_cast<uint4*>(smem + xoff + c * CHUNK_B + tc_tid() * 16) = v;
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
template <int HV = 1>
__device__ __forceinline__ void tc_epilogue_relu(unsigned char* smem, uint32_t tmem_lane_base, int dst, int half = 0) {
#pragma unroll
    for (int q0 = 0; q0 < 2 / HV; ++q0) {
        const int q = HV == 1 ? q0 : half;
        float v[32];
        tc::tmem_ld32(tmem_lane_base + TM_SCR + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            uint4 w;
            w.x = pack_relu_h2(v[8 * h + 0], v[8 * h + 1]);
            w.y = pack_relu_h2(v[8 * h + 2], v[8 * h + 3]);
            w.z = pack_relu_h2(v[8 * h + 4], v[8 * h + 5]);
            w.w = pack_relu_h2(v[8 * h + 6], v[8 * h + 7]);
            *reinterpret_cast<uint4*>(smem + dst + (4 * q + h) * CHUNK_B + tc_tid() * 16) = w;
        }
    }
}

// SCR accumulator masked by (activation image at `img` > 0) -> fp16 image at `dst`
template <int HV = 1>
__device__ __forceinline__ void tc_epilogue_mask(unsigned char* smem, uint32_t tmem_lane_base, int img, int dst, int half = 0) {
    const __half2 zero = __float2half2_rn(0.f);
#pragma unroll
    for (int q0 = 0; q0 < 2 / HV; ++q0) {
        const int q = HV == 1 ? q0 : half;
        float v[32];
        tc::tmem_ld32(tmem_lane_base + TM_SCR + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int cell = (4 * q + h) * CHUNK_B + tc_tid() * 16;
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
template <int HV = 1>
__device__ __forceinline__ void tc_sync_for_mma() {
    tc::fence_async_smem();
    tc::fence_before_sync();
    tc_group_sync<HV>();
    tc::fence_after_sync();
}

// ---------------------------------------------------------------------------------------------
struct TcFwdArgs {
    const unsigned char* image;
    const float* x;
    const int64_t* rows;
    int64_t n;
    int in_dim, out_dim, n_act, value_index, mode, scatter, probs;
    uint64_t seed, counter;
    int64_t rng_row0;  // added to the row index in the sampling key (a launch over rows [row0, row0 + n) of a larger batch)
    float* out;
    int32_t* act;
    float* logp;
    float* value;
    float* entropy;
};

#ifdef TSM_TC_TRACE
__device__ long long g_trace[4][64];
#define TRACE(row, slot) do { if (threadIdx.x == 0 && (row) >= 0) g_trace[row][slot] = clock64(); } while (0)
#define TRACE_H1(on, slot) do { if (threadIdx.x == 128 && (on)) g_trace[3][slot] = clock64(); } while (0)  // first half-1 thread of group 0
#else
#define TRACE(row, slot) do { } while (0)
#define TRACE_H1(on, slot) do { } while (0)
#endif

struct TcSetup {
    uint64_t* bar;   // bar[0]: waited phases; bar[1]: the unwaited weight-gradient commits of the backward
    uint32_t tmem, lane_base, sbase, phase;  // sbase: CTA smem base >> 4
    bool w0;         // this thread is in the group's first warp (shuffle broadcast: warp-uniform for the compiler)
};

// part A: nothing here depends on the parameters (runs ahead of pdl_wait()).  `smem` is the group's own map; TMEM is
// allocated once per CTA by its first warp (tmem_cols covers every group) into group 0's slot.
template <int HV = 1>
__device__ __forceinline__ void tc_setup_begin(unsigned char* smem, uint32_t tmem_cols) {
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    uint32_t* tmem_s = reinterpret_cast<uint32_t*>(smem + SM_BAR + 16);
    for (int i = tc_tid(); i < 2 * CHUNK_B / 16; i += TC_THREADS)  // G image (chunk 1 stays zero forever)
        *reinterpret_cast<uint4*>(smem + SM_G + i * 16) = make_uint4(0, 0, 0, 0);
    if ((threadIdx.x >> 5) == 0) tc::tmem_alloc(tmem_s, tmem_cols);
    if ((threadIdx.x & (HV * TC_THREADS - 1)) == 0) {
        tc::mbar_init(bar, 1);
        tc::mbar_init(bar + 1, 1);
        tc::mbar_init(bar + 3, 1);
        tc::fence_barrier_init();
    }
}

// the group's copy of the fp16 weight image: all loads first (one L2 round trip), then the stores
__device__ __forceinline__ void tc_load_image(unsigned char* smem, const unsigned char* image) {
    const uint4* src = reinterpret_cast<const uint4*>(image);
    uint4* dst = reinterpret_cast<uint4*>(smem + SM_IMG);
    constexpr int N16 = opt::IMG_BYTES / 16, PER = (N16 + TC_THREADS - 1) / TC_THREADS;
    uint4 v[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const int i = tc_tid() + u * TC_THREADS;
        if (i < N16) v[u] = __ldcg(src + i);
    }
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const int i = tc_tid() + u * TC_THREADS;
        if (i < N16) dst[i] = v[u];
    }
}

// part B: weight image (straight 16-byte copies; the image is rebuilt by the optimizer step, not here), then publish.
// tmem_slot: where the CTA's TMEM base was written (the group's own map for one-group kernels); this group owns
// TM_COLS_BWD columns from base + tmem_col0.
template <int HV = 1>
__device__ __forceinline__ TcSetup tc_setup_finish(unsigned char* smem, const unsigned char* image, const uint32_t* tmem_slot = nullptr,
                                                   uint32_t tmem_col0 = 0) {
    TcSetup s;
    s.bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    if (!tmem_slot) tmem_slot = reinterpret_cast<const uint32_t*>(smem + SM_BAR + 16);
    const int warp = tc_tid() >> 5;
    tc_load_image(smem, image);  // (HV = 2: both halves store the same values)
    tc_sync_for_mma<HV>();
    s.tmem = __shfl_sync(0xffffffffu, *tmem_slot + tmem_col0, 0);
    s.lane_base = s.tmem + ((uint32_t)(warp * 32) << 16);
    s.sbase = __shfl_sync(0xffffffffu, tc::smem_u32(smem) >> 4, 0);  // descriptor units
    s.phase = 0;
    s.w0 = tc_warp_u<HV>() == 0;
    return s;
}

__device__ __forceinline__ void tc_wait(TcSetup& s) {
    tc::mbar_wait(s.bar, s.phase);
    s.phase ^= 1;
    tc::fence_after_sync();
}

// [X|1] -> H1 -> H2 -> OUT for the tile in the X image at xoff; leaves OUT in TMEM (waited)
struct TcNoHook { __device__ __forceinline__ void operator()(int) const {} };
// in_wait(stage): run by every thread between the issue of layer `stage`'s MMAs and the wait for them (the epoch kernel
// slips the next tile's record gather into these waits)
template <int HV = 1, typename Hook = TcNoHook>
__device__ __forceinline__ void tc_forward_tile(unsigned char* smem, TcSetup& s, int xoff, int trow = -1, int half = 0, Hook in_wait = Hook()) {
    constexpr uint32_t ID64 = tc::make_idesc_f16(128, 64, 0, 0, 0, 0);
    constexpr uint32_t ID16 = tc::make_idesc_f16(128, 16, 0, 0, 0, 0);
    TRACE(trow, 1);
    tc_sync_for_mma<HV>();
    TRACE(trow, 2);
    if (tc_issuer(s.w0)) {
#pragma unroll
        for (int ks = 0; ks < TC_XF / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, xoff, ks), desc_w_k(s.sbase, SM_W1, TC_H, ks), ID64, ks > 0);
        tc::commit(s.bar);
    }
    TRACE(trow, 3);
    in_wait(0);
    tc_wait(s);
    TRACE(trow, 4);
    tc_epilogue_relu<HV>(smem, s.lane_base, SM_H1, half);
    TRACE(trow, 5);
    tc_sync_for_mma<HV>();
    TRACE(trow, 6);
    if (tc_issuer(s.w0)) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_H1, ks), desc_w_k(s.sbase, SM_W2, TC_H, ks), ID64, ks > 0);
        tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, xoff, 1), desc_w_k(s.sbase, SM_B2I, TC_H, 0), ID64, 1);  // + b2
        tc::commit(s.bar);
    }
    TRACE(trow, 7);
    in_wait(1);
    tc_wait(s);
    TRACE(trow, 8);
    tc_epilogue_relu<HV>(smem, s.lane_base, SM_H2, half);
    TRACE(trow, 9);
    tc_sync_for_mma<HV>();
    TRACE(trow, 10);
    if (tc_issuer(s.w0)) {
#pragma unroll
        for (int ks = 0; ks < TC_H / 16; ++ks)
            tc::mma_f16(s.tmem + TM_OUT, desc_act_k(s.sbase, SM_H2, ks), desc_w_k(s.sbase, SM_W3, TC_OF, ks), ID16, ks > 0);
        tc::mma_f16(s.tmem + TM_OUT, desc_act_k(s.sbase, xoff, 1), desc_w_k(s.sbase, SM_B3I, TC_OF, 0), ID16, 1);  // + b3
        tc::commit(s.bar);
    }
    TRACE(trow, 11);
    in_wait(2);
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
    // Rollout chain (collect.cu): launched with programmatic stream serialization behind the world step that writes this
    // step's observations, so TMEM allocation, barrier setup and the weight image load above run under that kernel's tail.
    // The successor (the world step reading the actions) may be scheduled from here on: its own prologue reads state that the
    // PREVIOUS world step wrote, which has completed once this wait returns.
    pdl_wait();
    pdl_launch_dependents();
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
                const Soft sm = softmax_row<TC_FAST_LOSS>(z, a.n_act, a.probs != 0);
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
    const int32_t* slots;          // optional slot table: slots[p] = record slot of batch position p (tsm_ppo_slot_table)
    BatchView view;
    int64_t mb_start, mb_len;
    const float* mb_stats;
    PpoCfg cfg;
    int in_dim, out_dim, n_act, value_index;
    float gscale, inv_gscale;  // S and 1/S (power of two)
    double* loss_sums;
    int debug;  // experiments only (TSM_TC_DEBUG): 1 skip flush, 2 skip backward GEMM phases, 4 skip the loss math
};

// The batch gather.  The 96-byte records of the NEXT tile are fetched a whole tile ahead of their use by TMA gather4
// copies (four rows per instruction, issued by every fourth lane with the slots of its three neighbours) into the group's
// record staging area (SM_REC).  Plain loads did not work as a prefetch here: the records of a tile sit on ~128 different
// pages, and a warp stalls at ISSUE of a global load until its address is translated (1.7k - 4.6k cycles per tile measured,
// against ~600 for contiguous records); one bulk copy per row costs as much, because the instruction takes uniform
// operands and the compiler serialises the 32 lanes of a warp.  With a slot table (TcPpoArgs::slots) the table entry is
// fetched one more tile ahead, so neither the permutation arithmetic nor a dependent load sits on the tile's critical path.
struct RowScalars {
    float adv, ret, logp_old, v_old;
    int act;
};

// slot of batch position p: from the table when there is one, else computed here
__device__ __forceinline__ int32_t tc_slot(const TcPpoArgs& a, int64_t p) {
    if (a.slots) return __ldg(a.slots + p);
    return (int32_t)((a.debug & 8) ? p : slot_of(a.view, p));
}

// issue (all 128 threads of the group): slot = record of this thread's row, any valid slot (0) for a row past the end
__device__ __forceinline__ void tc_gather_issue(unsigned char* smem, const CUtensorMap* tmap, int32_t slot) {
    uint64_t* rbar = reinterpret_cast<uint64_t*>(smem + SM_BAR) + 3;
    const int lane = tc_tid() & 31, base = lane & ~3;
    const int s0 = __shfl_sync(0xffffffffu, slot, base), s1 = __shfl_sync(0xffffffffu, slot, base + 1);
    const int s2 = __shfl_sync(0xffffffffu, slot, base + 2), s3 = __shfl_sync(0xffffffffu, slot, base + 3);
    if (tc_tid() == 0) tc::mbar_expect_tx(rbar, TC_ROWS * REC_BYTES);
    if ((lane & 3) == 0) tc::tma_gather4(smem + SM_REC + tc_tid() * REC_BYTES, tmap, 0, s0, s1, s2, s3, rbar);
}

// take: wait for the tile's records, move this thread's row into the X image at xoff (fp16 chunks, as packed) and its scalars
// into registers.  Afterwards the thread's staging row is free for the next issue (it is only ever touched by this thread).
// HV = 2: half h moves chunks 2h, 2h + 1 of the row; only half 0 reads the scalars.  The staging row is then shared by two
// threads: the next issue has to come after a group barrier (tc_forward_tile's after_sync hook).
template <int HV = 1>
__device__ __forceinline__ void tc_gather_take(unsigned char* smem, int xoff, bool valid, uint32_t& rphase, RowScalars& d, int half = 0) {
    uint64_t* rbar = reinterpret_cast<uint64_t*>(smem + SM_BAR) + 3;
    tc::mbar_wait(rbar, rphase);
    rphase ^= 1;
    const uint4* rec = reinterpret_cast<const uint4*>(smem + SM_REC + tc_tid() * REC_BYTES);
    const uint4 zero = make_uint4(0, 0, 0, 0);
    constexpr int NC = 4 / HV;
    const int c0 = HV == 1 ? 0 : NC * half;
    uint4 x[NC], sc = zero;
    uint32_t vo = 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) x[c] = valid ? rec[c0 + c] : zero;
    if (valid && half == 0) {
        sc = rec[4];
        vo = *reinterpret_cast<const uint32_t*>(rec + 5);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) *reinterpret_cast<uint4*>(smem + xoff + (c0 + c) * CHUNK_B + tc_tid() * 16) = x[c];
    d.adv = __uint_as_float(sc.x); d.ret = __uint_as_float(sc.y); d.logp_old = __uint_as_float(sc.z);
    d.act = (int)sc.w;
    d.v_old = __uint_as_float(vo);
}

// Staging copy of the flat gradient (fp32, in the group's own smem at SM_STAGE).  The W2 block is XOR-swizzled at
// float4 granularity so that both the per-lane row writes (stride 64 floats) and the linear 16-byte read-out are
// bank-conflict free.  Every block of the flat layout up to b3 starts on a multiple of 4 floats.
__device__ __forceinline__ int stage_pos4(int i, int w2_lo, int w2_hi) {  // i = 4 * j: float offset of logical float4 j
    if (i >= w2_lo && i < w2_hi) {
        const int rel = i - w2_lo, rr = rel >> 6, c4 = (rel & 63) >> 2;
        return w2_lo + (rr << 6) + ((c4 ^ (rr & 15)) << 2);
    }
    return i;
}

// TMEM weight-gradient accumulators of this group -> its staging copy
template <int HV = 1>
__device__ __forceinline__ void tc_stage_grads(unsigned char* smem, const TcSetup& s, const Offsets& o, int in_dim, int out_dim, int half = 0) {
    float* stage = reinterpret_cast<float*>(smem + SM_STAGE);
    const int r = tc_tid();
    if constexpr (HV == 2) {  // the same map, each half of a row's columns by its own thread
        if (r < TC_H) {
            float v[32];
            tc::tmem_ld32(s.lane_base + TM_GW2 + 32 * half, v);  // dW2[r][32 half ..]
#pragma unroll
            for (int h = 0; h < 8; ++h)
                *reinterpret_cast<float4*>(stage + o.w2 + (r << 6) + (((8 * half + h) ^ (r & 15)) << 2)) =
                    make_float4(v[4 * h], v[4 * h + 1], v[4 * h + 2], v[4 * h + 3]);
            float w[16];
            if (half == 0) {
                tc::tmem_ld16(s.lane_base + TM_GW3, w);  // dW3^T[r][o]
#pragma unroll
                for (int k = 0; k < OUTP; ++k)
                    if (k < out_dim) stage[o.w3 + k * TC_H + r] = w[k];
            } else {
                tc::tmem_ld16(s.lane_base + TM_GW1 + 16, w);  // D2^T [X|1], column 31 = db2[r]
                stage[o.b2 + r] = w[15];
            }
        } else {
            float v[16];
            tc::tmem_ld16(s.lane_base + TM_GW1 + 16 * half, v);  // dW1[r-64][16 half ..], db1[r-64] (column 31)
            const int rr = r - TC_H;
#pragma unroll
            for (int k = 0; k < 16; ++k)
                if (16 * half + k < in_dim) stage[o.w1 + rr * in_dim + 16 * half + k] = v[k];
            if (half == 1) stage[o.b1 + rr] = v[15];
        }
        return;
    }
    if (r < TC_H) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {  // dW2[r][32q..]
            float v[32];
            tc::tmem_ld32(s.lane_base + TM_GW2 + 32 * q, v);
#pragma unroll
            for (int h = 0; h < 8; ++h)
                *reinterpret_cast<float4*>(stage + o.w2 + (r << 6) + (((8 * q + h) ^ (r & 15)) << 2)) =
                    make_float4(v[4 * h], v[4 * h + 1], v[4 * h + 2], v[4 * h + 3]);
        }
        float v[16];
        tc::tmem_ld16(s.lane_base + TM_GW3, v);  // dW3^T[r][o]
#pragma unroll
        for (int k = 0; k < OUTP; ++k)
            if (k < out_dim) stage[o.w3 + k * TC_H + r] = v[k];
        tc::tmem_ld16(s.lane_base + TM_GW1 + 16, v);  // D2^T [X|1], column 31 = db2[r]
        stage[o.b2 + r] = v[15];
    } else {
        float v[32];
        tc::tmem_ld32(s.lane_base + TM_GW1, v);  // dW1[r-64][k], db1[r-64] (column 31)
        const int rr = r - TC_H;
#pragma unroll
        for (int k = 0; k < 32; ++k)
            if (k < in_dim) stage[o.w1 + rr * in_dim + k] = v[k];
        stage[o.b1 + rr] = v[31];
    }
}

__device__ __forceinline__ void red_add4(float* dst, float4 v) {  // one 16-byte reduction at L2, no return value
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void __launch_bounds__(TC_THREADS) tc_ppo_minibatch_kernel(TcPpoArgs a, const __grid_constant__ CUtensorMap rec_map) {
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

    __syncthreads();  // barriers initialised before any thread's bulk copy names them
    uint32_t rphase = 0;
    int64_t tile = blockIdx.x;
    if (tile < n_tiles) {  // batch records only: safe ahead of the dependency wait
        const int64_t q = tile * TC_ROWS + r;
        tc_gather_issue(smem, &rec_map, q < a.mb_len ? tc_slot(a, a.mb_start + q) : 0);
    }
    int32_t sl_next = 0;  // table path: slot of this thread's row in tile + gridDim.x, fetched a tile before its gather is issued
    bool v_next = false;
    if (a.slots) {
        const int64_t q = (tile + gridDim.x) * TC_ROWS + r;
        v_next = q < a.mb_len;
        if (v_next) sl_next = __ldg(a.slots + a.mb_start + q);
    }
    TRACE(krow, 32);
    pdl_wait();  // parameters (image), gradients and mb_stats of the preceding launches are now visible
    TcSetup s = tc_setup_finish(smem, a.image);
    TRACE(krow, 33);
    float mean = 0.f, rstd = 1.f;
    if ((a.cfg.adv_norm || a.cfg.plain_pg == 2) && a.mb_stats) { mean = __ldcg(a.mb_stats); rstd = __ldcg(a.mb_stats + 1); }
    uint32_t it = 0;
    for (; tile < n_tiles; tile += gridDim.x, ++it) {
        const bool valid = tile * TC_ROWS + r < a.mb_len;
        const int xoff = SM_X + (int)(it & 1) * 4 * CHUNK_B;
        const int trow = (it < 2 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? (blockIdx.x != 0) * 2 + (int)it : -1;
        TRACE(trow, 0);
        RowScalars d;
        tc_gather_take(smem, xoff, valid, rphase, d);
        const float d_adv = d.adv, d_ret = d.ret, d_logp_old = d.logp_old, d_v_old = d.v_old;
        const int d_act = d.act;
        if (tile + gridDim.x < n_tiles) {
            const int64_t q = (tile + gridDim.x) * TC_ROWS + r;
            const bool nv = q < a.mb_len;
            tc_gather_issue(smem, &rec_map, !nv ? 0 : (a.slots ? sl_next : tc_slot(a, a.mb_start + q)));
            if (a.slots) {
                const int64_t q2 = (tile + 2 * (int64_t)gridDim.x) * TC_ROWS + r;
                v_next = q2 < a.mb_len;
                if (v_next) sl_next = __ldg(a.slots + a.mb_start + q2);
            }
        }
        tc_forward_tile(smem, s, xoff, trow);  // its first wait also covers the previous tile's weight-gradient GEMMs
        // ---- loss and G = S * d(loss)/d(OUT) for this thread's row ----
        float z[OUTP], dz[OUTP];
        tc_read_out(s, z);
        RowLoss rl{0.f, 0.f, 0.f};
        if (a.debug & 4) {
#pragma unroll
            for (int j = 0; j < OUTP; ++j) dz[j] = z[j] * inv_n;
        } else {
            rl = row_loss<TC_FAST_LOSS>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz);
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
            if (tc_issuer(s.w0)) tc::commit(s.bar + 1);
            continue;
        }
        TRACE(trow, 13);
        tc_sync_for_mma();
        TRACE(trow, 14);
        if (tc_issuer(s.w0)) {
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
        if (tc_issuer(s.w0)) {
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
        if (tc_issuer(s.w0)) {
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
        // ---- flush: TMEM gradient accumulators -> smem staging (flat gradient layout) -> coalesced 16-byte reductions ----
        tc::mbar_wait(s.bar + 1, (it - 1) & 1);
        tc::fence_after_sync();
        tc_stage_grads(smem, s, o, a.in_dim, a.out_dim);
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
        for (int i = 4 * r; i < n_stage; i += 4 * TC_THREADS) {
            float4 v = *reinterpret_cast<const float4*>(stage + stage_pos4(i, w2_lo, w2_hi));
            v.x *= inv_s; v.y *= inv_s; v.z *= inv_s; v.w *= inv_s;
            red_add4(a.grads + i, v);
        }
        if (r < a.out_dim) atomicAdd(a.grads + o.b3 + r, red_f[r] + red_f[OUTP + r] + red_f[2 * OUTP + r] + red_f[3 * OUTP + r]);
        if (a.loss_sums && r < 3) atomicAdd(a.loss_sums + r, red_d[r] + red_d[3 + r] + red_d[6 + r] + red_d[9 + r]);
    }
    TRACE(krow, 38);
}

// ---------------------------------------------------------------------------------------------
// One PERSISTENT cooperative launch per epoch (uniform minibatch size).  Same tile pipeline as tc_ppo_minibatch_kernel,
// but one CTA per SM holding TC_GROUPS independent 128-thread groups (each with its own shared-memory map, TMEM columns
// and mbarriers -- two groups per SM overlap one tile's CUDA-core phases with the other's tensor-core phases, the job
// two co-resident CTAs do in the launch-per-minibatch path; the driver's occupancy calculator only ever admits ONE
// cooperative CTA per SM for a kernel that allocates TMEM).  Every group keeps its TMEM / barriers / smem across all
// minibatches and the optimizer step runs inside the kernel, spread over all CTAs, between two grid-wide barriers:
//   tiles of minibatch m -> the groups' staged gradients are summed in shared memory -> 16-byte reductions into
//                           gradient buffer m % 3; the CTA's copy of params / exp_avg / exp_avg_sq is fetched into registers
//   grid.sync -> EVERY CTA runs the WHOLE optimizer step on its own copy of the state: squared norm of the gradient,
//                clip coefficient, Adam for all n parameters, fp16 image entries written straight into both groups'
//                shared-memory images.  Same instructions on the same values in every CTA -> bit-identical copies, so
//                there is no second grid barrier, no image round trip through L2 and no slice ownership.  The copy
//                never touches memory: a thread owns 7 float4 of the flat parameter vector, keeps the fp32 masters in
//                registers and exp_avg / exp_avg_sq in the 64 TMEM columns its group leaves unused (148 copies kept in
//                L2 cost 20-30 MB of traffic per step: measured 2x slower than the two-barrier version).
//                Buffer (m - 1) % 3, which nobody reads any more, is cleared by slice owners off the critical path;
//                CTA 0 writes the net's flat arrays and global image after the last minibatch.
// No launch, TMEM allocation, barrier init or cold record gather per minibatch (the first tile of minibatch m + 1 is
// prefetched before the flush of m).  cooperative_groups grid barriers; the launch is refused unless all CTAs are
// co-resident (cudaLaunchCooperativeKernel).
struct TcEpochArgs {
    TcPpoArgs p;           // mb_start = first batch position of the epoch; grads = gradient buffer 0
    float* grads1;         // gradient buffers 1 and 2 (minibatch m reduces into buffer m % 3; all zero on entry and on exit)
    float* grads2;
    uint2* q_table;        // [threads][float4 per thread + 1] scratch (see the kernel): EP_Q_WORDS entries
    float* g_total;        // [EP_G_STRIDE] sum over ranks (world > 1 only)
    uint32_t* sync_words;  // [0], [1] arrival counters of the in-kernel grid barriers (zero between launches)
    int64_t n_mb, mb_size, total;  // minibatch m = [m * mb_size, (m + 1) * mb_size), the last one runs to `total`
    const float* mb_stats;         // [n_mb][2] or NULL
    double* loss_sums;             // [n_mb][4] or NULL
    opt::AdamArgs adam;
};

constexpr int EP_THREADS = TC_THREADS * TC_GROUPS;  // HV = 1; the kernel runs HV * EP_THREADS threads
constexpr int EP_HV = 2;                            // threads per tile row of the default instantiation
constexpr int SM_GROUP = (SM_TOTAL + 1023) & ~1023;   // stride between the groups' maps
constexpr int SM_EP_SHARED = TC_GROUPS * SM_GROUP;    // CTA-wide scratch: 8 norm partials, 3 coefficients, TC_GROUPS flags
constexpr int SM_EP_TOTAL = SM_EP_SHARED + 128;
constexpr int EP_G_STRIDE = 7168;                     // floats per gradient buffer (largest net: 6 728 parameters)
constexpr int EP_OPT = EP_G_STRIDE / 4 / EP_THREADS;  // HV = 1: 7 float4 of the flat parameter vector per thread: j = t + u * EP_THREADS
constexpr int EP_OPT2 = 4;                            // HV = 2: 4 float4 per thread (512 threads; entries past n are dead)
constexpr int EP_Q_WORDS = 2 * EP_THREADS * (EP_OPT2 + 1) > EP_THREADS * (EP_OPT + 1) ? 2 * EP_THREADS * (EP_OPT2 + 1) : EP_THREADS * (EP_OPT + 1);  // q_table entries
// Spare TMEM columns of a group (lanes = the thread's row).  HV = 1: exp_avg in 28 of the 32 columns at TM_OPT_M, exp_avg_sq in
// 16 + 12 columns at V0 / V1.  HV = 2: 16 + 16 columns per thread: half 0 exp_avg at M, exp_avg_sq at V0; half 1 at M + 16 / V1.
constexpr uint32_t TM_OPT_M = 208;
constexpr uint32_t TM_OPT_V0 = 240, TM_OPT_V1 = 80;
static_assert(EP_OPT == 7, "column map below is for 7 float4 per thread");
static_assert(4 * EP_OPT2 * 2 * EP_THREADS >= EP_G_STRIDE, "HV = 2 must cover the whole gradient buffer");

// 32-bit opt::image_index
__device__ __forceinline__ int ep_image_index(int i, const Offsets& o, int in_dim) {
    if (i < (int)o.b1) {
        const int out = i / in_dim, in = i - out * in_dim;
        return opt::IMG_W1 + ((in >> 3) * 64 + out) * 8 + (in & 7);
    }
    if (i < (int)o.w2) return opt::IMG_W1 + (3 * 64 + (i - (int)o.b1)) * 8 + 7;
    if (i < (int)o.b2) {
        const int rel = i - (int)o.w2;
        return opt::IMG_W2 + (((rel & 63) >> 3) * 64 + (rel >> 6)) * 8 + (rel & 7);
    }
    if (i < (int)o.w3) return opt::IMG_B2 + (64 + (i - (int)o.b2)) * 8 + 7;
    if (i < (int)o.b3) {
        const int rel = i - (int)o.w3;
        return opt::IMG_W3 + (((rel & 63) >> 3) * 16 + (rel >> 6)) * 8 + (rel & 7);
    }
    return opt::IMG_B3 + (16 + (i - (int)o.b3)) * 8 + 7;
}
// exp_avg / exp_avg_sq of float4 u: 4 TMEM columns each
__device__ __forceinline__ void ep_ld4(uint32_t taddr, float4& v) {
    uint32_t r0, r1, r2, r3;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n\ttcgen05.wait::ld.sync.aligned;"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(taddr) : "memory");
    v = make_float4(__uint_as_float(r0), __uint_as_float(r1), __uint_as_float(r2), __uint_as_float(r3));
}
__device__ __forceinline__ void ep_st4(uint32_t taddr, const float4& v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(__float_as_uint(v.x)),
                 "r"(__float_as_uint(v.y)), "r"(__float_as_uint(v.z)), "r"(__float_as_uint(v.w)) : "memory");
}
template <int HV>
__device__ __forceinline__ uint32_t ep_m_col(int u, int half) { return HV == 1 ? TM_OPT_M + 4 * u : TM_OPT_M + 16 * half + 4 * u; }
template <int HV>
__device__ __forceinline__ uint32_t ep_v_col(int u, int half) {
    if (HV == 1) return u < 4 ? TM_OPT_V0 + 4 * u : TM_OPT_V1 + 4 * (u - 4);
    return (half ? TM_OPT_V1 : TM_OPT_V0) + 4 * u;
}

template <int HV>
__global__ void __launch_bounds__(HV * EP_THREADS, 1) tc_ppo_epoch_kernel(TcEpochArgs e, const __grid_constant__ CUtensorMap rec_map) {
    extern __shared__ __align__(1024) unsigned char smem_all[];
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    constexpr int EPT = HV * EP_THREADS;           // threads of the CTA
    constexpr int OPT = HV == 1 ? EP_OPT : EP_OPT2;  // float4 of the flat parameter vector per thread: j = t + u * EPT
    const TcPpoArgs& a = e.p;  // (never written: the argument struct stays in the constant bank)
    const int grp = (int)(threadIdx.x / (HV * TC_THREADS)), r = tc_tid(), rc = (int)threadIdx.x;
    const int half = HV == 1 ? 0 : (int)((threadIdx.x >> 7) & 1);  // warp-uniform
    unsigned char* smem = smem_all + grp * SM_GROUP;
    float* s_red = reinterpret_cast<float*>(smem_all + SM_EP_SHARED);
    float* s_coef = s_red + EPT / 32;
    int* s_has = reinterpret_cast<int*>(s_coef + 4);
    static_assert(EPT / 32 + 4 + TC_GROUPS <= 32, "CTA-wide scratch is 128 bytes");
    const int64_t vb = (int64_t)blockIdx.x * TC_GROUPS + grp, vg = (int64_t)gridDim.x * TC_GROUPS;  // virtual CTA = group
    tc_setup_begin<HV>(smem, TM_COLS_BWD * TC_GROUPS);
    constexpr uint32_t ID_DH = tc::make_idesc_f16(128, 64, 0, 0, 0, 1);
    constexpr uint32_t ID_G64 = tc::make_idesc_f16(128, 64, 0, 0, 1, 1);
    constexpr uint32_t ID_G32 = tc::make_idesc_f16(128, 32, 0, 0, 1, 1);
    constexpr uint32_t ID_G16 = tc::make_idesc_f16(128, 16, 0, 0, 1, 1);
    const Offsets o = offsets(a.in_dim, TC_H, a.out_dim);
    const int64_t epoch_start = a.mb_start;
    const bool simple_loss = a.cfg.probs_param == 0 && a.cfg.value_heads <= 1 && a.cfg.plain_pg == 0 && a.n_act > 0 && a.value_index >= 0;
    const bool simple5 = simple_loss && a.n_act == 5 && a.value_index == 5;
    const opt::AdamArgs& ad = e.adam;
    const int64_t n = ad.n;
    const int64_t step0 = *ad.step;  // rewritten by CTA 0 only after the last grid barrier
    // this group's tile sequence over the whole epoch: (m, tile) -> next (m, tile); false at the end.  Tile counts and
    // in-minibatch row numbers in 32 bits (this bookkeeping sits on the tile's dependent chain); positions in 64.
    const int n_mb32 = (int)e.n_mb, vb32 = (int)vb, vg32 = (int)vg;
    const int len_full = (int)e.mb_size, len_last = (int)(e.total - (e.n_mb - 1) * e.mb_size);
    const int tiles_full = (len_full + TC_ROWS - 1) / TC_ROWS, tiles_last = (len_last + TC_ROWS - 1) / TC_ROWS;
    auto advance = [&](int64_t& m_, int64_t& t_) -> bool {
        const int m32 = (int)m_, t32 = (int)t_ + vg32;
        if (t32 < (m32 == n_mb32 - 1 ? tiles_last : tiles_full)) { t_ = t32; return true; }
        for (int mm = m32 + 1; mm < n_mb32; ++mm)
            if (vb32 < (mm == n_mb32 - 1 ? tiles_last : tiles_full)) { m_ = mm; t_ = vb32; return true; }
        return false;
    };
    auto pos_of = [&](int64_t m_, int64_t t_, bool& valid) -> int64_t {  // batch position of this thread's row
        const int q = (int)t_ * TC_ROWS + r;
        valid = q < ((int)m_ == n_mb32 - 1 ? len_last : len_full);
        return epoch_start + m_ * e.mb_size + q;
    };
    auto len_of = [&](int64_t m_) { return m_ == e.n_mb - 1 ? e.total - m_ * e.mb_size : e.mb_size; };
    uint32_t rphase = 0;
    __syncthreads();  // barriers initialised before any thread's bulk copy names them
    int64_t nm = 0, nt = vb;       // the tile whose records are in flight
    bool n_ok = vb * TC_ROWS < len_of(0);
    if (!n_ok) n_ok = advance(nm, nt);
    // HV = 2: the record prefetch (tile bookkeeping, slot loads and the gather itself) belongs to the half-1 warps
    const bool pf = HV == 1 || half == 1;
    if (n_ok && pf) {
        bool v;
        const int64_t p = pos_of(nm, nt, v);
        tc_gather_issue(smem, &rec_map, v ? tc_slot(a, p) : 0);
    }
    int64_t am = nm, at = nt;      // table path: the tile after that, whose slots are in flight (sl_next / v_next)
    bool a_ok = n_ok && advance(am, at);
    int32_t sl_next = 0;
    bool v_next = false;
    if (a.slots && a_ok && pf) {
        const int64_t p = pos_of(am, at, v_next);
        if (v_next) sl_next = __ldg(a.slots + p);
    }
    tc::fence_before_sync();
    __syncthreads();  // the CTA's TMEM base (group 0's slot) is visible to every group
    tc::fence_after_sync();
    TcSetup s = tc_setup_finish<HV>(smem, a.image, reinterpret_cast<const uint32_t*>(smem_all + SM_BAR + 16), grp * TM_COLS_BWD);
    uint32_t xb = 0, wg_commits = 0;  // X buffer toggle and number of commits to bar[1] so far (its phase parity)
    // ---- this CTA's copy of the optimizer state: fp32 masters in registers, moments in the group's spare TMEM columns ----
    float4 Pr[OPT];
    // image positions of each float4's four parameters (0xffff: none) and, in word EP_OPT, bit u = "four consecutive
    // positions" (W2 / W3 rows: one 8-byte store).  Computed once and parked in global memory (the same values from
    // every CTA; a thread only reads back what it wrote itself): 15 registers less across the tile loop.
    uint2* q_tab = e.q_table + (size_t)rc * (OPT + 1);
    {
    uint2 Q[OPT];
    uint32_t q_contig = 0;
#pragma unroll
    for (int u = 0; u < OPT; ++u) {
        const int i = 4 * (rc + u * EPT);
        {
            uint32_t q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = i + k < n ? (uint32_t)ep_image_index(i + k, o, a.in_dim) : 0xffffu;
            Q[u] = make_uint2(q[0] | (q[1] << 16), q[2] | (q[3] << 16));
            if (q[0] != 0xffffu && (q[0] & 3) == 0 && q[1] == q[0] + 1 && q[2] == q[0] + 2 && q[3] == q[0] + 3) q_contig |= 1u << u;
        }
        auto get = [&](const float* src, int k) { return i + k < n ? __ldcg(src + i + k) : 0.f; };
        Pr[u] = make_float4(get(ad.p, 0), get(ad.p, 1), get(ad.p, 2), get(ad.p, 3));
        ep_st4(s.lane_base + ep_m_col<HV>(u, half), make_float4(get(ad.m, 0), get(ad.m, 1), get(ad.m, 2), get(ad.m, 3)));
        ep_st4(s.lane_base + ep_v_col<HV>(u, half), make_float4(get(ad.v, 0), get(ad.v, 1), get(ad.v, 2), get(ad.v, 3)));
    }
#pragma unroll
    for (int u = 0; u < OPT; ++u) q_tab[u] = Q[u];
    q_tab[OPT] = make_uint2(q_contig, 0u);
    }
    tc::tmem_st_wait();

    for (int64_t m = 0; m < e.n_mb; ++m) {
        const int64_t mb_len = (m == e.n_mb - 1) ? e.total - m * e.mb_size : e.mb_size;
        float* g_cur = (m % 3 == 0) ? a.grads : (m % 3 == 1 ? e.grads1 : e.grads2);
        [[maybe_unused]] const int erow = (m == 2 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? (blockIdx.x != 0) * 2 : -1;
        TRACE(erow, 40);
        const int64_t n_tiles = (mb_len + TC_ROWS - 1) / TC_ROWS;
        const float inv_n = 1.0f / (float)mb_len;
        float gscale = 1.0f;
        while (gscale < (float)mb_len && gscale < 1048576.0f) gscale *= 2.0f;
        const float inv_s = 1.0f / gscale;
        float mean = 0.f, rstd = 1.f;
        if ((a.cfg.adv_norm || a.cfg.plain_pg == 2) && e.mb_stats) { mean = __ldcg(e.mb_stats + 2 * m); rstd = __ldcg(e.mb_stats + 2 * m + 1); }
        float gb3[OUTP];
#pragma unroll
        for (int j = 0; j < OUTP; ++j) gb3[j] = 0.f;
        float f_clip = 0.f, f_vf = 0.f, f_ent = 0.f;
        double l_clip = 0.0, l_vf = 0.0, l_ent = 0.0;
        uint32_t it = 0;
        for (int64_t tile = vb; tile < n_tiles; tile += vg, ++it) {
            const bool valid = tile * TC_ROWS + r < mb_len;
            const int xoff = SM_X + (int)(xb & 1) * 4 * CHUNK_B;
            ++xb;
            [[maybe_unused]] const int trow = (erow >= 0 && it < 2) ? erow + (int)it : -1;
            TRACE(trow, 0);
            RowScalars d;
            tc_gather_take<HV>(smem, xoff, valid, rphase, d, half);
            const float d_adv = d.adv, d_ret = d.ret, d_logp_old = d.logp_old, d_v_old = d.v_old;
            const int d_act = d.act;
            TRACE(trow, 24);
            // gather of this group's next tile (possibly the first one of a later minibatch: it then overlaps the flush and
            // the barriers); with a slot table its entries were fetched a tile ago and the ones after are fetched now.
            // HV = 1: here, ahead of the forward.  HV = 2: by the half-1 warps after the forward (every thread of the group
            // is then long done with the staging rows), in the shadow of the loss stage.  (The 32 gather4 copies of a tile
            // keep the SM's TMA unit busy for ~900 cycles and block their issuing warps that long: spreading them over the
            // MMA waits of the other stages, one warp each, measured slower -- the late warp delays its stage.)
            auto prefetch_next = [&]() {
                TRACE_H1(trow == 0, 50);
                nm = am; nt = at; n_ok = a_ok;
                if (n_ok) {
                    bool v;
                    const int64_t p = pos_of(nm, nt, v);
                    TRACE_H1(trow == 0, 51);
                    tc_gather_issue(smem, &rec_map, !v ? 0 : (a.slots ? sl_next : tc_slot(a, p)));
                    TRACE_H1(trow == 0, 52);
                    a_ok = advance(am, at);
                    TRACE_H1(trow == 0, 53);
                    if (a.slots && a_ok) {
                        const int64_t p2 = pos_of(am, at, v_next);
                        if (v_next) sl_next = __ldg(a.slots + p2);
                    }
                    TRACE_H1(trow == 0, 54);
                }
            };
            if (HV == 1) prefetch_next();
            TRACE(trow, 25);
            tc_forward_tile<HV>(smem, s, xoff, trow, half);
            if (HV == 2 && half == 1) prefetch_next();
            if (half == 0) {  // the loss stage: one thread per row
            float z[OUTP], dz[OUTP];
            tc_read_out(s, z);
            // (three instantiations of the same arithmetic behind uniform branches: five actions + value, the shape of every
            // discrete MPE scenario; plain PPO of any shape; everything else)
            const RowLoss rl = simple5 ? row_loss<TC_FAST_LOSS, true, 5>(z, 5, 5, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz)
                             : simple_loss ? row_loss<TC_FAST_LOSS, true>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz)
                                           : row_loss<TC_FAST_LOSS>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz);
            if (valid) { f_clip += rl.clip; f_vf += rl.vf; f_ent += rl.ent; }
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
                    g[j] = fminf(fmaxf(dz[j] * gscale, -65000.f), 65000.f);
                }
                uint4 w;
                w.x = pack_h2(g[0], g[1]); w.y = pack_h2(g[2], g[3]); w.z = pack_h2(g[4], g[5]); w.w = pack_h2(g[6], g[7]);
                *reinterpret_cast<uint4*>(smem + SM_G + r * 16) = w;
            }
            }
            TRACE(trow, 13);
            tc_sync_for_mma<HV>();
            TRACE(trow, 14);
            // The three weight-gradient products are issued as soon as their operands exist, each behind the waited MMA of
            // its stage (commit first: the wait does not cover them): the issuing lane spends the time it would otherwise
            // wait, instead of 24 MMA issues at the end of the tile with the rest of the group idle.
            const uint32_t acc0 = it > 0;
            if (tc_issuer(s.w0)) {
                tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_G, 0), desc_w_mn(s.sbase, SM_W3, TC_OF, 0), ID_DH, 0);
                tc::commit(s.bar);
#pragma unroll
                for (int ks = 0; ks < TC_ROWS / 16; ++ks)  // [H2; H1]^T G: rows 0..63 = dW3^T
                    tc::mma_f16(s.tmem + TM_GW3, desc_act_mn(s.sbase, SM_H2, ks), desc_act_mn(s.sbase, SM_G, ks), ID_G16, ks > 0 ? 1u : acc0);
            }
            TRACE(trow, 15);
            tc_wait(s);
            TRACE(trow, 16);
            tc_epilogue_mask<HV>(smem, s.lane_base, SM_H2, SM_D2, half);
            TRACE(trow, 17);
            tc_sync_for_mma<HV>();
            TRACE(trow, 18);
            if (tc_issuer(s.w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_H / 16; ++ks)
                    tc::mma_f16(s.tmem + TM_SCR, desc_act_k(s.sbase, SM_D2, ks), desc_w_mn(s.sbase, SM_W2, TC_H, ks), ID_DH, ks > 0);
                tc::commit(s.bar);
                // [D2; D1]^T H1: rows 0..63 = dW2.  Rows 64..127 (the D1 half of the stacked operand, which the next stage is
                // about to write) land in accumulator lanes nobody reads.
#pragma unroll
                for (int ks = 0; ks < TC_ROWS / 16; ++ks)
                    tc::mma_f16(s.tmem + TM_GW2, desc_act_mn(s.sbase, SM_D2, ks), desc_act_mn(s.sbase, SM_H1, ks), ID_G64, ks > 0 ? 1u : acc0);
            }
            TRACE(trow, 19);
            tc_wait(s);
            TRACE(trow, 20);
            tc_epilogue_mask<HV>(smem, s.lane_base, SM_H1, SM_D1, half);
            TRACE(trow, 21);
            tc_sync_for_mma<HV>();
            TRACE(trow, 22);
            if (tc_issuer(s.w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_ROWS / 16; ++ks)
                    tc::mma_f16(s.tmem + TM_GW1, desc_act_mn(s.sbase, SM_D2, ks), desc_act_mn(s.sbase, xoff, ks), ID_G32, ks > 0 ? 1u : acc0);
                tc::commit(s.bar + 1);
            }
            ++wg_commits;
            TRACE(trow, 23);
        }
        TRACE(erow, 41);
        // ---- flush: each group stages its accumulators, the CTA adds the stagings and issues one set of reductions ----
        // (the register reductions first: they overlap the last tile's weight-gradient MMAs, which the staging has to wait for)
        l_clip += (double)f_clip; l_vf += (double)f_vf; l_ent += (double)f_ent;
        {
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
            if ((r & 31) == 0 && half == 0) {
                const int w = r >> 5;
#pragma unroll
                for (int j = 0; j < OUTP; ++j) red_f[w * OUTP + j] = gb3[j];
                red_d[w * 3 + 0] = l_clip; red_d[w * 3 + 1] = l_vf; red_d[w * 3 + 2] = l_ent;
            }
            if (r == 0 && half == 0) s_has[grp] = it > 0;
        }
        if (it > 0) {
            tc::mbar_wait(s.bar + 1, (wg_commits - 1) & 1);
            tc::fence_after_sync();
            tc_stage_grads<HV>(smem, s, o, a.in_dim, a.out_dim, half);
        }
        tc::fence_before_sync();
        __syncthreads();
        tc::fence_after_sync();
        {
            const int w2_lo = (int)o.w2, w2_hi = (int)o.b2, n_stage = (int)o.b3;
            bool has[TC_GROUPS];
            bool any = false;
#pragma unroll
            for (int g = 0; g < TC_GROUPS; ++g) { has[g] = s_has[g] != 0; any |= has[g]; }
            if (any) {
                for (int i = 4 * rc; i < n_stage; i += 4 * EPT) {
                    const int pos = stage_pos4(i, w2_lo, w2_hi);
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int g = 0; g < TC_GROUPS; ++g)
                        if (has[g]) {
                            const float4 u = *reinterpret_cast<const float4*>(smem_all + g * SM_GROUP + SM_STAGE + 4 * pos);
                            v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
                        }
                    v.x *= inv_s; v.y *= inv_s; v.z *= inv_s; v.w *= inv_s;
                    red_add4(g_cur + i, v);
                }
                if (rc < a.out_dim) {
                    float t = 0.f;
#pragma unroll
                    for (int g = 0; g < TC_GROUPS; ++g) {
                        const float* rf = reinterpret_cast<const float*>(smem_all + g * SM_GROUP + SM_RED);
                        t += rf[rc] + rf[OUTP + rc] + rf[2 * OUTP + rc] + rf[3 * OUTP + rc];
                    }
                    atomicAdd(g_cur + o.b3 + rc, t);
                }
                if (e.loss_sums && rc < 3) {
                    double t = 0.0;
#pragma unroll
                    for (int g = 0; g < TC_GROUPS; ++g) {
                        const double* rd = reinterpret_cast<const double*>(smem_all + g * SM_GROUP + SM_RED + 192);
                        t += rd[rc] + rd[3 + rc] + rd[6 + rc] + rd[9 + rc];
                    }
                    atomicAdd(e.loss_sums + 4 * m + rc, t);
                }
            }
        }
        TRACE(erow, 42);
        // ---- optimizer step: the whole of it in every CTA, on the CTA's own copy of the state ----
        // ---- grid barrier with the cross-rank gradient exchange folded into it ----
        // Arrive: one atomic per CTA on a device counter; every CTA waits until the counter shows all local arrivals: the
        // local gradient is complete.  With data-parallel ranks (ad.world > 1) every CTA then PUSHES its slice of it into
        // slot [seq & 1][rank] of every peer's receive area over NVLink as 16-byte stores of {value, seq, value, seq}:
        // every 8-byte (value, step) pair is self-validating, so no fence, release or flag follows the data (a release
        // after the stores costs a full NVLink round trip, a flag one more hop; one remote atomic per CTA cost ~10 us;
        // one CTA pushing everything is fine at 2 ranks but store-issue bound at 8: 322 KB from one SM).  The same CTA
        // polls the pairs of its slice from every peer.  Receive areas are double buffered by step parity: a peer
        // overwrites parity b at step seq + 2, after it has finished step seq + 1, which needed this rank's data of step
        // seq + 1, pushed only after every CTA here was done with step seq.
        uint2 Q[OPT];  // (in flight across the barrier)
#pragma unroll
        for (int u = 0; u < OPT; ++u) Q[u] = q_tab[u];
        const uint32_t q_contig = q_tab[OPT].x;
        const uint32_t seq = ad.seq + (uint32_t)m;
        __threadfence();
        __syncthreads();
        const uint32_t arrivals = (uint32_t)(m + 1) * gridDim.x;  // counter value once every CTA has arrived at this barrier
        if (rc == 0) atomicAdd(e.sync_words, 1u);
        if (rc == 64) {  // the bias corrections of this step (double-precision powers): in the shadow of the barrier wait
            const int64_t step = step0 + m + 1;
            s_coef[1] = ad.lr / (float)(1.0 - opt::ipow((double)ad.beta1, step));
            s_coef[2] = (float)(1.0 / sqrt(1.0 - opt::ipow((double)ad.beta2, step)));
        }
        if (rc == 0) {
            const uint32_t* w = e.sync_words;
            const uint32_t want = arrivals;
            const long long t0 = clock64();
            uint32_t v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(w) : "memory");
                if ((int32_t)(v - want) >= 0) break;
                if (clock64() - t0 > 4000000000LL) {  // ~2 s: a CTA or a peer rank is missing; report instead of hanging the device
                    if (ad.status) *ad.status = 1;
                    break;
                }
            } while (true);
        }
        __syncthreads();
        TRACE(erow, 43);
        const uint4* r_base = ad.world > 1 ? ad.peer_ll[ad.rank] + (int64_t)(seq & 1u) * ad.world * ad.n_stride / 2 : nullptr;
        const float* g_src = g_cur;  // where the (summed) gradient is read from below
        if (ad.world > 1) {
            // Slice owners wait for the (value, step) pairs of their slice from every peer, sum the `world` contributions
            // (rank order) into g_total, one more grid barrier, and everybody reads the total once.  Every CTA summing all slots itself would move world x 28 KB x grid per step,
            // and reads of the peer-written receive area are slow (they miss L2: +3 us per step at two ranks when all 148
            // CTAs read the whole slot, +1 us for one slice each); the extra barrier costs ~1.3 us.
            const int64_t n4s = ad.n_stride / 4, per4 = (n4s + gridDim.x - 1) / gridDim.x;
            const int64_t lo4 = (int64_t)blockIdx.x * per4, hi4 = lo4 + per4 < n4s ? lo4 + per4 : n4s;
            for (int64_t j = lo4 + rc; j < hi4; j += EPT) {
                float4 own;
                if (4 * j + 3 < n) {
                    own = __ldcg(reinterpret_cast<const float4*>(g_cur) + j);
                } else {
                    auto get = [&](int k) { return 4 * j + k < n ? __ldcg(g_cur + 4 * j + k) : 0.f; };
                    own = make_float4(get(0), get(1), get(2), get(3));
                }
                {   // this CTA's slice of the local gradient -> slot [seq & 1][rank] of every peer
                    const uint4 plo = make_uint4(__float_as_uint(own.x), seq, __float_as_uint(own.y), seq);
                    const uint4 phi = make_uint4(__float_as_uint(own.z), seq, __float_as_uint(own.w), seq);
                    const int64_t off = ((int64_t)(seq & 1u) * ad.world + ad.rank) * ad.n_stride + 4 * j;
                    for (int q = 0; q < ad.world; ++q)
                        if (q != ad.rank) {
                            uint4* dst = ad.peer_ll[q] + off / 2;
                            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(plo.x), "r"(plo.y), "r"(plo.z), "r"(plo.w) : "memory");
                            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 1), "r"(phi.x), "r"(phi.y), "r"(phi.z), "r"(phi.w) : "memory");
                        }
                }
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int q = 0; q < ad.world; ++q) {
                    float4 y = own;
                    if (q != ad.rank) {
                        const uint4* src = r_base + ((int64_t)q * ad.n_stride + 4 * j) / 2;
                        uint4 lo, hi;
                        const long long t0 = clock64();
                        while (true) {
                            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo.x), "=r"(lo.y), "=r"(lo.z), "=r"(lo.w) : "l"(src) : "memory");
                            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w) : "l"(src + 1) : "memory");
                            if (lo.y == seq && lo.w == seq && hi.y == seq && hi.w == seq) break;
                            if (clock64() - t0 > 4000000000LL) {  // ~2 s: a peer rank is missing; report instead of hanging the device
                                if (ad.status) *ad.status = 1;
                                break;
                            }
                        }
                        y = make_float4(__uint_as_float(lo.x), __uint_as_float(lo.z), __uint_as_float(hi.x), __uint_as_float(hi.z));
                    }
                    sum.x += y.x; sum.y += y.y; sum.z += y.z; sum.w += y.w;
                }
                __stcg(reinterpret_cast<float4*>(e.g_total) + j, sum);
            }
            __threadfence();
            __syncthreads();
            if (rc == 0) {
                atomicAdd(e.sync_words + 1, 1u);
                const long long t0 = clock64();
                uint32_t v;
                do {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(e.sync_words + 1) : "memory");
                    if ((int32_t)(v - arrivals) >= 0) break;
                    if (clock64() - t0 > 4000000000LL) {
                        if (ad.status) *ad.status = 1;
                        break;
                    }
                } while (true);
            }
            __syncthreads();
            g_src = e.g_total;
        }
        TRACE(erow, 48);
        float4 G[OPT];  // all loads first (one L2 round trip)
        {
            const float gs = ad.grad_scale;
#pragma unroll
            for (int u = 0; u < OPT; ++u) {
                const int i = 4 * (rc + u * EPT);
                if (i + 3 < n) {
                    G[u] = __ldcg(reinterpret_cast<const float4*>(g_src) + rc + u * EPT);
                } else {  // (buffer 0 is the net's own gradient array: exactly n floats)
                    auto get = [&](int k) { return i + k < n ? __ldcg(g_src + i + k) : 0.f; };
                    G[u] = make_float4(get(0), get(1), get(2), get(3));
                }
            }
            float sq = 0.f;
#pragma unroll
            for (int u = 0; u < OPT; ++u) {
                sq = fmaf(G[u].x * gs, G[u].x * gs, sq); sq = fmaf(G[u].y * gs, G[u].y * gs, sq);
                sq = fmaf(G[u].z * gs, G[u].z * gs, sq); sq = fmaf(G[u].w * gs, G[u].w * gs, sq);
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) sq += __shfl_down_sync(0xffffffffu, sq, d);
            if ((rc & 31) == 0) s_red[rc >> 5] = sq;
            __syncthreads();
            if (rc == 0) {
                float t = 0.f;
                for (int w = 0; w < EPT / 32; ++w) t += s_red[w];
                const float norm = sqrtf(t);
                float c = 1.f;
                if (ad.max_norm > 0.f) c = fminf(ad.max_norm / (norm + 1e-6f), 1.f);
                s_coef[0] = c * gs;
                if (ad.gnorm_out && blockIdx.x == 0) *ad.gnorm_out = norm;
            }
            __syncthreads();
        }
        TRACE(erow, 44);
        {
            const float c0 = s_coef[0], c1 = s_coef[1], c2 = s_coef[2];
            const bool last = m == e.n_mb - 1 && blockIdx.x == 0;
            constexpr int NMV = HV == 1 ? 32 : 16;
            float Mo[NMV], Vo[NMV];  // one TMEM round trip for all moments of this thread (HV = 1: entries 28..31 unused)
            tc::tmem_st_wait();  // the previous step's stores of these columns (waited here, not there)
            if constexpr (HV == 1) {
                float v0[16], v1[16];
                tc::tmem_ld32(s.lane_base + TM_OPT_M, Mo);
                tc::tmem_ld16(s.lane_base + TM_OPT_V0, v0);
                tc::tmem_ld16(s.lane_base + TM_OPT_V1, v1);
#pragma unroll
                for (int k = 0; k < 16; ++k) { Vo[k] = v0[k]; Vo[16 + k] = v1[k]; }
            } else {
                tc::tmem_ld16(s.lane_base + ep_m_col<HV>(0, half), Mo);
                tc::tmem_ld16(s.lane_base + ep_v_col<HV>(0, half), Vo);
            }
#pragma unroll
            for (int u = 0; u < OPT; ++u) {
                const int i = 4 * (rc + u * EPT);
                const bool live = i < n;  // TMEM traffic stays warp-uniform: dead entries are all-zero and stay so
                float4 mm = make_float4(Mo[4 * u], Mo[4 * u + 1], Mo[4 * u + 2], Mo[4 * u + 3]);
                float4 vv = make_float4(Vo[4 * u], Vo[4 * u + 1], Vo[4 * u + 2], Vo[4 * u + 3]);
                Pr[u].x = opt::adam_elem_fast(Pr[u].x, G[u].x, mm.x, vv.x, c0, c1, c2, ad.beta1, ad.beta2, ad.eps);
                Pr[u].y = opt::adam_elem_fast(Pr[u].y, G[u].y, mm.y, vv.y, c0, c1, c2, ad.beta1, ad.beta2, ad.eps);
                Pr[u].z = opt::adam_elem_fast(Pr[u].z, G[u].z, mm.z, vv.z, c0, c1, c2, ad.beta1, ad.beta2, ad.eps);
                Pr[u].w = opt::adam_elem_fast(Pr[u].w, G[u].w, mm.w, vv.w, c0, c1, c2, ad.beta1, ad.beta2, ad.eps);
                Mo[4 * u] = mm.x; Mo[4 * u + 1] = mm.y; Mo[4 * u + 2] = mm.z; Mo[4 * u + 3] = mm.w;
                Vo[4 * u] = vv.x; Vo[4 * u + 1] = vv.y; Vo[4 * u + 2] = vv.z; Vo[4 * u + 3] = vv.w;
                if ((q_contig >> u) & 1u) {  // four consecutive inputs of one output unit: 8 contiguous bytes of the image
                    const uint32_t q = Q[u].x & 0xffffu;
                    const uint2 h = make_uint2(pack_h2(Pr[u].x, Pr[u].y), pack_h2(Pr[u].z, Pr[u].w));
#pragma unroll
                    for (int g = 0; g < TC_GROUPS; ++g) *reinterpret_cast<uint2*>(smem_all + g * SM_GROUP + SM_IMG + 2 * q) = h;
                    if (last) *reinterpret_cast<uint2*>(ad.image + q) = h;
                } else if (live) {  // W1 rows (in_dim is not a multiple of 8) and the biases: one position at a time
                    const float pp[4] = {Pr[u].x, Pr[u].y, Pr[u].z, Pr[u].w};
                    const uint32_t qq[4] = {Q[u].x & 0xffffu, Q[u].x >> 16, Q[u].y & 0xffffu, Q[u].y >> 16};
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (qq[k] != 0xffffu) {
                            const __half h = __float2half_rn(pp[k]);
#pragma unroll
                            for (int g = 0; g < TC_GROUPS; ++g) *reinterpret_cast<__half*>(smem_all + g * SM_GROUP + SM_IMG + 2 * qq[k]) = h;
                            if (last) ad.image[qq[k]] = h;
                        }
                }
                if (last) {  // the net's own arrays, once per launch
                    const float pp[4] = {Pr[u].x, Pr[u].y, Pr[u].z, Pr[u].w}, m4[4] = {mm.x, mm.y, mm.z, mm.w}, v4[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (i + k < n) { ad.p[i + k] = pp[k]; ad.m[i + k] = m4[k]; ad.v[i + k] = v4[k]; }
                }
            }
#pragma unroll
            for (int c = 0; c < NMV / 8; ++c) {
                float t8[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) t8[k] = Mo[8 * c + k];
                tc::tmem_st8(s.lane_base + ep_m_col<HV>(0, half) + 8 * c, t8);
#pragma unroll
                for (int k = 0; k < 8; ++k) t8[k] = Vo[8 * c + k];
                tc::tmem_st8(s.lane_base + (HV == 1 ? (c < 2 ? TM_OPT_V0 + 8 * c : TM_OPT_V1 + 8 * (c - 2)) : ep_v_col<HV>(0, half) + 8 * c), t8);
            }
        }
        TRACE(erow, 45);
        {   // buffer (m - 1) % 3 has been read by everybody (all CTAs are past barrier m); it is next reduced into after barrier m + 1
            float* g_old = ((m + 2) % 3 == 0) ? a.grads : ((m + 2) % 3 == 1 ? e.grads1 : e.grads2);
            const int64_t per = (n + gridDim.x - 1) / gridDim.x;
            const int64_t lo = (int64_t)blockIdx.x * per, hi = lo + per < n ? lo + per : n;
            for (int64_t i = lo + rc; i < hi; i += EPT) g_old[i] = 0.f;
        }
        TRACE(erow, 46);
        tc::fence_async_smem();  // the image entries written above are read by both groups' MMAs
        __syncthreads();
        TRACE(erow, 47);
    }
    {   // the last buffer: clear it once nobody reads it any more; publish the step count
        __threadfence();
        grid.sync();
        float* g_last = ((e.n_mb - 1) % 3 == 0) ? a.grads : ((e.n_mb - 1) % 3 == 1 ? e.grads1 : e.grads2);
        const int64_t per = (n + gridDim.x - 1) / gridDim.x;
        const int64_t lo = (int64_t)blockIdx.x * per, hi = lo + per < n ? lo + per : n;
        for (int64_t i = lo + rc; i < hi; i += EPT) g_last[i] = 0.f;
        if (blockIdx.x == 0 && rc == 0) *ad.step = step0 + e.n_mb;
        if (blockIdx.x == 0 && rc < 2) e.sync_words[rc] = 0u;  // arrival counters: zero between launches
    }
    tc::tmem_st_wait();
    tc::fence_before_sync();
    __syncthreads();
    if ((rc >> 5) == 0) tc::tmem_dealloc(*reinterpret_cast<const uint32_t*>(smem_all + SM_BAR + 16), TM_COLS_BWD * TC_GROUPS);
}
