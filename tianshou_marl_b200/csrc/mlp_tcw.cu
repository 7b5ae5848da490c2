// mlp_tcw.cu -- engine 2: the "wide" tcgen05 engine for the CTDE / MAPPO nets of BASELINE.json configs[2]
// (reference multiagent/ctde.py:346-414: DecentralizedActor obs -> 128 -> 128 -> n_act, CentralizedCritic
// global state (A * obs = 600 wide at N = 10) -> 128 -> 128 -> one value per agent), sm_100a.
//
// Net shape in(<= 1023) -> 128 -> 128 -> out(<= 16).  Every GEMM is tcgen05.mma kind::f16 (fp16 operands in shared
// memory, fp32 accumulators in TMEM); descriptors and the chunked operand layout [feature/8][row 0..127][feature%8] are
// those of tc.cuh (pinned by tests/test_gpu_tc.py).  Differences to engine 1 (mlp_tc_kernels.cuh):
//   * inputs arrive as PACKED fp16 tile images (tsm_tcw_pack_x): one 128-row tile = KP/8 chunks of 2 KB, so a tile (or a
//     64-feature slice of it) is ONE contiguous TMA bulk copy, and the same bytes serve as K-major A operand of the
//     first layer and as MN-major B operand of its weight gradient;
//   * KP <= 64 ("resident": the actor): W1 and a double-buffered X tile live in shared memory, the W1 gradient
//     accumulates in TMEM across all tiles of the CTA;
//     KP  > 64 ("streaming": the centralized critic, KP = 608): X and W1 stream through a 2-stage ring in 64-feature
//     slices (16 KB each) for the forward GEMM, X streams a second time for the W1 gradient, which is produced in
//     64-column TMEM blocks (double buffered) and flushed per tile with 16-byte global reductions;
//   * hidden activations carry a constant ones feature (index 128, chunk pair 16/17), so b2 / b3 are row 128 of the
//     W2 / W3 images and db2 is column 128 of the W2 gradient -- no bias epilogue anywhere; b1 rides on input feature KP-1;
//   * warp roles: 8 epilogue warps (thread = row x column half), one TMA producer warp, one MMA issuer warp; epilogue
//     -> issuer hand-over through a named barrier (bar.arrive / bar.sync), issuer -> epilogue through tcgen05.commit
//     on an mbarrier.
// Minibatches are sets of whole tiles (tile list per launch); the optimizer (tcw_adam_kernel) steps up to two nets with
// one joint gradient norm (PPO clips over actor + critic parameters together, a2c.py:102-105).
#include <algorithm>
#include <cstdlib>
#include <cstdio>
#include "common.cuh"
#include "philox.cuh"
#include <cuda_fp16.h>
#include "feistel.cuh"
#include "ppo_math.cuh"
#include "adam.cuh"
#include "tc.cuh"

namespace tsm {
namespace tcw {

using namespace tsm::ppo;

constexpr int ROWS = 128;          // rows per tile
constexpr int H = 128;             // hidden width
constexpr int HP = 144;            // hidden width + the chunk pair that holds the ones feature (index 128)
constexpr int OP = 16;             // padded head outputs
constexpr int CH = ROWS * 16;      // bytes of one 8-feature chunk of a 128-row image (2048)
constexpr int EPI_THREADS = 256;   // 8 epilogue warps
constexpr int THREADS = 320;       // + producer warp (8) + MMA issuer warp (9)
constexpr int W2_BYTES = (HP / 8) * H * 16;   // [18][128][8] halves = 36 864
constexpr int W3_BYTES = (HP / 8) * OP * 16;  // [18][16][8] halves  =  4 608
constexpr int RES_KP = 64;         // resident path: KP <= 64
constexpr int SLICE = 64;          // streaming path: features per ring slot
constexpr int SLICE_B = (SLICE / 8) * CH;  // 16 384 bytes of X (or W1) per ring slot
constexpr int STAGES = 2;

// shared-memory map (bytes)
constexpr int SM_W2 = 0;
constexpr int SM_W3 = SM_W2 + W2_BYTES;
constexpr int SM_H1 = SM_W3 + W3_BYTES;        // 18 chunks (16, 17: the constant ones feature)
constexpr int SM_H2 = SM_H1 + 18 * CH;         // 18 chunks
constexpr int SM_D1 = SM_H2;                   // D1 overwrites H2's first 16 chunks once the W3 gradient has been read
constexpr int SM_D2 = SM_H2 + 18 * CH;         // 16 chunks
constexpr int SM_G = SM_D2 + 16 * CH;          // 2 chunks
constexpr int SM_BAR = SM_G + 2 * CH;          // mbarriers (16 x 8 bytes), TMEM base at +128, reduction scratch at +256
constexpr int SM_DYN = SM_BAR + 1024;          // kernel-specific area
// resident: W1 [KP/8][128][8] (<= 16 KB) then two X tiles (<= 16 KB each)
constexpr int SM_RW1 = SM_DYN;
constexpr int SM_RX = SM_RW1 + (RES_KP / 8) * CH;
constexpr int SM_RES_TOTAL = SM_RX + 2 * (RES_KP / 8) * CH;
// streaming: STAGES x { X slice 16 KB, W1 slice 16 KB }
constexpr int SM_RING = SM_DYN;
constexpr int SM_STR_TOTAL = SM_RING + STAGES * 2 * SLICE_B;
static_assert(SM_RES_TOTAL <= 227 * 1024 && SM_STR_TOTAL <= 227 * 1024, "shared-memory map exceeds an SM");

// mbarrier slots
enum { B_MAIN = 0, B_XFULL = 1 /*2*/, B_XEMPTY = 3 /*2*/, B_FULL = 5 /*STAGES*/, B_EMPTY = 7 /*STAGES*/, B_DONE = 9 /*2*/, B_FREE = 11 /*2*/, B_W = 13 };

// TMEM columns (fp32, 128 lanes)
constexpr uint32_t TM_SCR = 0;     // 128: H1pre / H2pre / DH2 / DH1 in turn
constexpr uint32_t TM_OUT = 128;   // 16
constexpr uint32_t TM_GW3 = 144;   // 16: lanes = H2 feature, columns = head output
constexpr uint32_t TM_GW2 = 160;   // 144: lanes = D2 feature (out), columns = [H1 | 1] feature (in; 128 = bias)
constexpr uint32_t TM_GW1 = 304;   // resident: KP columns; streaming: 2 blocks of 64
constexpr uint32_t TM_COLS = 512;

struct Img {  // byte offsets inside the weight image of a net with padded input width KP
    int w1, w2, w3, total;
};
__host__ __device__ inline int pad_in(int in_dim) { return (in_dim + 1 + 15) & ~15; }
__host__ __device__ inline Img img_of(int KP) {
    Img i;
    i.w1 = 0;
    i.w2 = KP * H * 2;
    i.w3 = i.w2 + W2_BYTES;
    i.total = i.w3 + W3_BYTES;
    return i;
}
// halves index of flat parameter i inside the image
__host__ __device__ inline int64_t image_index(int64_t i, const Offsets& o, int in_dim, int KP) {
    const Img im = img_of(KP);
    if (i < o.b1) {
        const int out = (int)(i / in_dim), in = (int)(i - (int64_t)out * in_dim);
        return im.w1 / 2 + ((int64_t)(in >> 3) * H + out) * 8 + (in & 7);
    }
    if (i < o.w2) return im.w1 / 2 + ((int64_t)((KP - 1) >> 3) * H + (i - o.b1)) * 8 + ((KP - 1) & 7);
    if (i < o.b2) {
        const int rel = (int)(i - o.w2), out = rel >> 7, in = rel & 127;
        return im.w2 / 2 + ((int64_t)(in >> 3) * H + out) * 8 + (in & 7);
    }
    if (i < o.w3) return im.w2 / 2 + ((int64_t)(H >> 3) * H + (i - o.b2)) * 8;
    if (i < o.b3) {
        const int rel = (int)(i - o.w3), out = rel >> 7, in = rel & 127;
        return im.w3 / 2 + ((int64_t)(in >> 3) * OP + out) * 8 + (in & 7);
    }
    return im.w3 / 2 + ((int64_t)(H >> 3) * OP + (i - o.b3)) * 8;
}

// the same with 32-bit arithmetic (device hot path of the optimizer kernel; n_params < 2^31)
__device__ __forceinline__ int image_index32(int i, const Offsets& o, int in_dim, int KP) {
    const int b1 = (int)o.b1, w2 = (int)o.w2, b2 = (int)o.b2, w3 = (int)o.w3, b3 = (int)o.b3;
    const int iw2 = KP * H, iw3 = iw2 + W2_BYTES / 2;  // halves offsets of the W2 / W3 images
    if (i < b1) {
        const int out = i / in_dim, in = i - out * in_dim;
        return ((in >> 3) * H + out) * 8 + (in & 7);
    }
    if (i < w2) return (((KP - 1) >> 3) * H + (i - b1)) * 8 + ((KP - 1) & 7);
    if (i < b2) {
        const int rel = i - w2, out = rel >> 7, in = rel & 127;
        return iw2 + ((in >> 3) * H + out) * 8 + (in & 7);
    }
    if (i < w3) return iw2 + ((H >> 3) * H + (i - b2)) * 8;
    if (i < b3) {
        const int rel = i - w3, out = rel >> 7, in = rel & 127;
        return iw3 + ((in >> 3) * OP + out) * 8 + (in & 7);
    }
    return iw3 + ((H >> 3) * OP + (i - b3)) * 8;
}

// ---------------------------------------------------------------------------------------------
// small PTX helpers on top of tc.cuh
__device__ __forceinline__ void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* dst, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ uint32_t h2_bits(__half2 h) { return *reinterpret_cast<const uint32_t*>(&h); }
__device__ __forceinline__ uint32_t pack_h2(float a, float b) { return h2_bits(__floats2half2_rn(a, b)); }
__device__ __forceinline__ uint32_t pack_relu_h2(float a, float b) {
    return h2_bits(__hmax2(__floats2half2_rn(a, b), __float2half2_rn(0.f)));
}
constexpr int BAR_READY = 1;                   // named barrier: 256 epilogue arrivals + the issuer warp
constexpr int READY_COUNT = EPI_THREADS + 32;

__device__ __forceinline__ uint64_t desc_at(uint32_t sb4, uint32_t off_bytes, uint32_t lbo, uint32_t sbo) {
    const uint32_t lo = sb4 + ((off_bytes >> 4) + ((lbo >> 4) << 16));
    const uint32_t hi = (sbo >> 4) | (1u << 14);
    return ((uint64_t)hi << 32) | lo;
}
// activation image (128 rows): K-major (K = features) / MN-major (K = rows)
__device__ __forceinline__ uint64_t d_act_k(uint32_t sb4, int off, int ks) { return desc_at(sb4, off + ks * 2 * CH, CH, 128); }
__device__ __forceinline__ uint64_t d_act_mn(uint32_t sb4, int off, int ks) { return desc_at(sb4, off + ks * 256, 128, CH); }
// weight image [in/8][n_out][8]: K-major (N = out, K = in) / MN-major (N = in, K = out)
__device__ __forceinline__ uint64_t d_w_k(uint32_t sb4, int off, int n_out, int ks) { return desc_at(sb4, off + ks * 2 * n_out * 16, n_out * 16, 128); }
__device__ __forceinline__ uint64_t d_w_mn(uint32_t sb4, int off, int n_out, int ks) { return desc_at(sb4, off + ks * 256, 128, n_out * 16); }

// ---------------------------------------------------------------------------------------------
struct Args {
    const unsigned char* image;   // weight image of the net
    float* grads;                 // flat gradient buffer (train only)
    int in_dim, KP, out_dim, n_act;
    // tiles of this launch: logical tile p in [0, n_tiles) -> physical tile of the packed input image
    const unsigned char* ximg;
    int64_t n_rows;               // rows behind ximg (the last tile may be ragged)
    const int32_t* super_ids;     // optional: physical tile = super_ids[(tile0 + p) / sub] * sub + (tile0 + p) % sub
    int sub;
    int64_t tile0, n_tiles;
    // forward outputs (mode >= 0) -- indexed by the PHYSICAL row
    int mode;                     // -1: train; 0 raw outputs; 1 sample; 2 argmax; 3 evaluate given actions
    uint64_t seed, counter;
    int64_t rng_row0;
    float* out;                   // [n_rows][out_dim] (mode 0) or logits [n_rows][n_act]
    int32_t* act;
    float* logp;
    float* entropy;
    // train
    int loss_kind;                // 0: policy loss on n_act logits (row scalars adv / logp_old / act)
                                  // 1: per-head value loss on `heads` outputs (targets [row][heads])
                                  // 2: value = mean of the heads against one target per row (ctde.py:153-166)
    const float* adv;             // loss 0: [n_rows]
    const float* logp_old;        // loss 0: [n_rows]
    const float* target;          // loss 1: [n_rows][heads]; loss 2: [n_rows]
    const float* v_old;           // loss 1, value_clip: [n_rows][heads]
    const float* weight;          // loss 1, optional: [n_rows][heads] (0 / 1)
    int heads;
    PpoCfg cfg;
    float inv_n, gscale, inv_gscale;
    const float* mb_stats;        // optional (mean, rstd) of the minibatch advantages
    double* loss_sums;            // [4]: clip, vf, ent, rows
};

// Instrumented builds only (tools/tcw_trace.py, -DTSM_TCW_TRACE): clock64 stamps of CTA 0 -- slots 0..63 by the first epilogue
// thread, 64..127 by the MMA issuer lane, 128..191 by the producer lane.
#ifdef TSM_TCW_TRACE
__device__ long long g_tcw_trace[192];
#define TCW_TRACE(slot) do { if (blockIdx.x == 0 && (slot) < 192) g_tcw_trace[slot] = clock64(); } while (0)
#else
#define TCW_TRACE(slot) do { } while (0)
#endif

struct Ctx {
    unsigned char* smem;
    uint64_t* bar;
    uint32_t tmem, lane_base, sb4;
    int tid, row, half, warp;
    uint32_t ph_main;
};

__device__ __forceinline__ int64_t phys_tile(const Args& a, int64_t p) {
    const int64_t q = a.tile0 + p;
    if (!a.super_ids) return q;
    return (int64_t)a.super_ids[q / a.sub] * a.sub + q % a.sub;
}

__device__ __forceinline__ void wait_main(Ctx& c) {
    tc::mbar_wait(c.bar + B_MAIN, c.ph_main);
    c.ph_main ^= 1;
    tc::fence_after_sync();
}
// epilogue threads: smem writes + TMEM reads of this stage are done -> hand over to the issuer warp
__device__ __forceinline__ void ready(const Ctx&) {
    tc::fence_async_smem();
    tc::fence_before_sync();
    bar_arrive(BAR_READY, READY_COUNT);
}

// relu(SCR) -> fp16 image at dst (this thread: its row, 64 of the 128 columns)
__device__ __forceinline__ void epi_relu(const Ctx& c, int dst) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(c.lane_base + TM_SCR + 64 * c.half + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            uint4 w;
            w.x = pack_relu_h2(v[8 * h + 0], v[8 * h + 1]);
            w.y = pack_relu_h2(v[8 * h + 2], v[8 * h + 3]);
            w.z = pack_relu_h2(v[8 * h + 4], v[8 * h + 5]);
            w.w = pack_relu_h2(v[8 * h + 6], v[8 * h + 7]);
            *reinterpret_cast<uint4*>(c.smem + dst + (8 * c.half + 4 * q + h) * CH + c.row * 16) = w;
        }
    }
}
// SCR masked by (activation image `img` > 0) -> fp16 image at dst
__device__ __forceinline__ void epi_mask(const Ctx& c, int img, int dst) {
    const __half2 zero = __float2half2_rn(0.f);
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(c.lane_base + TM_SCR + 64 * c.half + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int cell = (8 * c.half + 4 * q + h) * CH + c.row * 16;
            const uint4 m = *reinterpret_cast<const uint4*>(c.smem + img + cell);
            const __half2* mh = reinterpret_cast<const __half2*>(&m);
            uint4 w;
            w.x = pack_h2(v[8 * h + 0], v[8 * h + 1]) & __hgt2_mask(mh[0], zero);
            w.y = pack_h2(v[8 * h + 2], v[8 * h + 3]) & __hgt2_mask(mh[1], zero);
            w.z = pack_h2(v[8 * h + 4], v[8 * h + 5]) & __hgt2_mask(mh[2], zero);
            w.w = pack_h2(v[8 * h + 6], v[8 * h + 7]) & __hgt2_mask(mh[3], zero);
            *reinterpret_cast<uint4*>(c.smem + dst + cell) = w;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// set-up shared by all kernels: constant ones chunks, G zero, weight images W2 / W3 (+ W1 when resident), barriers, TMEM
template <bool STREAM>
__device__ __forceinline__ Ctx setup(unsigned char* smem, const Args& a) {
    Ctx c;
    c.smem = smem;
    c.bar = reinterpret_cast<uint64_t*>(smem + SM_BAR);
    c.tid = threadIdx.x;
    c.warp = threadIdx.x >> 5;
    c.row = threadIdx.x & 127;
    c.half = (threadIdx.x >> 7) & 1;
    uint32_t* tmem_s = reinterpret_cast<uint32_t*>(smem + SM_BAR + 128);
    if (c.warp == 0) tc::tmem_alloc(tmem_s, TM_COLS);
    if (c.tid == 32) {
        tc::mbar_init(c.bar + B_MAIN, 1);
        tc::mbar_init(c.bar + B_W, 1);
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(c.bar + B_XFULL + i, 1);
            tc::mbar_init(c.bar + B_XEMPTY + i, 1);
            tc::mbar_init(c.bar + B_FULL + i, 1);
            tc::mbar_init(c.bar + B_EMPTY + i, 1);
            tc::mbar_init(c.bar + B_DONE + i, 1);
            tc::mbar_init(c.bar + B_FREE + i, EPI_THREADS);
        }
        tc::fence_barrier_init();
        tc::fence_async_smem();
        // weight images by TMA bulk copies (W2 | W3 are adjacent in the image and in shared memory; W1 when resident): the
        // issuer warp waits for B_W before its first MMA, nobody else reads them.  (Per-thread 16-byte copy loops cost 4.4 k
        // cycles of dependent L2 round trips here.)
        const Img im = img_of(a.KP);
        const uint32_t w1b = STREAM ? 0u : (uint32_t)(a.KP * H * 2);
        tc::mbar_expect_tx(c.bar + B_W, (uint32_t)(W2_BYTES + W3_BYTES) + w1b);
        tc::bulk_g2s(smem + SM_W2, a.image + im.w2, (uint32_t)(W2_BYTES + W3_BYTES), c.bar + B_W);
        if (!STREAM) tc::bulk_g2s(smem + SM_RW1, a.image + im.w1, w1b, c.bar + B_W);
    }
    // ones feature (index 128 = element 0 of chunk 16) of H1 and H2, chunk 17 zero; G zero
    for (int i = c.tid; i < 2 * ROWS; i += THREADS) {
        const uint4 v = i < ROWS ? make_uint4(0x00003C00u, 0, 0, 0) : make_uint4(0, 0, 0, 0);  // fp16 1.0 in element 0
        *reinterpret_cast<uint4*>(smem + SM_H1 + 16 * CH + i * 16) = v;
        *reinterpret_cast<uint4*>(smem + SM_H2 + 16 * CH + i * 16) = v;
        *reinterpret_cast<uint4*>(smem + SM_G + i * 16) = make_uint4(0, 0, 0, 0);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    c.tmem = *tmem_s;
    c.lane_base = c.tmem + ((uint32_t)((c.warp & 3) * 32) << 16);
    c.sb4 = tc::smem_u32(smem) >> 4;
    c.ph_main = 0;
    return c;
}

// ---- issuer-side pieces (one elected lane of warp 9) ---------------------------------------------------------------------
__device__ __forceinline__ void issue_l2(const Ctx& c) {
    constexpr uint32_t ID = tc::make_idesc_f16(128, H, 0, 0, 0, 0);
#pragma unroll
    for (int ks = 0; ks < HP / 16; ++ks) tc::mma_f16(c.tmem + TM_SCR, d_act_k(c.sb4, SM_H1, ks), d_w_k(c.sb4, SM_W2, H, ks), ID, ks > 0);
}
__device__ __forceinline__ void issue_l3(const Ctx& c) {
    constexpr uint32_t ID = tc::make_idesc_f16(128, OP, 0, 0, 0, 0);
#pragma unroll
    for (int ks = 0; ks < HP / 16; ++ks) tc::mma_f16(c.tmem + TM_OUT, d_act_k(c.sb4, SM_H2, ks), d_w_k(c.sb4, SM_W3, OP, ks), ID, ks > 0);
}
__device__ __forceinline__ void issue_dh2(const Ctx& c) {  // DH2 = G W3 : A = G (K = 16 outputs), B = W3 MN-major (N = 128 inputs)
    constexpr uint32_t ID = tc::make_idesc_f16(128, H, 0, 0, 0, 1);
    tc::mma_f16(c.tmem + TM_SCR, d_act_k(c.sb4, SM_G, 0), d_w_mn(c.sb4, SM_W3, OP, 0), ID, 0);
}
__device__ __forceinline__ void issue_gw3(const Ctx& c, bool first) {  // [H2]^T G : lanes = H2 feature, columns = head output
    constexpr uint32_t ID = tc::make_idesc_f16(128, OP, 0, 0, 1, 1);
#pragma unroll
    for (int ks = 0; ks < ROWS / 16; ++ks)
        tc::mma_f16(c.tmem + TM_GW3, d_act_mn(c.sb4, SM_H2, ks), d_act_mn(c.sb4, SM_G, ks), ID, (ks > 0 || !first) ? 1u : 0u);
}
__device__ __forceinline__ void issue_dh1(const Ctx& c) {  // DH1 = D2 W2 : B = W2 MN-major (N = 128 inputs, K = 128 outputs)
    constexpr uint32_t ID = tc::make_idesc_f16(128, H, 0, 0, 0, 1);
#pragma unroll
    for (int ks = 0; ks < H / 16; ++ks) tc::mma_f16(c.tmem + TM_SCR, d_act_k(c.sb4, SM_D2, ks), d_w_mn(c.sb4, SM_W2, H, ks), ID, ks > 0);
}
__device__ __forceinline__ void issue_gw2(const Ctx& c, bool first) {  // D2^T [H1 | 1] : N = 144 (column 128 = db2)
    constexpr uint32_t ID = tc::make_idesc_f16(128, HP, 0, 0, 1, 1);
#pragma unroll
    for (int ks = 0; ks < ROWS / 16; ++ks)
        tc::mma_f16(c.tmem + TM_GW2, d_act_mn(c.sb4, SM_D2, ks), d_act_mn(c.sb4, SM_H1, ks), ID, (ks > 0 || !first) ? 1u : 0u);
}

// ---- loss stage (epilogue threads of column half 0: thread = row) ---------------------------------------------------------
struct LossAcc {
    float clip, vf, ent, rows;
    float db3[OP];
};

__device__ __forceinline__ void loss_stage(const Ctx& c, const Args& a, int64_t prow, bool valid, LossAcc& acc, const float* side /*regs*/,
                                           int act_reg) {
    float o[16];
    tc::tmem_ld16(c.lane_base + TM_OUT, o);
    float dz[OP];
#pragma unroll
    for (int j = 0; j < OP; ++j) dz[j] = 0.f;
    if (valid) {
        if (a.loss_kind == 0) {
            float z[OUTP], d8[OUTP];
#pragma unroll
            for (int j = 0; j < OUTP; ++j) z[j] = o[j];
            float mean = 0.f, rstd = 1.f;
            if (a.mb_stats) { mean = a.mb_stats[0]; rstd = a.mb_stats[1]; }
            const RowLoss rl = row_loss<true>(z, a.n_act, -1, act_reg, side[0], 0.f, side[1], 0.f, mean, rstd, a.cfg, a.inv_n, d8);
#pragma unroll
            for (int j = 0; j < OUTP; ++j) dz[j] = d8[j];
            acc.clip += rl.clip;
            acc.ent += rl.ent;
            acc.rows += 1.f;
        } else if (a.loss_kind == 1) {
            const float* tg = a.target + prow * a.heads;
            const float* vo = a.v_old ? a.v_old + prow * a.heads : nullptr;
            const float* wt = a.weight ? a.weight + prow * a.heads : nullptr;
#pragma unroll
            for (int j = 0; j < OP; ++j) {
                if (j < a.heads) {
                    const float w = wt ? wt[j] : 1.f;
                    const float ret = tg[j], v = o[j];
                    float vf = (ret - v) * (ret - v), dv = 2.f * (v - ret);
                    if (a.cfg.value_clip) {
                        const float vold = vo[j];
                        const float d = v - vold;
                        const float vc = vold + fminf(fmaxf(d, -a.cfg.eps_clip), a.cfg.eps_clip);
                        const float vf2 = (ret - vc) * (ret - vc);
                        if (vf2 > vf) {
                            vf = vf2;
                            dv = (d >= -a.cfg.eps_clip && d <= a.cfg.eps_clip) ? 2.f * (vc - ret) : 0.f;
                        }
                    }
                    acc.vf += w * vf;
                    acc.rows += w;
                    dz[j] = w * a.cfg.vf_coef * a.inv_n * dv;
                }
            }
        } else {
            float v = 0.f;
#pragma unroll
            for (int j = 0; j < OP; ++j)
                if (j < a.heads) v += o[j];
            v /= (float)a.heads;
            const float ret = a.target[prow];
            acc.vf += (ret - v) * (ret - v);
            acc.rows += 1.f;
            const float dvs = a.cfg.vf_coef * a.inv_n * 2.f * (v - ret) / (float)a.heads;
#pragma unroll
            for (int j = 0; j < OP; ++j)
                if (j < a.heads) dz[j] = dvs;
        }
    }
#pragma unroll
    for (int j = 0; j < OP; ++j) acc.db3[j] += dz[j];
    uint4 g0, g1;
    const float S = a.gscale;
    g0.x = pack_h2(dz[0] * S, dz[1] * S); g0.y = pack_h2(dz[2] * S, dz[3] * S); g0.z = pack_h2(dz[4] * S, dz[5] * S); g0.w = pack_h2(dz[6] * S, dz[7] * S);
    g1.x = pack_h2(dz[8] * S, dz[9] * S); g1.y = pack_h2(dz[10] * S, dz[11] * S); g1.z = pack_h2(dz[12] * S, dz[13] * S); g1.w = pack_h2(dz[14] * S, dz[15] * S);
    *reinterpret_cast<uint4*>(c.smem + SM_G + c.row * 16) = g0;
    *reinterpret_cast<uint4*>(c.smem + SM_G + CH + c.row * 16) = g1;
}

// forward output stage (threads of column half 0): same semantics as tc_forward_kernel (mlp_tc_kernels.cuh)
__device__ __forceinline__ void output_stage(const Ctx& c, const Args& a, int64_t prow, bool valid) {
    float o[16];
    tc::tmem_ld16(c.lane_base + TM_OUT, o);
    if (!valid) return;
    if (a.mode == 0) {
        if (a.out) {
#pragma unroll
            for (int j = 0; j < OP; ++j)
                if (j < a.out_dim) a.out[prow * a.out_dim + j] = o[j];
        }
        return;
    }
    float z[OUTP];
#pragma unroll
    for (int j = 0; j < OUTP; ++j) z[j] = o[j];
    const Soft sm = softmax_row<true>(z, a.n_act, a.cfg.probs_param != 0);
    int act = 0;
    if (a.mode == 1) {
        const int64_t key = prow + a.rng_row0;
        const Philox4 rnd = philox4x32((uint32_t)key, (uint32_t)((uint64_t)key >> 32), (uint32_t)a.counter, (uint32_t)(a.counter >> 32),
                                       (uint32_t)a.seed, (uint32_t)(a.seed >> 32) ^ STREAM_ACT);
        const float u = u01(rnd.x);
        float cum = 0.f;
        act = a.n_act - 1;
        bool found = false;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j < a.n_act) {
                cum += sm.p[j];
                if (!found && u < cum) { act = j; found = true; }
            }
    } else if (a.mode == 2) {
        float best = -INFINITY;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j < a.n_act && z[j] > best) { best = z[j]; act = j; }
    } else {
        act = a.act[prow];
    }
    float lp = 0.f;
#pragma unroll
    for (int j = 0; j < OUTP; ++j)
        if (j == act) lp = sm.logp[j];
    if (a.mode != 3 && a.act) a.act[prow] = act;
    if (a.logp) a.logp[prow] = lp;
    if (a.entropy) a.entropy[prow] = sm.ent;
    if (a.out) {
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j < a.n_act) a.out[prow * a.n_act + j] = z[j];
    }
}

// ---------------------------------------------------------------------------------------------
// flush of the TMEM-resident gradient accumulators (after the last tile of the CTA, all MMAs complete: shared memory is free).
// The accumulators are laid down in shared memory at their FLAT gradient positions, then added to the global gradient with
// fully coalesced 16-byte reductions (a warp instruction = 512 contiguous bytes = four L2 lines).  Measured on B200 before
// this staging (per-thread row-strided reductions straight from TMEM registers): 43 k of the actor kernel's 79 k cycles.
// The W2 block (lane = row, 32 float4 per row) is stored with its float4 column XORed by (row & 7): conflict-free on both sides.
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 2, 256;" ::: "memory"); }

template <bool STREAM>
__device__ __forceinline__ void flush_grads(const Ctx& c, const Args& a, const LossAcc& acc) {
    const Offsets o = offsets(a.in_dim, H, a.out_dim);
    const float s = a.inv_gscale;
    const int64_t base = STREAM ? o.w2 : 0;  // the streaming path has flushed the W1 / b1 part tile by tile
    float* stg = reinterpret_cast<float*>(c.smem);
    float4* stg4 = reinterpret_cast<float4*>(c.smem);
    const int w2q = (int)((o.w2 - base) >> 2);  // float4 index of the W2 block
    // W2 gradient: lane = out, columns = in (0..127), column 128 = b2
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(c.lane_base + TM_GW2 + 64 * c.half + 32 * q, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c4 = 16 * c.half + 8 * q + j;
            stg4[w2q + c.row * 32 + (c4 ^ (c.row & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
    }
    if (c.half == 1) {
        float v[8];
        tc::tmem_ld8(c.lane_base + TM_GW2 + H, v);
        stg[o.b2 - base + c.row] = v[0];
        if (c.row < a.out_dim) stg[o.b3 - base + c.row] = 0.f;  // b3 goes straight to global memory below
    } else {
        // W3 gradient: lane = in (H2 feature), column = output j
        float v[16];
        tc::tmem_ld16(c.lane_base + TM_GW3, v);
#pragma unroll
        for (int j = 0; j < OP; ++j)
            if (j < a.out_dim) stg[o.w3 - base + j * H + c.row] = v[j];
        // b3 gradient: fp32 register sums of the loss stage (unscaled), reduced over the warp
#pragma unroll
        for (int j = 0; j < OP; ++j) {
            float t = acc.db3[j];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
            if ((c.tid & 31) == 0 && j < a.out_dim) atomicAdd(a.grads + o.b3 + j, t);
        }
    }
    if (!STREAM) {
        // W1 gradient: lane = out, columns = input feature (KP - 1 = b1)
        const int half_cols = a.KP / 2;  // 8, 16, 24 or 32
        for (int k0 = 0; k0 < half_cols; k0 += 8) {
            float v[8];
            const int col = half_cols * c.half + k0;
            tc::tmem_ld8(c.lane_base + TM_GW1 + col, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int k = col + j;
                if (k < a.in_dim) stg[c.row * a.in_dim + k] = v[j];
                else if (k == a.KP - 1) stg[o.b1 + c.row] = v[j];
            }
        }
    }
    epi_sync();
    const int total = (int)(o.total - base), n4 = total >> 2;
    float* g = a.grads + base;
    // every CTA adds to the same addresses: start each CTA at its own offset so that they do not queue on the same L2 lines
    const int rot = (int)((blockIdx.x * 1031u) % (unsigned)max(n4, 1));
    for (int F0 = c.tid; F0 < n4; F0 += EPI_THREADS) {
        int F = F0 + rot;
        if (F >= n4) F -= n4;
        int idx = F;
        const int rel = F - w2q;
        if (rel >= 0 && rel < H * H / 4) {
            const int r = rel >> 5, c4 = rel & 31;
            idx = w2q + r * 32 + (c4 ^ (r & 7));
        }
        const float4 v = stg4[idx];
        red_add_v4(g + 4 * F, v.x * s, v.y * s, v.z * s, v.w * s);
    }
    for (int i = 4 * n4 + c.tid; i < total; i += EPI_THREADS) atomicAdd(g + i, stg[i] * s);
}

__device__ __forceinline__ void flush_losses(const Ctx& c, const Args& a, const LossAcc& acc) {
    if (!a.loss_sums || c.half != 0) return;
    float v[4] = {acc.clip, acc.vf, acc.ent, acc.rows};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], d);
        if ((c.tid & 31) == 0 && v[k] != 0.f) atomicAdd(a.loss_sums + k, (double)v[k]);
    }
}

// ---------------------------------------------------------------------------------------------
// the kernel: persistent over the launch's tiles (tile = blockIdx.x, += gridDim.x).  TRAIN: forward + loss + backward;
// otherwise forward + output stage.
template <bool STREAM, bool TRAIN>
__global__ void __launch_bounds__(THREADS, 1) tcw_kernel(Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    if (threadIdx.x == 0) TCW_TRACE(0);
    Ctx c = setup<STREAM>(smem, a);
    if (threadIdx.x == 0) TCW_TRACE(1);
    const int KP = a.KP;
    const int n_slices = (KP + SLICE - 1) / SLICE;
    const int64_t tile_bytes = (int64_t)(KP / 8) * CH;
    const int64_t my_tiles = a.n_tiles > (int64_t)blockIdx.x ? (a.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const Img im = img_of(KP);
    // streaming path, W1-gradient pass: every CTA walks the 64-column slices from its own starting slice, so that the CTAs
    // of a minibatch do not all reduce into the same 32 KB of the gradient at the same time
    const int rot = (int)(blockIdx.x % (unsigned)n_slices);

    if (c.warp == 8) {
        // ================= TMA producer =================
        if (tc::elect_one()) {
            if (!STREAM) {
                for (int64_t i = 0; i < my_tiles; ++i) {
                    const int b = (int)(i & 1);
                    if (i >= 2) tc::mbar_wait(c.bar + B_XEMPTY + b, (uint32_t)(((i >> 1) - 1) & 1));
                    const int64_t pt = phys_tile(a, blockIdx.x + i * gridDim.x);
                    tc::mbar_expect_tx(c.bar + B_XFULL + b, (uint32_t)tile_bytes);
                    tc::bulk_g2s(smem + SM_RX + b * (RES_KP / 8) * CH, a.ximg + pt * tile_bytes, (uint32_t)tile_bytes, c.bar + B_XFULL + b);
                }
            } else {
                uint32_t it = 0;  // ring slot uses so far
                for (int64_t i = 0; i < my_tiles; ++i) {
                    const int64_t pt = phys_tile(a, blockIdx.x + i * gridDim.x);
                    const unsigned char* xt = a.ximg + pt * tile_bytes;
                    for (int pass = 0; pass < (TRAIN ? 2 : 1); ++pass) {
                        for (int sl0 = 0; sl0 < n_slices; ++sl0, ++it) {
                            const int sl = pass == 1 ? (sl0 + rot) % n_slices : sl0;  // W1-gradient pass: rotated slice order
                            const int s = (int)(it % STAGES);
                            if (it >= STAGES) tc::mbar_wait(c.bar + B_EMPTY + s, (uint32_t)((it / STAGES - 1) & 1));
                            const int nk = min(SLICE, KP - sl * SLICE);
                            const uint32_t bytes = (uint32_t)(nk / 8) * CH;
                            unsigned char* dst = smem + SM_RING + s * 2 * SLICE_B;
                            if (pass == 0) {
                                tc::mbar_expect_tx(c.bar + B_FULL + s, 2 * bytes);
                                tc::bulk_g2s(dst, xt + (int64_t)sl * SLICE_B, bytes, c.bar + B_FULL + s);
                                tc::bulk_g2s(dst + SLICE_B, a.image + im.w1 + (int64_t)sl * SLICE_B, bytes, c.bar + B_FULL + s);
                            } else {
                                tc::mbar_expect_tx(c.bar + B_FULL + s, bytes);
                                tc::bulk_g2s(dst, xt + (int64_t)sl * SLICE_B, bytes, c.bar + B_FULL + s);
                            }
                        }
                    }
                }
            }
        }
    } else if (c.warp == 9) {
        // ================= MMA issuer =================
        constexpr uint32_t ID_H = tc::make_idesc_f16(128, H, 0, 0, 0, 0);
        uint32_t it = 0, blk = 0;
        tc::mbar_wait(c.bar + B_W, 0);  // weight images have landed
        for (int64_t i = 0; i < my_tiles; ++i) {
            const int ib = 64 + 12 * (int)i;
            (void)ib;
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib);
            // ---- layer 1 ----
            if (!STREAM) {
                const int b = (int)(i & 1);
                tc::mbar_wait(c.bar + B_XFULL + b, (uint32_t)((i >> 1) & 1));
                tc::fence_after_sync();
                if (tc::elect_one()) {
                    const int xoff = SM_RX + b * (RES_KP / 8) * CH;
                    for (int ks = 0; ks < KP / 16; ++ks)
                        tc::mma_f16(c.tmem + TM_SCR, d_act_k(c.sb4, xoff, ks), d_w_k(c.sb4, SM_RW1, H, ks), ID_H, ks > 0);
                    tc::commit(c.bar + B_MAIN);
                }
                __syncwarp();
            } else {
                for (int sl = 0; sl < n_slices; ++sl, ++it) {
                    const int s = (int)(it % STAGES);
                    tc::mbar_wait(c.bar + B_FULL + s, (uint32_t)((it / STAGES) & 1));
                    tc::fence_after_sync();
                    if (tc::elect_one()) {
                        const int off = SM_RING + s * 2 * SLICE_B;
                        const int nks = min(SLICE, KP - sl * SLICE) / 16;
                        for (int ks = 0; ks < nks; ++ks)
                            tc::mma_f16(c.tmem + TM_SCR, d_act_k(c.sb4, off, ks), d_w_k(c.sb4, off + SLICE_B, H, ks), ID_H, (sl > 0 || ks > 0) ? 1u : 0u);
                        tc::commit(c.bar + B_EMPTY + s);
                        if (sl == n_slices - 1) tc::commit(c.bar + B_MAIN);
                    }
                    __syncwarp();
                }
            }
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 1);  // layer 1 issued
            // ---- layer 2 ----
            bar_sync(BAR_READY, READY_COUNT);
            tc::fence_after_sync();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 2);  // H1 ready seen
            if (tc::elect_one()) { issue_l2(c); tc::commit(c.bar + B_MAIN); }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 3);  // layer 2 issued
            // ---- layer 3 ----
            bar_sync(BAR_READY, READY_COUNT);
            tc::fence_after_sync();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 4);
            if (tc::elect_one()) { issue_l3(c); tc::commit(c.bar + B_MAIN); }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 5);
            if (!TRAIN) {
                bar_sync(BAR_READY, READY_COUNT);  // OUT has been read: the next tile may overwrite SCR / OUT
                tc::fence_after_sync();
                if (!STREAM && tc::elect_one()) tc::commit(c.bar + B_XEMPTY + (int)(i & 1));
                __syncwarp();
                continue;
            }
            // ---- G ready: DH2 ----
            bar_sync(BAR_READY, READY_COUNT);
            tc::fence_after_sync();
            if (tc::elect_one()) { issue_dh2(c); tc::commit(c.bar + B_MAIN); }
            __syncwarp();
            // ---- D2 ready: W3 gradient, DH1 (waited), W2 gradient ----
            bar_sync(BAR_READY, READY_COUNT);
            tc::fence_after_sync();
            if (tc::elect_one()) {
                issue_gw3(c, i == 0);
                issue_dh1(c);
                tc::commit(c.bar + B_MAIN);
                issue_gw2(c, i == 0);
            }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 6);  // W3 gradient, DH1, W2 gradient issued
            // ---- D1 ready: W1 gradient ----
            bar_sync(BAR_READY, READY_COUNT);
            tc::fence_after_sync();
            if ((threadIdx.x & 31) == 0) TCW_TRACE(ib + 7);
            if (!STREAM) {
                if (tc::elect_one()) {
                    const int b = (int)(i & 1);
                    const int xoff = SM_RX + b * (RES_KP / 8) * CH;
                    const uint32_t ID = tc::make_idesc_f16(128, KP, 0, 0, 1, 1);
                    for (int ks = 0; ks < ROWS / 16; ++ks)
                        tc::mma_f16(c.tmem + TM_GW1, d_act_mn(c.sb4, SM_D1, ks), d_act_mn(c.sb4, xoff, ks), ID, (ks > 0 || i > 0) ? 1u : 0u);
                    tc::commit(c.bar + B_XEMPTY + b);
                }
                __syncwarp();
            } else {
                for (int sl0 = 0; sl0 < n_slices; ++sl0, ++it, ++blk) {
                    const int sl = (sl0 + rot) % n_slices;
                    const int s = (int)(it % STAGES);
                    const int bk = (int)(blk & 1);
                    tc::mbar_wait(c.bar + B_FULL + s, (uint32_t)((it / STAGES) & 1));
                    if ((threadIdx.x & 31) == 0 && i == 0) TCW_TRACE(160 + 3 * sl0);
                    if (blk >= 2) tc::mbar_wait(c.bar + B_FREE + bk, (uint32_t)(((blk >> 1) - 1) & 1));
                    tc::fence_after_sync();
                    if ((threadIdx.x & 31) == 0 && i == 0) TCW_TRACE(161 + 3 * sl0);
                    if (tc::elect_one()) {
                        const int off = SM_RING + s * 2 * SLICE_B;
                        const int nk = min(SLICE, KP - sl * SLICE);
                        const uint32_t ID = tc::make_idesc_f16(128, nk, 0, 0, 1, 1);
                        for (int ks = 0; ks < ROWS / 16; ++ks)
                            tc::mma_f16(c.tmem + TM_GW1 + 64 * bk, d_act_mn(c.sb4, SM_D1, ks), d_act_mn(c.sb4, off, ks), ID, ks > 0);
                        tc::commit(c.bar + B_EMPTY + s);
                        tc::commit(c.bar + B_DONE + bk);
                    }
                    __syncwarp();
                    if ((threadIdx.x & 31) == 0 && i == 0) TCW_TRACE(162 + 3 * sl0);
                }
            }
        }
        if ((threadIdx.x & 31) == 0) TCW_TRACE(127);
        if (TRAIN) {  // everything issued: one last commit the epilogue threads wait for before the flush
            if (tc::elect_one()) tc::commit(c.bar + B_MAIN);
            __syncwarp();
        }
    } else {
        // ================= epilogue warps (0..7) =================
        LossAcc acc;
        acc.clip = acc.vf = acc.ent = acc.rows = 0.f;
#pragma unroll
        for (int j = 0; j < OP; ++j) acc.db3[j] = 0.f;
        uint32_t blk = 0;
        const Offsets o = offsets(a.in_dim, H, a.out_dim);
        for (int64_t i = 0; i < my_tiles; ++i) {
            const int64_t pt = phys_tile(a, blockIdx.x + i * gridDim.x);
            const int64_t prow = pt * ROWS + c.row;
            const bool valid = prow < a.n_rows;
            float side[2] = {0.f, 0.f};
            int act_reg = 0;
            if (TRAIN && c.half == 0 && valid && a.loss_kind == 1) {  // pull the row's targets towards L1 while the forward runs
                const float* tg = a.target + prow * a.heads;
                asm volatile("prefetch.global.L1 [%0];" ::"l"(tg));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(tg + a.heads - 1));
                if (a.v_old) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.v_old + prow * a.heads));
            }
            if (TRAIN && c.half == 0 && valid && a.loss_kind == 0) {  // the loss stage's row scalars: in flight during the forward
                side[0] = __ldg(a.adv + prow);
                side[1] = a.logp_old ? __ldg(a.logp_old + prow) : 0.f;
                act_reg = __ldg(a.act + prow);
            }
            const int tb = 2 + 12 * (int)i;  // trace slots of this tile (first 5 tiles)
            (void)tb;
            if (c.tid == 0) TCW_TRACE(tb);
            wait_main(c);  // layer 1 done
            if (c.tid == 0) TCW_TRACE(tb + 1);
            epi_relu(c, SM_H1);
            ready(c);
            if (c.tid == 0) TCW_TRACE(tb + 2);
            wait_main(c);  // layer 2 done
            if (c.tid == 0) TCW_TRACE(tb + 3);
            epi_relu(c, SM_H2);
            ready(c);
            if (c.tid == 0) TCW_TRACE(tb + 4);
            wait_main(c);  // layer 3 done
            if (c.tid == 0) TCW_TRACE(tb + 5);
            if (!TRAIN) {
                if (c.half == 0) output_stage(c, a, prow, valid);
                ready(c);
                continue;
            }
            if (c.half == 0) loss_stage(c, a, prow, valid, acc, side, act_reg);
            ready(c);
            if (c.tid == 0) TCW_TRACE(tb + 6);
            wait_main(c);  // DH2 done
            if (c.tid == 0) TCW_TRACE(tb + 7);
            epi_mask(c, SM_H2, SM_D2);
            ready(c);
            if (c.tid == 0) TCW_TRACE(tb + 8);
            wait_main(c);  // W3 gradient + DH1 done: H2 may be overwritten by D1
            if (c.tid == 0) TCW_TRACE(tb + 9);
            epi_mask(c, SM_H1, SM_D1);
            ready(c);
            if (c.tid == 0) TCW_TRACE(tb + 10);
            if (STREAM) {
                // W1 gradient of this tile, 64 input features per TMEM block: flush straight to the gradient buffer
                for (int sl0 = 0; sl0 < n_slices; ++sl0, ++blk) {
                    const int sl = (sl0 + rot) % n_slices;
                    const int bk = (int)(blk & 1);
                    tc::mbar_wait(c.bar + B_DONE + bk, (uint32_t)((blk >> 1) & 1));
                    tc::fence_after_sync();
                    if (c.tid == 0 && i == 0) TCW_TRACE(128 + 3 * sl0);
                    const int nk = min(SLICE, KP - sl * SLICE);
                    float v[32];
                    const bool mine = 32 * c.half < nk;
                    if (mine) tc::tmem_ld32(c.lane_base + TM_GW1 + 64 * bk + 32 * c.half, v);
                    tc::fence_before_sync();
                    mbar_arrive(c.bar + B_FREE + bk);
                    // stage the block (row = out, 16 float4 slots per row, column XORed by row & 7), then coalesced reductions.
                    // Two staging buffers (H1's first 16 chunks / D2: both free once the first block is done, because every MMA
                    // that read them was issued before it): the barrier of block b + 1 also fences the readers of block b.
                    float4* st4 = reinterpret_cast<float4*>(smem + (bk ? SM_D2 : SM_H1));
                    if (mine) {
                        const float s = a.inv_gscale;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            st4[c.row * 16 + ((8 * c.half + j) ^ (c.row & 7))] = make_float4(v[4 * j] * s, v[4 * j + 1] * s, v[4 * j + 2] * s, v[4 * j + 3] * s);
                    }
                    epi_sync();
                    if (c.tid == 0 && i == 0) TCW_TRACE(129 + 3 * sl0);
                    const int q4 = nk >> 2;
                    const bool vec = (a.in_dim & 3) == 0;
                    const int rrot = (int)(blockIdx.x * 37u) & (ROWS - 1);  // CTAs walk the rows from different starting points
                    for (int F = c.tid; F < ROWS * q4; F += EPI_THREADS) {
                        const int r0 = q4 == 16 ? F >> 4 : (q4 == 8 ? F >> 3 : F / q4);
                        const int c4 = F - r0 * q4, r = (r0 + rrot) & (ROWS - 1);
                        const float4 val = st4[r * 16 + (c4 ^ (r & 7))];
                        const int k0 = sl * SLICE + 4 * c4;
                        float* dst = a.grads + o.w1 + (int64_t)r * a.in_dim + k0;
                        if (vec && k0 + 4 <= a.in_dim) {
#ifdef TSM_TCW_EXPERIMENT_STORE
                            *reinterpret_cast<float4*>(dst + (int64_t)blockIdx.x * 0) = val;  // timing experiment only: plain stores
#else
                            red_add_v4(dst, val.x, val.y, val.z, val.w);
#endif
                        } else {
                            const float e[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                if (k0 + u < a.in_dim) atomicAdd(dst + u, e[u]);
                                else if (k0 + u == KP - 1) atomicAdd(a.grads + o.b1 + r, e[u]);
                            }
                        }
                    }
                    if (c.tid == 0 && i == 0) TCW_TRACE(130 + 3 * sl0);
                }
                epi_sync();  // the staging buffers alias H1 / D2: nobody writes the next tile's activations while a block is being read
            }
        }
        if (TRAIN) {
            if (c.tid == 0) TCW_TRACE(62);
            if (my_tiles > 0) {
                wait_main(c);  // all MMAs of this CTA have completed
                if (c.tid == 0) TCW_TRACE(63);
                flush_grads<STREAM>(c, a, acc);
            }
            flush_losses(c, a, acc);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (threadIdx.x == 0) TCW_TRACE(191);
    if (c.warp == 0) tc::tmem_dealloc(c.tmem, TM_COLS);
}

// ---------------------------------------------------------------------------------------------
// f32 rows -> packed fp16 tile images [tile][KP/8][128][8]; feature KP-1 = 1 for valid rows, everything else of the padding 0.
// One CTA = one tile x 64 features, transposed through shared memory so that reads (row segments) and writes (2 KB chunks)
// are both coalesced.
__global__ void __launch_bounds__(256) pack_x_kernel(const float* __restrict__ x, int64_t ld, int in_dim, int KP, const int64_t* __restrict__ idx,
                                                     int64_t n_rows, unsigned char* __restrict__ ximg) {
    __shared__ float s[ROWS][SLICE + 1];
    const int64_t tile = blockIdx.x;
    const int f0 = blockIdx.y * SLICE;
    const int nf = min(SLICE, KP - f0);
    for (int e = threadIdx.x; e < ROWS * SLICE; e += 256) {
        const int r = e / SLICE, k = e % SLICE;
        const int64_t row = tile * ROWS + r;
        float v = 0.f;
        if (row < n_rows && k < nf) {
            const int f = f0 + k;
            if (f < in_dim) v = x[(idx ? idx[row] : row) * ld + f];
            else if (f == KP - 1) v = 1.f;
        }
        s[r][k] = v;
    }
    __syncthreads();
    unsigned char* dst = ximg + tile * (int64_t)(KP / 8) * CH + (int64_t)(f0 / 8) * CH;
    for (int e = threadIdx.x; e < (nf / 8) * ROWS; e += 256) {
        const int ch = e / ROWS, r = e % ROWS;
        const float* p = &s[r][8 * ch];
        uint4 v;
        v.x = pack_h2(p[0], p[1]); v.y = pack_h2(p[2], p[3]); v.z = pack_h2(p[4], p[5]); v.w = pack_h2(p[6], p[7]);
        *reinterpret_cast<uint4*>(dst + ch * CH + r * 16) = v;
    }
}

// fp32 parameters -> weight image (standalone; the optimizer kernel keeps the image current afterwards)
__global__ void pack_image_kernel(const float* __restrict__ p, int64_t n, int in_dim, int out_dim, int KP, __half* __restrict__ image) {
    const Offsets o = offsets(in_dim, H, out_dim);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        image[image_index(i, o, in_dim, KP)] = __float2half_rn(p[i]);
}

// ---------------------------------------------------------------------------------------------
// optimizer: joint gradient-norm clip over up to two nets (torch.nn.utils.clip_grad_norm_ over ActorCritic parameters,
// algorithm_base.py:485-498, a2c.py:102-105) + torch.optim.Adam per element + fp16 image entry + zeroed gradient.
struct AdamNet {
    float *p, *g, *m, *v;
    int64_t n;
    __half* image;
    int in_dim, out_dim, KP;
    int64_t* step_out;
};
struct AdamWs {      // TSM_TCW_ADAM_WS_BYTES, zeroed once by the caller
    unsigned int count;      // arrivals of all launches so far (the grid barrier counts monotonically)
    unsigned int pad[3];
    double partial[254];     // one squared-norm partial per CTA: plain stores, summed in CTA order by everybody after the barrier
};
struct AdamJob {
    AdamNet net[2];
    int n_nets;
    // data-parallel ranks of one node (world == 1: unused).  Every rank's gradient arena ([2 parities][stride] floats) and
    // flag pad live in symmetric memory: rank r announces "my gradient of exchange xseq is complete" with one flag store per
    // peer, waits for all ranks' flags, and then READS the gradients of all ranks over NVLink (coalesced peer loads, summed
    // in rank order: the same bits everywhere).  The arena of the other parity is zeroed here for the minibatch after next.
    int world, rank;
    const float* peer_g[8];   // rank q's arena of this exchange's parity
    uint32_t* peer_flag[8];   // rank q's flag pad
    int64_t goff[2];          // offset of each net's gradient inside a parity block
    float* gsum;              // local scratch [stride]: the rank-summed gradient (phase 1 -> phase 2)
    float* zero_next;         // local arena of the other parity
    uint32_t xseq;
    int* status;
    int64_t step;            // 1-based step number of THIS update
    float lr, beta1, beta2, eps, max_norm, grad_scale;
    AdamWs* ws;
    uint32_t seq;            // number of earlier launches on this workspace (the grid barrier counts monotonically)
    float* gnorm_out;
};

__global__ void __launch_bounds__(256) tcw_adam_kernel(AdamJob j) {
    __shared__ double s_part[8];
    __shared__ float s_coef;
    float coef = j.grad_scale;
    const bool multi = j.world > 1;
    if (multi) {
        // (a) announce; (b) wait for every rank (bounded: a missing peer sets *status instead of hanging the GPU)
        if (blockIdx.x == 0 && (int)threadIdx.x < j.world)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(j.peer_flag[threadIdx.x] + j.rank), "r"(j.xseq) : "memory");
        if ((int)threadIdx.x < j.world) {
            const uint32_t* f = j.peer_flag[j.rank] + threadIdx.x;
            const long long t0 = clock64();
            uint32_t v;
            do {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                if ((int)(v - j.xseq) >= 0) break;
                if (clock64() - t0 > 6000000000LL) { *j.status = 1; break; }  // ~3 s
            } while (true);
        }
        __syncthreads();
        // (c) rank-ordered sum of the peers' gradients -> gsum (and the other parity's arena zeroed for the minibatch after next)
        for (int t = 0; t < j.n_nets; ++t) {
            const AdamNet& n = j.net[t];
            const int cnt = (int)n.n, stride = (int)(gridDim.x * blockDim.x);
            const int64_t off = j.goff[t];
            for (int i0 = (int)(blockIdx.x * blockDim.x + threadIdx.x); i0 < cnt; i0 += 4 * stride) {
                // all ranks' loads are issued before the first add (in-order issue: an add behind each rank's loads would cost
                // one NVLink round trip per rank instead of one in total)
                float v[8][4];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        v[r][u] = 0.f;
                        if (r < j.world && i0 + u * stride < cnt)
                            asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v[r][u]) : "l"(j.peer_g[r] + off + i0 + u * stride));
                    }
                }
                float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int r = 0; r < 8; ++r) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) g[u] += v[r][u];  // rank order; absent ranks add +0
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (i0 + u * stride < cnt) {
                        j.gsum[off + i0 + u * stride] = g[u];
                        j.zero_next[off + i0 + u * stride] = 0.f;
                    }
            }
        }
    }
    if (j.max_norm > 0.f) {
        double acc = 0.0;
        for (int t = 0; t < j.n_nets; ++t) {
            const AdamNet& n = j.net[t];
            const int cnt = (int)n.n, stride = (int)(gridDim.x * blockDim.x);
            const float* gp = multi ? j.gsum + j.goff[t] : n.g;  // (each thread re-reads only what it wrote itself)
            for (int i0 = (int)(blockIdx.x * blockDim.x + threadIdx.x); i0 < cnt; i0 += 4 * stride) {
                float g[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) g[u] = i0 + u * stride < cnt ? __ldcg(gp + i0 + u * stride) * j.grad_scale : 0.f;
#pragma unroll
                for (int u = 0; u < 4; ++u) acc += (double)g[u] * (double)g[u];
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < 8; ++w) t += s_part[w];
            // (148 atomicAdd(double) on one address serialise in L2 for ~3 us; a slot per CTA costs one store)
            *reinterpret_cast<volatile double*>(&j.ws->partial[blockIdx.x]) = t;
            __threadfence();
            atomicAdd(&j.ws->count, 1u);
            const unsigned int target = (j.seq + 1u) * gridDim.x;
            while ((int)(*reinterpret_cast<volatile unsigned int*>(&j.ws->count) - target) < 0) {
            }
            __threadfence();
        }
        __syncthreads();
        if (threadIdx.x < 32) {  // the same sum in the same order in every CTA: one clip coefficient, bit for bit
            double t = 0.0;
            for (unsigned int b = threadIdx.x; b < gridDim.x; b += 32) t += __ldcg(&j.ws->partial[b]);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
            if (threadIdx.x == 0) {
                const float norm = (float)sqrt(t);
                const float cc = j.max_norm / (norm + 1e-6f);  // clip_grad_norm_: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1
                s_coef = j.grad_scale * (cc < 1.f ? cc : 1.f);
                if (blockIdx.x == 0 && j.gnorm_out) *j.gnorm_out = norm;
            }
        }
        __syncthreads();
        coef = s_coef;
    }
    const double bc1 = 1.0 - opt::ipow((double)j.beta1, j.step), bc2 = 1.0 - opt::ipow((double)j.beta2, j.step);
    const float step_size = (float)((double)j.lr / bc1), bc2_sqrt = (float)sqrt(bc2);
    const int stride = (int)(gridDim.x * blockDim.x);
    for (int t = 0; t < j.n_nets; ++t) {
        const AdamNet& n = j.net[t];
        const Offsets o = offsets(n.in_dim, H, n.out_dim);
        const int cnt = (int)n.n;
        const float* gp = multi ? j.gsum + j.goff[t] : n.g;
        // four independent elements per thread and pass: the loads of all four are in flight together
        for (int i0 = (int)(blockIdx.x * blockDim.x + threadIdx.x); i0 < cnt; i0 += 4 * stride) {
            float p[4], g[4], m[4], v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * stride;
                if (i < cnt) { p[u] = n.p[i]; g[u] = __ldcg(gp + i); m[u] = n.m[i]; v[u] = n.v[i]; }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * stride;
                if (i < cnt) {
                    const float pn = opt::adam_elem(p[u], g[u], m[u], v[u], coef, step_size, bc2_sqrt, j.beta1, j.beta2, j.eps);
                    n.p[i] = pn; n.m[i] = m[u]; n.v[i] = v[u];
                    if (!multi) n.g[i] = 0.f;  // (multi: peers may still be reading this parity; the other one was zeroed above)
                    if (n.image) n.image[image_index32(i, o, n.in_dim, n.KP)] = __float2half_rn(pn);
                }
            }
        }
        if (blockIdx.x == 0 && threadIdx.x == 0 && n.step_out) *n.step_out = j.step;
    }
}

// per-minibatch advantage statistics for tile-ordered minibatches: minibatch m = super-tiles [m * mb_super, (m+1) * mb_super)
// of the (optional) list, a super-tile = rows_per_super consecutive rows.  stats[m] = (mean, 1 / (std_unbiased + eps)),
// ppo.py:184-186.
__global__ void __launch_bounds__(512) adv_stats_kernel(const float* __restrict__ adv, int64_t n_rows, const int32_t* __restrict__ super_ids,
                                                        int64_t rows_per_super, int64_t n_super, int64_t mb_super, float eps,
                                                        float* __restrict__ stats) {
    __shared__ double s1[16], s2[16], sn[16];
    const int64_t m = blockIdx.x;
    const int64_t s_lo = m * mb_super, s_hi = min(n_super, s_lo + mb_super);
    double a1 = 0.0, a2 = 0.0, an = 0.0;
    for (int64_t sidx = s_lo; sidx < s_hi; ++sidx) {
        const int64_t r0 = (super_ids ? (int64_t)super_ids[sidx] : sidx) * rows_per_super;
        const int64_t r1 = min(n_rows, r0 + rows_per_super);
        for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
            const double v = (double)adv[r];
            a1 += v; a2 += v * v; an += 1.0;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        a1 += __shfl_xor_sync(0xffffffffu, a1, d);
        a2 += __shfl_xor_sync(0xffffffffu, a2, d);
        an += __shfl_xor_sync(0xffffffffu, an, d);
    }
    if ((threadIdx.x & 31) == 0) { s1[threadIdx.x >> 5] = a1; s2[threadIdx.x >> 5] = a2; sn[threadIdx.x >> 5] = an; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t1 = 0.0, t2 = 0.0, tn = 0.0;
        for (int w = 0; w < 16; ++w) { t1 += s1[w]; t2 += s2[w]; tn += sn[w]; }
        const double mean = tn > 0 ? t1 / tn : 0.0;
        const double var = tn > 1 ? fmax(0.0, (t2 - tn * mean * mean) / (tn - 1.0)) : 0.0;
        stats[2 * m] = (float)mean;
        stats[2 * m + 1] = 1.f / ((float)sqrt(var) + eps);
    }
}

}  // namespace tcw
}  // namespace tsm

using namespace tsm;

// ---------------------------------------------------------------------------------------------
static int tcw_check_net(const tsm_net* n, const char* fn) {
    if (!n) return fail(TSM_ERR_INVALID, "%s: net is NULL", fn);
    if (n->engine != 2) return fail(TSM_ERR_INVALID, "%s: net.engine must be 2 (wide tcgen05 engine)", fn);
    if (n->hidden != tcw::H || n->in_dim < 1 || n->in_dim > 1023 || n->out_dim < 1 || n->out_dim > tcw::OP)
        return fail(TSM_ERR_UNSUPPORTED, "%s: engine 2 needs hidden == 128, in_dim <= 1023, out_dim <= 16", fn);
    if (n->n_act < 0 || n->n_act > n->out_dim || n->n_act > ppo::OUTP) return fail(TSM_ERR_INVALID, "%s: 0 <= n_act <= min(out_dim, 8)", fn);
    if (!n->params || !n->tc_image) return fail(TSM_ERR_INVALID, "%s: params / tc_image is NULL", fn);
    if (n->n_params != ppo::offsets(n->in_dim, n->hidden, n->out_dim).total) return fail(TSM_ERR_INVALID, "%s: n_params does not match the layout", fn);
    if (reinterpret_cast<uintptr_t>(n->tc_image) & 15) return fail(TSM_ERR_INVALID, "%s: tc_image must be 16-byte aligned", fn);
    return TSM_OK;
}

extern "C" int64_t tsm_tcw_image_bytes(int in_dim) { return in_dim >= 1 ? (int64_t)tcw::img_of(tcw::pad_in(in_dim)).total : 0; }

static int64_t tcw_x_bytes_impl(int in_dim, int64_t n_rows) {
    if (in_dim < 1 || n_rows < 0) return 0;
    return ceil_div(n_rows, tcw::ROWS) * (int64_t)(tcw::pad_in(in_dim) / 8) * tcw::CH;
}
extern "C" int64_t tsm_tcw_x_bytes(int in_dim, int64_t n_rows) { return tcw_x_bytes_impl(in_dim, n_rows); }

namespace tsm {
int tcw_net_pack(const tsm_net* net, cudaStream_t s) {
    int rc = tcw_check_net(net, __func__);
    if (rc) return rc;
    const int KP = tcw::pad_in(net->in_dim);
    TSM_CUDA(cudaMemsetAsync(net->tc_image, 0, (size_t)tcw::img_of(KP).total, s));
    tcw::pack_image_kernel<<<(unsigned)std::min<int64_t>(ceil_div(net->n_params, 256), 296), 256, 0, s>>>(
        net->params, net->n_params, net->in_dim, net->out_dim, KP, reinterpret_cast<__half*>(net->tc_image));
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
}  // namespace tsm

extern "C" int tsm_tcw_pack_x(const float* x, int64_t ld, int in_dim, const int64_t* idx, int64_t n_rows, void* ximg, tsm_stream_t stream) {
    TSM_REQUIRE(x && ximg && in_dim >= 1 && in_dim <= 1023 && ld >= in_dim && n_rows >= 0, "x / ximg != NULL, 1 <= in_dim <= 1023, ld >= in_dim");
    TSM_REQUIRE((reinterpret_cast<uintptr_t>(ximg) & 15) == 0, "ximg must be 16-byte aligned");
    if (n_rows == 0) return TSM_OK;
    const int KP = tcw::pad_in(in_dim);
    const int64_t tiles = ceil_div(n_rows, tcw::ROWS);
    TSM_REQUIRE(tiles < ((int64_t)1 << 31), "n_rows too large");
    dim3 grid((unsigned)tiles, (unsigned)((KP + tcw::SLICE - 1) / tcw::SLICE));
    tcw::pack_x_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, ld, in_dim, KP, idx, n_rows, reinterpret_cast<unsigned char*>(ximg));
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

template <bool STREAM, bool TRAIN>
static int tcw_launch(const tcw::Args& a, cudaStream_t s, int max_ctas = 0) {
    const size_t smem = STREAM ? tcw::SM_STR_TOTAL : tcw::SM_RES_TOTAL;
    static bool attr = false;
    if (!attr) {
        TSM_CUDA(cudaFuncSetAttribute(tcw::tcw_kernel<STREAM, TRAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    const unsigned grid = (unsigned)std::min<int64_t>(a.n_tiles, (int64_t)(max_ctas > 0 ? max_ctas : sm_count()));
    tcw::tcw_kernel<STREAM, TRAIN><<<grid, tcw::THREADS, smem, s>>>(a);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

static void tcw_fill_net(tcw::Args& a, const tsm_net* net) {
    a.image = reinterpret_cast<const unsigned char*>(net->tc_image);
    a.grads = net->grads;
    a.in_dim = net->in_dim; a.KP = tcw::pad_in(net->in_dim); a.out_dim = net->out_dim; a.n_act = net->n_act;
}

extern "C" int tsm_tcw_forward(const tsm_net* net, const void* ximg, int64_t n_rows, int mode, uint64_t seed, uint64_t counter, int64_t rng_row0,
                               float* out, int32_t* act, float* logp, float* entropy, tsm_stream_t stream) {
    int rc = tcw_check_net(net, __func__);
    if (rc) return rc;
    TSM_REQUIRE(ximg && n_rows >= 0 && mode >= 0 && mode <= 3, "ximg != NULL, n_rows >= 0, 0 <= mode <= 3");
    TSM_REQUIRE(mode == 0 || net->n_act > 0, "sampling / evaluation needs n_act > 0");
    TSM_REQUIRE(mode != 3 || act != nullptr, "mode 3 (evaluate) needs act");
    if (n_rows == 0) return TSM_OK;
    tcw::Args a{};
    tcw_fill_net(a, net);
    a.ximg = reinterpret_cast<const unsigned char*>(ximg); a.n_rows = n_rows; a.super_ids = nullptr; a.sub = 1; a.tile0 = 0;
    a.n_tiles = ceil_div(n_rows, tcw::ROWS);
    a.mode = mode; a.seed = seed; a.counter = counter; a.rng_row0 = rng_row0; a.cfg.probs_param = net->probs_param;
    a.out = out; a.act = act; a.logp = logp; a.entropy = entropy;
    return a.KP <= tcw::RES_KP ? tcw_launch<false, false>(a, as_stream(stream)) : tcw_launch<true, false>(a, as_stream(stream));
}

namespace tsm {
// engines 0 / 1 signature (float32 rows in, tsm_net_forward / tsm_net_forward_range / tsm_collect_lockstep): pack into the net's
// scratch image, then the forward kernel.  Outputs are indexed like the input rows [0, n) (row lists gather; no scatter).
int tcw_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter, uint64_t seed, uint64_t counter,
                    float* out, int32_t* act, float* logp, float* value, float* entropy, int64_t rng_row0, cudaStream_t s) {
    int rc = tcw_check_net(net, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n >= 0 && mode >= 0 && mode <= 3, "n >= 0, 0 <= mode <= 3");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(x != nullptr && !scatter, "x != NULL; engine 2 has no scatter mode");
    TSM_REQUIRE(value == nullptr || (mode == 0 && net->n_act == 0 && net->out_dim == 1), "engine 2 nets are separate: no value head beside the logits");
    TSM_REQUIRE(rng_row0 % tcw::ROWS == 0, "engine 2: a row range starts on a 128-row tile boundary");
    const int64_t off = tcw_x_bytes_impl(net->in_dim, rng_row0);
    TSM_REQUIRE(net->x_scratch && net->x_scratch_bytes >= off + tcw_x_bytes_impl(net->in_dim, n), "engine 2: net.x_scratch too small for this call");
    unsigned char* ximg = reinterpret_cast<unsigned char*>(net->x_scratch) + off;
    rc = tsm_tcw_pack_x(x, net->in_dim, net->in_dim, rows, n, ximg, reinterpret_cast<tsm_stream_t>(s));
    if (rc) return rc;
    return tsm_tcw_forward(net, ximg, n, mode, seed, counter, rng_row0, mode == 0 && value ? value : out, act, logp, entropy,
                           reinterpret_cast<tsm_stream_t>(s));
}
}  // namespace tsm

static int tcw_train_impl(const tsm_net* net, const tsm_tcw_tiles* t, const tsm_tcw_loss* l, const tsm_ppo_cfg* cfg, float inv_n, float gscale,
                          const float* mb_stats, double* loss_sums, cudaStream_t stream, int max_ctas);

extern "C" int tsm_tcw_train(const tsm_net* net, const tsm_tcw_tiles* t, const tsm_tcw_loss* l, const tsm_ppo_cfg* cfg, float inv_n, float gscale,
                             const float* mb_stats, double* loss_sums, tsm_stream_t stream) {
    return tcw_train_impl(net, t, l, cfg, inv_n, gscale, mb_stats, loss_sums, as_stream(stream), 0);
}

static int tcw_train_impl(const tsm_net* net, const tsm_tcw_tiles* t, const tsm_tcw_loss* l, const tsm_ppo_cfg* cfg, float inv_n, float gscale,
                          const float* mb_stats, double* loss_sums, cudaStream_t stream, int max_ctas) {
    int rc = tcw_check_net(net, __func__);
    if (rc) return rc;
    TSM_REQUIRE(t && l && cfg && net->grads, "tiles / loss / cfg / net.grads != NULL");
    TSM_REQUIRE(t->ximg && t->n_rows > 0 && t->n_tiles >= 0 && t->tile0 >= 0 && t->sub >= 1, "tiles: ximg, n_rows > 0, n_tiles >= 0, sub >= 1");
    TSM_REQUIRE(l->kind >= 0 && l->kind <= 2, "loss.kind in 0..2");
    if (l->kind == 0) TSM_REQUIRE(net->n_act > 0 && l->adv && l->act, "policy loss needs n_act > 0, adv and act");
    if (l->kind != 0) TSM_REQUIRE(l->target && l->heads >= 1 && l->heads <= net->out_dim, "value loss needs target and 1 <= heads <= out_dim");
    if (l->kind == 1 && cfg->value_clip) TSM_REQUIRE(l->v_old != nullptr, "value_clip needs v_old");
    TSM_REQUIRE(gscale > 0.f && inv_n > 0.f, "gscale > 0, inv_n > 0");
    TSM_REQUIRE((reinterpret_cast<uintptr_t>(net->grads) & 15) == 0, "net.grads must be 16-byte aligned");
    if (t->n_tiles == 0) return TSM_OK;
    tcw::Args a{};
    tcw_fill_net(a, net);
    a.ximg = reinterpret_cast<const unsigned char*>(t->ximg); a.n_rows = t->n_rows; a.super_ids = t->super_ids; a.sub = t->sub;
    a.tile0 = t->tile0; a.n_tiles = t->n_tiles;
    a.mode = -1;
    a.loss_kind = l->kind; a.adv = l->adv; a.logp_old = l->logp_old; a.act = const_cast<int32_t*>(l->act);
    a.target = l->target; a.v_old = l->v_old; a.weight = l->weight; a.heads = l->heads;
    a.cfg.eps_clip = cfg->eps_clip; a.cfg.dual_clip = cfg->dual_clip; a.cfg.vf_coef = cfg->vf_coef; a.cfg.ent_coef = cfg->ent_coef;
    a.cfg.value_clip = cfg->value_clip; a.cfg.adv_norm = cfg->adv_norm; a.cfg.plain_pg = cfg->plain_pg; a.cfg.value_heads = 0; a.cfg.probs_param = net->probs_param;
    a.inv_n = inv_n; a.gscale = gscale; a.inv_gscale = 1.0f / gscale;
    a.mb_stats = mb_stats; a.loss_sums = loss_sums;
    return a.KP <= tcw::RES_KP ? tcw_launch<false, true>(a, stream, max_ctas) : tcw_launch<true, true>(a, stream, max_ctas);
}

static int tcw_adam_impl(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                         float grad_scale, void* workspace, uint32_t seq, float* grad_norm_out, const tsm_tcw_peers* peers, uint32_t xseq,
                         tsm_stream_t stream);

extern "C" int tsm_tcw_adam(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                            float grad_scale, void* workspace, uint32_t seq, float* grad_norm_out, tsm_stream_t stream) {
    return tcw_adam_impl(nets, n_nets, step, lr, beta1, beta2, eps, max_grad_norm, grad_scale, workspace, seq, grad_norm_out, nullptr, 0, stream);
}

extern "C" int tsm_tcw_adam_peers(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps,
                                  float max_grad_norm, void* workspace, uint32_t seq, const tsm_tcw_peers* peers, uint32_t xseq,
                                  tsm_stream_t stream) {
    TSM_REQUIRE(peers != nullptr && peers->world >= 1, "peers != NULL");
    return tcw_adam_impl(nets, n_nets, step, lr, beta1, beta2, eps, max_grad_norm, 1.0f / (float)peers->world, workspace, seq, nullptr, peers, xseq,
                         stream);
}

static int tcw_adam_impl(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                         float grad_scale, void* workspace, uint32_t seq, float* grad_norm_out, const tsm_tcw_peers* peers, uint32_t xseq,
                         tsm_stream_t stream) {
    TSM_REQUIRE(nets && n_nets >= 1 && n_nets <= 2 && step >= 1, "1 <= n_nets <= 2, step >= 1");
    TSM_REQUIRE(max_grad_norm <= 0.f || workspace != nullptr, "clipping needs the workspace (TSM_TCW_ADAM_WS_BYTES)");
    tcw::AdamJob j{};
    int64_t total = 0;
    for (int t = 0; t < n_nets; ++t) {
        int rc = tcw_check_net(&nets[t], __func__);
        if (rc) return rc;
        TSM_REQUIRE(nets[t].grads && nets[t].adam_m && nets[t].adam_v, "grads / adam state != NULL");
        tcw::AdamNet& n = j.net[t];
        n.p = nets[t].params; n.g = nets[t].grads; n.m = nets[t].adam_m; n.v = nets[t].adam_v; n.n = nets[t].n_params;
        n.image = reinterpret_cast<__half*>(nets[t].tc_image); n.in_dim = nets[t].in_dim; n.out_dim = nets[t].out_dim;
        n.KP = tcw::pad_in(nets[t].in_dim); n.step_out = nets[t].adam_step;
        total += n.n;
    }
    if (peers && peers->world > 1) {
        TSM_REQUIRE(peers->world <= 8 && peers->rank >= 0 && peers->rank < peers->world && xseq >= 1, "1 < world <= 8, 0 <= rank < world, xseq >= 1");
        TSM_REQUIRE(peers->gsum && peers->status && peers->stride > 0, "peers: gsum / status / stride");
        const int64_t par = (int64_t)(xseq & 1u);
        j.world = peers->world; j.rank = peers->rank; j.xseq = xseq; j.gsum = peers->gsum; j.status = peers->status;
        for (int q = 0; q < peers->world; ++q) {
            TSM_REQUIRE(peers->arena[q] && peers->flags[q], "peers: arena / flags of every rank");
            j.peer_g[q] = peers->arena[q] + par * peers->stride;
            j.peer_flag[q] = peers->flags[q];
        }
        j.zero_next = peers->arena[peers->rank] + (par ^ 1) * peers->stride;
        for (int t = 0; t < n_nets; ++t) {
            j.goff[t] = peers->net_off[t];
            TSM_REQUIRE(nets[t].grads == peers->arena[peers->rank] + par * peers->stride + peers->net_off[t], "net.grads must be this parity's arena slice");
        }
    } else {
        j.world = 1;
    }
    j.n_nets = n_nets; j.step = step; j.lr = lr; j.beta1 = beta1; j.beta2 = beta2; j.eps = eps; j.max_norm = max_grad_norm;
    j.grad_scale = grad_scale; j.ws = reinterpret_cast<tcw::AdamWs*>(workspace); j.seq = seq; j.gnorm_out = grad_norm_out;
    static_assert(sizeof(tcw::AdamWs) == TSM_TCW_ADAM_WS_BYTES, "workspace layout");
    TSM_REQUIRE(sm_count() <= 254, "optimizer grid exceeds the workspace's partial slots");
    // fixed grid (the grid barrier counts arrivals monotonically across launches on one workspace); all CTAs must be
    // co-resident: 148 CTAs of 256 threads always are when the kernel runs alone on its stream's turn, and the
    // cooperative launch makes the driver guarantee it
    const unsigned grid = (unsigned)sm_count();
    (void)total;
    void* args[] = {&j};
    if (max_grad_norm > 0.f) {
        TSM_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(tcw::tcw_adam_kernel), dim3(grid), dim3(256), args, 0, as_stream(stream)));
    } else {
        tcw::tcw_adam_kernel<<<grid, 256, 0, as_stream(stream)>>>(j);
        TSM_LAUNCH_CHECK();
    }
    return TSM_OK;
}

extern "C" int tsm_tcw_adv_stats(const float* adv, int64_t n_rows, const int32_t* super_ids, int64_t rows_per_super, int64_t n_super,
                                 int64_t mb_super, float eps, float* stats, tsm_stream_t stream) {
    TSM_REQUIRE(adv && stats && n_rows > 0 && rows_per_super > 0 && n_super > 0 && mb_super > 0, "adv / stats != NULL, sizes > 0");
    const int64_t n_mb = ceil_div(n_super, mb_super);
    tcw::adv_stats_kernel<<<(unsigned)n_mb, 512, 0, as_stream(stream)>>>(adv, n_rows, super_ids, rows_per_super, n_super, mb_super, eps, stats);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_tcw_ppo_epoch(const tsm_tcw_epoch* e, const tsm_ppo_cfg* cfg, tsm_stream_t stream) {
    TSM_REQUIRE(e && cfg && e->actor && e->critic && e->ximg_actor && e->ximg_critic, "epoch / cfg / nets / images != NULL");
    TSM_REQUIRE(e->n_agents >= 1 && e->n_rows > 0 && e->n_rows % e->n_agents == 0, "n_rows = worlds x n_agents");
    TSM_REQUIRE(e->n_super > 0 && e->mb_super > 0 && e->h_mb_rows && e->sums_actor && e->sums_critic, "n_super, mb_super > 0; h_mb_rows / sums != NULL");
    TSM_REQUIRE(e->act && e->adv && e->logp_old && e->ret, "act / adv / logp_old / ret != NULL");
    TSM_REQUIRE(e->critic->out_dim >= e->n_agents, "the critic needs one value head per agent");
    const int A = e->n_agents;
    const int64_t n_mb = ceil_div(e->n_super, e->mb_super);
    tsm_net nets[2] = {*e->actor, *e->critic};
    tsm_ppo_cfg ccfg = *cfg;
    uint32_t seq = e->adam_seq0;
    cudaStream_t s = as_stream(stream);
    const bool multi = e->peers != nullptr && e->peers->world > 1;
    // The critic's tiles of a minibatch (one per 128 worlds) occupy a third of the SMs for about as long as the actor's tiles
    // occupy the rest: the two launches run side by side, the critic on a forked stream, and join before the optimizer.
    static const bool two = !(getenv("TSM_TCW_STREAMS") && atoi(getenv("TSM_TCW_STREAMS")) == 1);
    static cudaStream_t side = nullptr;
    static cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    static int ev_device = -1;
    if (two) {
        int dev = 0;
        TSM_CUDA(cudaGetDevice(&dev));
        if (ev_device != dev) {  // (one device per process)
            TSM_CUDA(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
            TSM_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
            TSM_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
            ev_device = dev;
        }
    }
    for (int64_t m = 0; m < n_mb; ++m) {
        const int64_t s0 = m * e->mb_super, cnt = std::min(e->mb_super, e->n_super - s0);
        const int64_t rows = e->h_mb_rows[m];
        TSM_REQUIRE(rows > 0, "h_mb_rows > 0");
        float gs = 1.f;
        while (gs < (float)rows && gs < 1048576.f) gs *= 2.f;
        const float inv_n = 1.0f / (float)rows;
        const bool split = two && cnt * 2 <= sm_count();  // critic CTAs + actor CTAs fit side by side
        const uint32_t xseq = e->xseq0 + (uint32_t)m + 1u;
        if (multi) {  // this exchange's parity of the symmetric gradient arena
            float* blockp = e->peers->arena[e->peers->rank] + (int64_t)(xseq & 1u) * e->peers->stride;
            nets[0].grads = blockp + e->peers->net_off[0];
            nets[1].grads = blockp + e->peers->net_off[1];
        }
        tsm_tcw_tiles tc{e->ximg_critic, e->n_rows / A, e->super_ids, 1, s0, cnt};
        tsm_tcw_loss lc{};
        lc.kind = 1; lc.target = e->ret; lc.v_old = e->v_old; lc.weight = nullptr; lc.heads = A;
        if (split) {
            TSM_CUDA(cudaEventRecord(ev_fork, s));
            TSM_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        }
        int rc = tcw_train_impl(&nets[1], &tc, &lc, &ccfg, inv_n, gs, nullptr, e->sums_critic + 4 * m, split ? side : s, 0);
        if (rc) return rc;
        if (split) TSM_CUDA(cudaEventRecord(ev_join, side));
        tsm_tcw_tiles ta{e->ximg_actor, e->n_rows, e->super_ids, A, s0 * A, cnt * A};
        tsm_tcw_loss la{};
        la.kind = 0; la.act = e->act; la.adv = e->adv; la.logp_old = e->logp_old;
        rc = tcw_train_impl(&nets[0], &ta, &la, &ccfg, inv_n, gs, e->mb_stats ? e->mb_stats + 2 * m : nullptr, e->sums_actor + 4 * m, s,
                            split ? sm_count() - (int)cnt : 0);
        if (rc) return rc;
        if (split) TSM_CUDA(cudaStreamWaitEvent(s, ev_join, 0));
        if (multi)
            rc = tsm_tcw_adam_peers(nets, 2, e->step0 + m + 1, e->lr, e->beta1, e->beta2, e->eps, e->max_grad_norm, e->adam_ws, seq, e->peers, xseq, stream);
        else
            rc = tsm_tcw_adam(nets, 2, e->step0 + m + 1, e->lr, e->beta1, e->beta2, e->eps, e->max_grad_norm, 1.0f, e->adam_ws, seq, nullptr, stream);
        if (rc) return rc;
        if (e->max_grad_norm > 0.f) ++seq;
    }
    return TSM_OK;
}

#ifdef TSM_TCW_TRACE
extern "C" int tsm_tcw_trace_read(long long* h_out /*[192]*/) {
    TSM_CUDA(cudaDeviceSynchronize());
    TSM_CUDA(cudaMemcpyFromSymbol(h_out, tcw::g_tcw_trace, sizeof(long long) * 192));
    return TSM_OK;
}
#endif
