// mlp_tc_epoch4.cuh -- the persistent epoch kernel with FOUR tile groups per SM (included by mlp_tc.cu after mlp_tc_kernels.cuh,
// inside namespace tsm::mlptc).
//
// tc_ppo_epoch_kernel (mlp_tc_kernels.cuh) keeps two 128-row tiles in flight per SM: a 65 536-row minibatch is 512 tiles over
// 296 groups, i.e. TWO rounds of a ~9 k-cycle dependent chain (five waited GEMM stages whose hand-overs are latency, not work).
// This kernel keeps four tiles in flight, so the same minibatch is ONE round (592 slots): a group runs one tile per minibatch.
// What makes four groups fit:
//   shared memory (45 KB per group instead of 113 KB): the data gradients are written IN PLACE -- D2 over H2, D1 over H1 (the
//     stacked operands [H2;H1] and [D2;D1] keep their addresses) -- so every weight-gradient product is issued while its
//     activation operand still exists and is WAITED before the epilogue that overwrites it; one X buffer; the record staging
//     area aliases H1 (the next tile's records land between tiles: during the flush / barrier / optimizer phase of a minibatch);
//     one weight image per CTA instead of one per group.
//   TMEM (496 of 512 columns): 64 columns per group (the head outputs alias the first 16 of the stage accumulator), ONE set
//     of weight-gradient accumulators for the whole CTA (112 columns: accumulating tcgen05.mma issued by different threads of a
//     CTA into one accumulator do not race -- tools/tc_shared_acc_probe.py), the optimizer moments in 2 x 16 columns per group.
// Everything else (records, loss, flush through a swizzled staging copy, one grid barrier, the optimizer step in every CTA, the
// cross-rank exchange) is tc_ppo_epoch_kernel's.

constexpr int G4_GROUPS = 4;
constexpr int G4_THREADS = TC_THREADS * G4_GROUPS;   // 512
constexpr int G4_H2 = 0;                              // 16 KB: H2, then D2
constexpr int G4_H1 = G4_H2 + 8 * CHUNK_B;            // 16 KB: record staging (between tiles), H1, then D1
constexpr int G4_X = G4_H1 + 8 * CHUNK_B;             // 8 KB
constexpr int G4_G = G4_X + 4 * CHUNK_B;              // 4 KB (chunk 1 stays zero)
constexpr int G4_RED = G4_G + 2 * CHUNK_B;            // 4 warps x 8 floats + 4 x 3 doubles
constexpr int G4_BAR = G4_RED + 384;                  // mbarriers: +0 waited MMAs, +8 weight-gradient MMAs, +24 records
constexpr int G4_STRIDE = (G4_BAR + 64 + 1023) & ~1023;
constexpr int G4_REC = G4_H1;
constexpr int G4_IMG = G4_GROUPS * G4_STRIDE;         // the CTA's weight image (adam.cuh layout)
constexpr int G4_W1 = G4_IMG + opt::IMG_W1 * 2, G4_W2 = G4_IMG + opt::IMG_W2 * 2, G4_W3 = G4_IMG + opt::IMG_W3 * 2;
constexpr int G4_B2I = G4_IMG + opt::IMG_B2 * 2, G4_B3I = G4_IMG + opt::IMG_B3 * 2;
constexpr int G4_SHARED = (G4_IMG + opt::IMG_BYTES + 127) & ~127;  // 16 norm partials, 4 coefficients, 4 flags, TMEM base at +124
constexpr int G4_TOTAL = G4_SHARED + 128;
constexpr int G4_STAGE = G4_H2;                       // group 0's H2 + H1: fp32 copy of the flat gradient (<= 32 KB)
static_assert(G4_TOTAL <= 232448, "four groups must fit the 227 KB of one CTA");
static_assert(EP_G_STRIDE * 4 <= 16 * CHUNK_B, "the staging copy lives in one group's H2 + H1");
// TMEM columns
constexpr uint32_t G4_TM_GRP = 64;     // group g: stage accumulator at 64 g (head outputs = its first 16 columns)
constexpr uint32_t G4_TM_GW2 = 256;    // 64: [D2;D1]^T H1   (rows 0..63 = dW2)
constexpr uint32_t G4_TM_GW1 = 320;    // 32: [D2;D1]^T [X|1] (rows 64..127 = dW1 | db1, rows 0..63 column 31 = db2)
constexpr uint32_t G4_TM_GW3 = 352;    // 16: [H2;H1]^T G    (rows 0..63 = dW3^T)
constexpr uint32_t G4_TM_M = 368;      // exp_avg: 16 columns per group
constexpr uint32_t G4_TM_V = 432;      // exp_avg_sq: 16 columns per group
constexpr int G4_OPT = 4;              // float4 of the flat parameter vector per thread: j = t + u * 512

__device__ __forceinline__ void g4_group_sync(int grp) {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(TC_THREADS) : "memory");
}
__device__ __forceinline__ void g4_sync_for_mma(int grp) {
    tc::fence_async_smem();
    tc::fence_before_sync();
    g4_group_sync(grp);
    tc::fence_after_sync();
}
// all 128 threads of a group: gather the records of the group's next tile into the staging area (= H1: the caller knows that
// every MMA and every thread is done with it)
__device__ __forceinline__ void g4_gather_issue(unsigned char* smem, const CUtensorMap* tmap, int32_t slot) {
    uint64_t* rbar = reinterpret_cast<uint64_t*>(smem + G4_BAR) + 3;
    const int lane = tc_tid() & 31, base = lane & ~3;
    const int s0 = __shfl_sync(0xffffffffu, slot, base), s1 = __shfl_sync(0xffffffffu, slot, base + 1);
    const int s2 = __shfl_sync(0xffffffffu, slot, base + 2), s3 = __shfl_sync(0xffffffffu, slot, base + 3);
    if (tc_tid() == 0) tc::mbar_expect_tx(rbar, TC_ROWS * REC_BYTES);
    if ((lane & 3) == 0) tc::tma_gather4(smem + G4_REC + tc_tid() * REC_BYTES, tmap, 0, s0, s1, s2, s3, rbar);
}
__device__ __forceinline__ void g4_gather_take(unsigned char* smem, bool valid, uint32_t& rphase, RowScalars& d) {
    uint64_t* rbar = reinterpret_cast<uint64_t*>(smem + G4_BAR) + 3;
    tc::mbar_wait(rbar, rphase);
    rphase ^= 1;
    const uint4* rec = reinterpret_cast<const uint4*>(smem + G4_REC + tc_tid() * REC_BYTES);
    const uint4 zero = make_uint4(0, 0, 0, 0);
    uint4 x[4], sc = zero;
    uint32_t vo = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) x[c] = valid ? rec[c] : zero;
    if (valid) {
        sc = rec[4];
        vo = *reinterpret_cast<const uint32_t*>(rec + 5);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(smem + G4_X + c * CHUNK_B + tc_tid() * 16) = x[c];
    d.adv = __uint_as_float(sc.x); d.ret = __uint_as_float(sc.y); d.logp_old = __uint_as_float(sc.z);
    d.act = (int)sc.w;
    d.v_old = __uint_as_float(vo);
}
// relu / mask epilogues of the group's stage accumulator (64 columns of this thread's lane) with explicit TMEM address
__device__ __forceinline__ void g4_epilogue_relu(unsigned char* smem, uint32_t tacc, int dst) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(tacc + 32 * q, v);
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
// accumulator masked by (activation image > 0), written over that activation image.  before_store() runs once, after the first
// half of the row has been computed and before anything is stored: the caller waits there for the weight-gradient product that
// still reads the activation image (the wait then overlaps the TMEM load and the mask reads).
template <typename F>
__device__ __forceinline__ void g4_epilogue_mask_inplace(unsigned char* smem, uint32_t tacc, int img, F before_store) {
    const __half2 zero = __float2half2_rn(0.f);
    uint4 w0[4];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float v[32];
        tc::tmem_ld32(tacc + 32 * q, v);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            uint4* cell = reinterpret_cast<uint4*>(smem + img + (4 * q + h) * CHUNK_B + tc_tid() * 16);
            const uint4 m = *cell;
            const __half2* mh = reinterpret_cast<const __half2*>(&m);
            uint4 w;
            w.x = pack_h2(v[8 * h + 0], v[8 * h + 1]) & __hgt2_mask(mh[0], zero);
            w.y = pack_h2(v[8 * h + 2], v[8 * h + 3]) & __hgt2_mask(mh[1], zero);
            w.z = pack_h2(v[8 * h + 4], v[8 * h + 5]) & __hgt2_mask(mh[2], zero);
            w.w = pack_h2(v[8 * h + 6], v[8 * h + 7]) & __hgt2_mask(mh[3], zero);
            if (q == 0) w0[h] = w;
            else *cell = w;
        }
        if (q == 0) {
            before_store();
#pragma unroll
            for (int h = 0; h < 4; ++h) *reinterpret_cast<uint4*>(smem + img + h * CHUNK_B + tc_tid() * 16) = w0[h];
        }
    }
}

#ifdef TSM_TC_TRACE
__device__ unsigned long long g_trace_cta[160][8];  // globaltimer (ns) stamps of every CTA in minibatch 2
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define TRACE_CTA(m, k) do { if (threadIdx.x == 0 && (m) == 2 && blockIdx.x < 160) g_trace_cta[blockIdx.x][k] = gtime(); } while (0)
__device__ unsigned long long g_trace_grp[4][8];  // the four groups of CTA 0 in minibatch 2
#define TRACE_GRP(m, k) do { if ((threadIdx.x & 127) == 0 && (m) == 2 && blockIdx.x == 0) g_trace_grp[threadIdx.x >> 7][k] = gtime(); } while (0)
#else
#define TRACE_CTA(m, k) do { } while (0)
#define TRACE_GRP(m, k) do { } while (0)
#endif

__global__ void __launch_bounds__(G4_THREADS, 1) tc_ppo_epoch4_kernel(TcEpochArgs e, const __grid_constant__ CUtensorMap rec_map) {
    extern __shared__ __align__(1024) unsigned char smem_all[];
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    constexpr int EPT = G4_THREADS, OPT = G4_OPT;
    const TcPpoArgs& a = e.p;  // (never written: the argument struct stays in the constant bank)
    const int grp = (int)(threadIdx.x >> 7), r = tc_tid(), rc = (int)threadIdx.x;
    unsigned char* smem = smem_all + grp * G4_STRIDE;
    float* s_red = reinterpret_cast<float*>(smem_all + G4_SHARED);
    float* s_coef = s_red + 16;
    int* s_has = reinterpret_cast<int*>(s_coef + 4);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_all + G4_SHARED + 124);
    // virtual CTA = group; group g of CTA c is number g * gridDim + c, so that a minibatch of fewer than 4 * gridDim tiles gives
    // every SM three or four tiles instead of leaving SMs idle (the SMs with three flush early, ahead of the burst of reductions)
    const int64_t vb = (int64_t)grp * gridDim.x + blockIdx.x, vg = (int64_t)gridDim.x * G4_GROUPS;
    constexpr uint32_t ID64 = tc::make_idesc_f16(128, 64, 0, 0, 0, 0);
    constexpr uint32_t ID16 = tc::make_idesc_f16(128, 16, 0, 0, 0, 0);
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
    // ---- set-up: barriers, G image, TMEM, the CTA's weight image ----
    {
        uint64_t* bar = reinterpret_cast<uint64_t*>(smem + G4_BAR);
        for (int i = r; i < 2 * CHUNK_B / 16; i += TC_THREADS) *reinterpret_cast<uint4*>(smem + G4_G + i * 16) = make_uint4(0, 0, 0, 0);
        if ((threadIdx.x >> 5) == 0) tc::tmem_alloc(tmem_slot, 512);
        if (r == 0) {
            tc::mbar_init(bar, 1);
            tc::mbar_init(bar + 1, 4);  // every warp of the group issues (and commits) a quarter of each weight-gradient product
            tc::mbar_init(bar + 3, 1);
            tc::fence_barrier_init();
        }
        const uint4* src = reinterpret_cast<const uint4*>(a.image);
        uint4* dst = reinterpret_cast<uint4*>(smem_all + G4_IMG);
        for (int i = rc; i < opt::IMG_BYTES / 16; i += EPT) dst[i] = __ldcg(src + i);
    }
    // this group's tile sequence over the whole epoch: (m, tile) -> next (m, tile); false at the end (32-bit tile arithmetic)
    const int n_mb32 = (int)e.n_mb, vb32 = (int)vb, vg32 = (int)vg;
    const int len_full = (int)e.mb_size, len_last = (int)(e.total - (e.n_mb - 1) * e.mb_size);
    const int tiles_full = (len_full + TC_ROWS - 1) / TC_ROWS, tiles_last = (len_last + TC_ROWS - 1) / TC_ROWS;
    auto advance = [&](int& m_, int& t_) -> bool {
        const int t32 = t_ + vg32;
        if (t32 < (m_ == n_mb32 - 1 ? tiles_last : tiles_full)) { t_ = t32; return true; }
        for (int mm = m_ + 1; mm < n_mb32; ++mm)
            if (vb32 < (mm == n_mb32 - 1 ? tiles_last : tiles_full)) { m_ = mm; t_ = vb32; return true; }
        return false;
    };
    auto slot_at = [&](int m_, int t_) -> int32_t {  // record slot of this thread's row in tile (m_, t_); 0 for a row past the end
        const int q = t_ * TC_ROWS + r;
        if (q >= (m_ == n_mb32 - 1 ? len_last : len_full)) return 0;
        const int64_t p = epoch_start + (int64_t)m_ * e.mb_size + q;
        return a.slots ? __ldg(a.slots + p) : tc_slot(a, p);
    };
    uint32_t rphase = 0, gphase = 0;  // parities of the record barrier and of the weight-gradient barrier
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();  // barriers, TMEM base, weight image
    tc::fence_after_sync();
    // (nm, nt): the group's next tile; its records are gathered as soon as the staging area (= H1) is free
    int nm = 0, nt = vb32;
    bool n_ok = vb32 < (n_mb32 == 1 ? tiles_last : tiles_full);
    if (!n_ok) n_ok = advance(nm, nt);
    if (n_ok) g4_gather_issue(smem, &rec_map, slot_at(nm, nt));
    int32_t sl_next = 0;  // slot of this thread's row in the tile AFTER (nm, nt), fetched a tile ahead of its gather
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const uint32_t lane_base = tmem + ((uint32_t)(((threadIdx.x >> 5) & 3) * 32) << 16);
    const uint32_t tacc = lane_base + G4_TM_GRP * grp;   // this thread's lane of the group's stage accumulator
    const uint32_t dacc = tmem + G4_TM_GRP * grp;        // the same accumulator as an MMA destination
    const uint32_t sb4 = __shfl_sync(0xffffffffu, tc::smem_u32(smem) >> 4, 0);      // group map, descriptor units
    const uint32_t wb4 = __shfl_sync(0xffffffffu, tc::smem_u32(smem_all) >> 4, 0);  // CTA map (weight image)
    const int wq = __shfl_sync(0xffffffffu, (int)((threadIdx.x >> 5) & 3), 0);  // warp of the group (a broadcast: uniform for the compiler)
    const bool w0 = wq == 0;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + G4_BAR);
    uint32_t phase = 0;
    auto wait_mma = [&]() { tc::mbar_wait(bar, phase); phase ^= 1; tc::fence_after_sync(); };
    auto wait_gw = [&]() { tc::mbar_wait(bar + 1, gphase); gphase ^= 1; tc::fence_after_sync(); };
    // ---- this CTA's copy of the optimizer state: fp32 masters in registers, moments in TMEM (16 + 16 columns per group) ----
    float4 Pr[OPT];
    uint2* q_tab = e.q_table + (size_t)rc * (OPT + 1);
    {
        uint2 Q[OPT];
        uint32_t q_contig = 0;
        float Mo[16], Vo[16];
#pragma unroll
        for (int u = 0; u < OPT; ++u) {
            const int i = 4 * (rc + u * EPT);
            uint32_t q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = i + k < n ? (uint32_t)ep_image_index(i + k, o, a.in_dim) : 0xffffu;
            Q[u] = make_uint2(q[0] | (q[1] << 16), q[2] | (q[3] << 16));
            if (q[0] != 0xffffu && (q[0] & 3) == 0 && q[1] == q[0] + 1 && q[2] == q[0] + 2 && q[3] == q[0] + 3) q_contig |= 1u << u;
            auto get = [&](const float* src, int k) { return i + k < n ? __ldcg(src + i + k) : 0.f; };
            Pr[u] = make_float4(get(ad.p, 0), get(ad.p, 1), get(ad.p, 2), get(ad.p, 3));
#pragma unroll
            for (int k = 0; k < 4; ++k) { Mo[4 * u + k] = get(ad.m, k); Vo[4 * u + k] = get(ad.v, k); }
        }
#pragma unroll
        for (int u = 0; u < OPT; ++u) q_tab[u] = Q[u];
        q_tab[OPT] = make_uint2(q_contig, 0u);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            float t8[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) t8[k] = Mo[8 * c + k];
            tc::tmem_st8(lane_base + G4_TM_M + 16 * grp + 8 * c, t8);
#pragma unroll
            for (int k = 0; k < 8; ++k) t8[k] = Vo[8 * c + k];
            tc::tmem_st8(lane_base + G4_TM_V + 16 * grp + 8 * c, t8);
        }
        // the CTA's weight-gradient accumulators start at zero: group g clears 32 of the 112 columns (g = 3: the last 16)
        float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (grp < 3 || c < 2) tc::tmem_st8(lane_base + G4_TM_GW2 + 32 * grp + 8 * c, z8);
    }
    tc::tmem_st_wait();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();

    bool need_issue = false;  // (nm, nt) exists and its gather has not been issued yet
    float gscale_full = 1.0f, gscale_last = 1.0f;  // gradient scale S of a minibatch: the power of two >= its length (<= 2^20)
    while (gscale_full < (float)len_full && gscale_full < 1048576.0f) gscale_full *= 2.0f;
    while (gscale_last < (float)len_last && gscale_last < 1048576.0f) gscale_last *= 2.0f;
    for (int64_t m = 0; m < e.n_mb; ++m) {
        const int64_t mb_len = (m == e.n_mb - 1) ? e.total - m * e.mb_size : e.mb_size;
        float* g_cur = (m % 3 == 0) ? a.grads : (m % 3 == 1 ? e.grads1 : e.grads2);
        const int n_tiles = (int)((mb_len + TC_ROWS - 1) / TC_ROWS);
        const float inv_n = 1.0f / (float)mb_len;
        const float gscale = m == e.n_mb - 1 ? gscale_last : gscale_full;
        const float inv_s = 1.0f / gscale;
        float mean = 0.f, rstd = 1.f;
        if ((a.cfg.adv_norm || a.cfg.plain_pg == 2) && e.mb_stats) { mean = __ldcg(e.mb_stats + 2 * m); rstd = __ldcg(e.mb_stats + 2 * m + 1); }
        float gb3[OUTP];
#pragma unroll
        for (int j = 0; j < OUTP; ++j) gb3[j] = 0.f;
        float f_clip = 0.f, f_vf = 0.f, f_ent = 0.f;
        double l_clip = 0.0, l_vf = 0.0, l_ent = 0.0;
        uint32_t it = 0;
        bool gw_pending = false;  // the tile's last weight-gradient commit has not been waited yet
        [[maybe_unused]] const int erow = (m == 2 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? (blockIdx.x != 0) * 2 : -1;
        TRACE(erow, 40);
        TRACE_CTA(m, 0);
#ifdef TSM_TC_TRACE
        if (threadIdx.x == 0 && blockIdx.x == 0 && m < 64) g_trace[1][m] = clock64();  // start of every minibatch (CTA 0)
#endif
        for (int tile = vb32; tile < n_tiles; tile += vg32, ++it) {
            const bool valid = (int64_t)tile * TC_ROWS + r < mb_len;
            [[maybe_unused]] const int trow = (erow >= 0 && it < 2) ? erow + (int)it : -1;
            TRACE(trow, 0);
            TRACE_GRP(m, 0);
            if (gw_pending) { wait_gw(); gw_pending = false; }  // (only when the gather of this tile could not be issued earlier)
            RowScalars d;
#ifdef TSM_TC_TRACE
            tc::mbar_wait(reinterpret_cast<uint64_t*>(smem + G4_BAR) + 3, rphase);  // (stamp only: the wait inside the take then returns at once)
            TRACE(trow, 26);
#endif
            g4_gather_take(smem, valid, rphase, d);
            TRACE(trow, 27);
            const float d_adv = d.adv, d_ret = d.ret, d_logp_old = d.logp_old, d_v_old = d.v_old;
            const int d_act = d.act;
            // the group's next tile: its slot now (the load's issue stalls on address translation), its gather once H1 is free
            n_ok = advance(nm, nt);
            need_issue = n_ok;
            const bool same_mb = n_ok && nm == (int)m;
            TRACE(trow, 1);
            TRACE_GRP(m, 1);
            g4_sync_for_mma(grp);  // X complete; every thread is past its reads of the staging rows (= H1)
            TRACE(trow, 2);
            // ---- layer 1 ----
            if (tc_issuer(w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_XF / 16; ++ks)
                    tc::mma_f16(dacc, desc_act_k(sb4, G4_X, ks), desc_w_k(wb4, G4_W1, TC_H, ks), ID64, ks > 0);
                tc::commit(bar);
            }
            if (n_ok) sl_next = slot_at(nm, nt);  // (the load's issue stalls on address translation: inside the wait for layer 1)
            TRACE(trow, 3);
            wait_mma();
            TRACE(trow, 4);
            g4_epilogue_relu(smem, tacc, G4_H1);
            TRACE(trow, 5);
            g4_sync_for_mma(grp);
            TRACE(trow, 6);
            // ---- layer 2 ----
            if (tc_issuer(w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_H / 16; ++ks)
                    tc::mma_f16(dacc, desc_act_k(sb4, G4_H1, ks), desc_w_k(wb4, G4_W2, TC_H, ks), ID64, ks > 0);
                tc::mma_f16(dacc, desc_act_k(sb4, G4_X, 1), desc_w_k(wb4, G4_B2I, TC_H, 0), ID64, 1);  // + b2
                tc::commit(bar);
            }
            TRACE(trow, 7);
            wait_mma();
            TRACE(trow, 8);
            g4_epilogue_relu(smem, tacc, G4_H2);
            TRACE(trow, 9);
            g4_sync_for_mma(grp);
            TRACE(trow, 10);
            // ---- head: 16 columns over the (dead) stage accumulator ----
            if (tc_issuer(w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_H / 16; ++ks)
                    tc::mma_f16(dacc, desc_act_k(sb4, G4_H2, ks), desc_w_k(wb4, G4_W3, TC_OF, ks), ID16, ks > 0);
                tc::mma_f16(dacc, desc_act_k(sb4, G4_X, 1), desc_w_k(wb4, G4_B3I, TC_OF, 0), ID16, 1);  // + b3
                tc::commit(bar);
            }
            TRACE(trow, 11);
            wait_mma();
            TRACE(trow, 12);
            {
                float z[OUTP], dz[OUTP];
                {
                    float o16[16];
                    tc::tmem_ld16(tacc, o16);
#pragma unroll
                    for (int j = 0; j < OUTP; ++j) z[j] = o16[j];
                }
                const RowLoss rl = simple5 ? row_loss<TC_FAST_LOSS, true, 5>(z, 5, 5, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz)
                                 : simple_loss ? row_loss<TC_FAST_LOSS, true>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz)
                                               : row_loss<TC_FAST_LOSS>(z, a.n_act, a.value_index, d_act, d_adv, d_ret, d_logp_old, d_v_old, mean, rstd, a.cfg, inv_n, dz);
                if (valid) { f_clip += rl.clip; f_vf += rl.vf; f_ent += rl.ent; }
                if ((it & 15) == 15) {
                    l_clip += (double)f_clip; l_vf += (double)f_vf; l_ent += (double)f_ent;
                    f_clip = f_vf = f_ent = 0.f;
                }
                float g[OUTP];
#pragma unroll
                for (int j = 0; j < OUTP; ++j) {
                    if (!valid) dz[j] = 0.f;
                    gb3[j] += dz[j];
                    g[j] = fminf(fmaxf(dz[j] * gscale, -65000.f), 65000.f);
                }
                uint4 w;
                w.x = pack_h2(g[0], g[1]); w.y = pack_h2(g[2], g[3]); w.z = pack_h2(g[4], g[5]); w.w = pack_h2(g[6], g[7]);
                *reinterpret_cast<uint4*>(smem + G4_G + r * 16) = w;
            }
            TRACE(trow, 13);
            g4_sync_for_mma(grp);
            TRACE(trow, 14);
            // ---- DH2 = G W3, and [H2;H1]^T G while H2 still exists (waited before D2 replaces it) ----
            if (tc_issuer(w0)) {
                tc::mma_f16(dacc, desc_act_k(sb4, G4_G, 0), desc_w_mn(wb4, G4_W3, TC_OF, 0), ID_DH, 0);
                tc::commit(bar);
            }
            // The weight-gradient products accumulate (order does not matter) and accumulating MMAs of different threads do not
            // race: every warp issues two of the eight K-steps, so no warp is late for the epilogue by a whole product's issue.
            if (tc::elect_one()) {
#pragma unroll
                for (int k2 = 0; k2 < 2; ++k2)
                    tc::mma_f16(tmem + G4_TM_GW3, desc_act_mn(sb4, G4_H2, 2 * wq + k2), desc_act_mn(sb4, G4_G, 2 * wq + k2), ID_G16, 1);
                tc::commit(bar + 1);
            }
            __syncwarp();
            TRACE(trow, 15);
            wait_mma();
            TRACE(trow, 16);
            g4_epilogue_mask_inplace(smem, tacc, G4_H2, wait_gw);
            TRACE(trow, 17);
            g4_sync_for_mma(grp);
            TRACE(trow, 18);
            // ---- DH1 = D2 W2, and [D2;.]^T H1 while H1 still exists ----
            if (tc_issuer(w0)) {
#pragma unroll
                for (int ks = 0; ks < TC_H / 16; ++ks)
                    tc::mma_f16(dacc, desc_act_k(sb4, G4_H2, ks), desc_w_mn(wb4, G4_W2, TC_H, ks), ID_DH, ks > 0);
                tc::commit(bar);
            }
            if (tc::elect_one()) {
#pragma unroll
                for (int k2 = 0; k2 < 2; ++k2)
                    tc::mma_f16(tmem + G4_TM_GW2, desc_act_mn(sb4, G4_H2, 2 * wq + k2), desc_act_mn(sb4, G4_H1, 2 * wq + k2), ID_G64, 1);
                tc::commit(bar + 1);
            }
            __syncwarp();
            TRACE(trow, 19);
            wait_mma();
            TRACE(trow, 20);
            g4_epilogue_mask_inplace(smem, tacc, G4_H1, wait_gw);
            TRACE(trow, 21);
            g4_sync_for_mma(grp);
            TRACE(trow, 22);
            // ---- [D2;D1]^T [X|1] ----
            if (tc::elect_one()) {
#pragma unroll
                for (int k2 = 0; k2 < 2; ++k2)
                    tc::mma_f16(tmem + G4_TM_GW1, desc_act_mn(sb4, G4_H2, 2 * wq + k2), desc_act_mn(sb4, G4_X, 2 * wq + k2), ID_G32, 1);
                tc::commit(bar + 1);
            }
            __syncwarp();
            gw_pending = true;
            TRACE(trow, 23);
            TRACE_GRP(m, 2);
            if (same_mb) {  // more tiles of this minibatch for this group (minibatch > 4 tiles per SM): the gather waits for H1
                wait_gw();
                gw_pending = false;
                g4_gather_issue(smem, &rec_map, sl_next);
                need_issue = false;
            }
        }
        TRACE(erow, 41);
        TRACE_CTA(m, 1);
        // ---- flush: ONE set of accumulators per CTA; all 512 threads stage a share of it, then one set of reductions ----
        l_clip += (double)f_clip; l_vf += (double)f_vf; l_ent += (double)f_ent;
        {
            float* red_f = reinterpret_cast<float*>(smem + G4_RED);
            double* red_d = reinterpret_cast<double*>(smem + G4_RED + 192);
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
            if (r == 0) s_has[grp] = it > 0;
        }
        TRACE(erow, 49);
        if (gw_pending) { wait_gw(); gw_pending = false; }  // this group's weight-gradient MMAs have completed
        TRACE(erow, 50);
        TRACE_GRP(m, 3);
        tc::fence_before_sync();
        __syncthreads();  // ... and so have every group's
        tc::fence_after_sync();
        TRACE_GRP(m, 4);
        const bool any = (s_has[0] | s_has[1] | s_has[2] | s_has[3]) != 0;
        if (any) {
            // staging copy (fp32, flat gradient layout, W2 block XOR-swizzled: stage_pos4) in group 0's H2 + H1.  Lanes 0..63:
            // group g takes dW2 columns 16 g.., group 0 also dW3^T, group 1 also db2; lanes 64..127: groups 0 / 1 take dW1
            // columns 0..15 / 16..31 (column 31 = db1).
            float* stage = reinterpret_cast<float*>(smem_all + G4_STAGE);
            float v[16];
            if (r < TC_H) {
                tc::tmem_ld16(lane_base + G4_TM_GW2 + 16 * grp, v);
#pragma unroll
                for (int h = 0; h < 4; ++h)
                    *reinterpret_cast<float4*>(stage + o.w2 + (r << 6) + (((4 * grp + h) ^ (r & 15)) << 2)) =
                        make_float4(v[4 * h], v[4 * h + 1], v[4 * h + 2], v[4 * h + 3]);
                if (grp == 0) {
                    tc::tmem_ld16(lane_base + G4_TM_GW3, v);
#pragma unroll
                    for (int k = 0; k < OUTP; ++k)
                        if (k < a.out_dim) stage[o.w3 + k * TC_H + r] = v[k];
                } else if (grp == 1) {
                    tc::tmem_ld16(lane_base + G4_TM_GW1 + 16, v);
                    stage[o.b2 + r] = v[15];
                }
            } else if (grp < 2) {
                tc::tmem_ld16(lane_base + G4_TM_GW1 + 16 * grp, v);
                const int rr = r - TC_H;
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (16 * grp + k < a.in_dim) stage[o.w1 + rr * a.in_dim + 16 * grp + k] = v[k];
                if (grp == 1) stage[o.b1 + rr] = v[15];
            }
            tc::fence_before_sync();
            __syncthreads();
            tc::fence_after_sync();
            TRACE_GRP(m, 5);
            {   // the accumulators restart at zero for the next minibatch (their columns were all read above)
                float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (grp < 3 || c < 2) tc::tmem_st8(lane_base + G4_TM_GW2 + 32 * grp + 8 * c, z8);
            }
            const int w2_lo = (int)o.w2, w2_hi = (int)o.b2, n_stage = (int)o.b3;
            for (int i = 4 * rc; i < n_stage; i += 4 * EPT) {
                float4 u = *reinterpret_cast<const float4*>(stage + stage_pos4(i, w2_lo, w2_hi));
                u.x *= inv_s; u.y *= inv_s; u.z *= inv_s; u.w *= inv_s;
                red_add4(g_cur + i, u);
            }
            if (rc < a.out_dim) {
                float t = 0.f;
#pragma unroll
                for (int g = 0; g < G4_GROUPS; ++g) {
                    const float* rf = reinterpret_cast<const float*>(smem_all + g * G4_STRIDE + G4_RED);
                    t += rf[rc] + rf[OUTP + rc] + rf[2 * OUTP + rc] + rf[3 * OUTP + rc];
                }
                atomicAdd(g_cur + o.b3 + rc, t);
            }
            if (e.loss_sums && rc < 3) {
                double t = 0.0;
#pragma unroll
                for (int g = 0; g < G4_GROUPS; ++g) {
                    const double* rd = reinterpret_cast<const double*>(smem_all + g * G4_STRIDE + G4_RED + 192);
                    t += rd[rc] + rd[3 + rc] + rd[6 + rc] + rd[9 + rc];
                }
                atomicAdd(e.loss_sums + 4 * m + rc, t);
            }
        }
        TRACE(erow, 42);
        TRACE_CTA(m, 2);
        TRACE_GRP(m, 6);
        // ---- grid barrier with the cross-rank gradient exchange folded into it (tc_ppo_epoch_kernel's, see there) ----
        uint2 Q[OPT];  // (in flight across the barrier)
#pragma unroll
        for (int u = 0; u < OPT; ++u) Q[u] = q_tab[u];
        const uint32_t q_contig = q_tab[OPT].x;
        const uint32_t seq = ad.seq + (uint32_t)m;
        __threadfence();
        tc::fence_async_smem();
        __syncthreads();  // the staging copy has been read: group 0's H1 is free again
        const uint32_t arrivals = (uint32_t)(m + 1) * gridDim.x;  // counter value once every CTA has arrived at this barrier
        TRACE_CTA(m, 3);
        if (rc == 0) atomicAdd(e.sync_words, 1u);
        // the records of every group's next tile land while the barrier, the exchange and the optimizer step run (issued after
        // the arrival: the 32 gather4 copies of a group block their warps for ~900 cycles)
        if (need_issue) {
            g4_gather_issue(smem, &rec_map, sl_next);
            need_issue = false;
        }
        if (rc == 64) {  // the bias corrections of this step (double-precision powers): in the shadow of the barrier wait
            const int64_t step = step0 + m + 1;
            s_coef[1] = ad.lr / (float)(1.0 - opt::ipow((double)ad.beta1, step));
            s_coef[2] = (float)(1.0 / sqrt(1.0 - opt::ipow((double)ad.beta2, step)));
        }
        if (rc == 0) {
            const uint32_t* w = e.sync_words;
            const long long t0 = clock64();
            uint32_t v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(w) : "memory");
                if ((int32_t)(v - arrivals) >= 0) break;
                if (clock64() - t0 > 4000000000LL) {  // ~2 s: a CTA or a peer rank is missing; report instead of hanging the device
                    if (ad.status) *ad.status = 1;
                    break;
                }
            } while (true);
        }
        __syncthreads();
        TRACE(erow, 43);
        TRACE_CTA(m, 4);
        const uint4* r_base = ad.world > 1 ? ad.peer_ll[ad.rank] + (int64_t)(seq & 1u) * ad.world * ad.n_stride / 2 : nullptr;
        const float* g_src = g_cur;  // where the (summed) gradient is read from below
        if (ad.world > 1) {
            // Slice owners wait for the (value, step) pairs of their slice from every peer, sum the `world` contributions
            // (rank order) into g_total, one more grid barrier, and everybody reads the total once.  Every CTA summing all slots itself would move world x 28 KB x grid per step,
            // and reads of the peer-written receive area are slow (they miss L2: +3 us per step at two ranks when all 148
            // CTAs read the whole slot, +1 us for one slice each); the extra barrier costs ~1.3 us.
            // One thread per (float4 of the slice, rank): the thread of this rank loads the local value and pushes it to every
            // peer, the thread of peer q polls q's pairs -- all peers' polls (each an L2 miss) run side by side instead of one
            // after the other in one thread (+6.7 us per minibatch at 8 ranks before).  The contributions meet in shared memory
            // (group 0's H2: free between the flush and the next tile) and are summed in rank order.
            const int64_t n4s = ad.n_stride / 4, per4 = (n4s + gridDim.x - 1) / gridDim.x;
            const int64_t lo4 = (int64_t)blockIdx.x * per4, hi4 = lo4 + per4 < n4s ? lo4 + per4 : n4s;
            const int cnt = hi4 > lo4 ? (int)(hi4 - lo4) : 0;
            float4* xch = reinterpret_cast<float4*>(smem_all + G4_H2);  // [cnt][world]
            for (int idx = rc; idx < cnt * ad.world; idx += EPT) {
                const int jl = idx / ad.world, q = idx - jl * ad.world;
                const int64_t j = lo4 + jl;
                float4 y;
                if (q == ad.rank) {
                    if (4 * j + 3 < n) {
                        y = __ldcg(reinterpret_cast<const float4*>(g_cur) + j);
                    } else {
                        auto get = [&](int k) { return 4 * j + k < n ? __ldcg(g_cur + 4 * j + k) : 0.f; };
                        y = make_float4(get(0), get(1), get(2), get(3));
                    }
                    // this CTA's slice of the local gradient -> slot [seq & 1][rank] of every peer
                    const uint4 plo = make_uint4(__float_as_uint(y.x), seq, __float_as_uint(y.y), seq);
                    const uint4 phi = make_uint4(__float_as_uint(y.z), seq, __float_as_uint(y.w), seq);
                    const int64_t off = ((int64_t)(seq & 1u) * ad.world + ad.rank) * ad.n_stride + 4 * j;
                    for (int p = 0; p < ad.world; ++p)
                        if (p != ad.rank) {
                            uint4* dst = ad.peer_ll[p] + off / 2;
                            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(plo.x), "r"(plo.y), "r"(plo.z), "r"(plo.w) : "memory");
                            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 1), "r"(phi.x), "r"(phi.y), "r"(phi.z), "r"(phi.w) : "memory");
                        }
                } else {
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
                xch[idx] = y;
            }
            __syncthreads();
            for (int jl = rc; jl < cnt; jl += EPT) {
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int q = 0; q < ad.world; ++q) {  // rank order: the same sum on every rank
                    const float4 y = xch[jl * ad.world + q];
                    sum.x += y.x; sum.y += y.y; sum.z += y.z; sum.w += y.w;
                }
                __stcg(reinterpret_cast<float4*>(e.g_total) + lo4 + jl, sum);
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
            float Mo[16], Vo[16];  // one TMEM round trip for all moments of this thread
            tc::tmem_st_wait();  // the previous step's stores of these columns and the cleared accumulators (waited here, not there)
            tc::tmem_ld16(lane_base + G4_TM_M + 16 * grp, Mo);
            tc::tmem_ld16(lane_base + G4_TM_V + 16 * grp, Vo);
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
                    *reinterpret_cast<uint2*>(smem_all + G4_IMG + 2 * q) = h;
                    if (last) *reinterpret_cast<uint2*>(ad.image + q) = h;
                } else if (live) {  // W1 rows (in_dim is not a multiple of 8) and the biases: one position at a time
                    const float pp[4] = {Pr[u].x, Pr[u].y, Pr[u].z, Pr[u].w};
                    const uint32_t qq[4] = {Q[u].x & 0xffffu, Q[u].x >> 16, Q[u].y & 0xffffu, Q[u].y >> 16};
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (qq[k] != 0xffffu) {
                            const __half h = __float2half_rn(pp[k]);
                            *reinterpret_cast<__half*>(smem_all + G4_IMG + 2 * qq[k]) = h;
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
            for (int c = 0; c < 2; ++c) {
                float t8[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) t8[k] = Mo[8 * c + k];
                tc::tmem_st8(lane_base + G4_TM_M + 16 * grp + 8 * c, t8);
#pragma unroll
                for (int k = 0; k < 8; ++k) t8[k] = Vo[8 * c + k];
                tc::tmem_st8(lane_base + G4_TM_V + 16 * grp + 8 * c, t8);
            }
        }
        TRACE(erow, 45);
        {   // buffer (m - 1) % 3 has been read by everybody (all CTAs are past barrier m); it is next reduced into after barrier m + 1
            float* g_old = ((m + 2) % 3 == 0) ? a.grads : ((m + 2) % 3 == 1 ? e.grads1 : e.grads2);
            const int64_t per = (n + gridDim.x - 1) / gridDim.x;
            const int64_t lo = (int64_t)blockIdx.x * per, hi = lo + per < n ? lo + per : n;
            for (int64_t i = lo + rc; i < hi; i += EPT) g_old[i] = 0.f;
        }
        tc::fence_async_smem();  // the image entries written above are read by every group's MMAs
        tc::tmem_st_wait();      // the cleared weight-gradient accumulators are in place before any group's next product
        tc::fence_before_sync();
        __syncthreads();
        tc::fence_after_sync();
        TRACE(erow, 47);
        TRACE_CTA(m, 5);
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
    if ((rc >> 5) == 0) tc::tmem_dealloc(*tmem_slot, 512);
}

