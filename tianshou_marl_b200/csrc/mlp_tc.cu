// mlp_tc.cu -- K4/K5 on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Same contract as mlp.cu (forward / sample / evaluate, fused PPO minibatch fwd+loss+bwd) for the
// net shape in(<=31) -> 64 -> 64 -> out(<=8), with every GEMM issued as tcgen05.mma kind::f16
// (fp16 operands in shared memory, fp32 accumulators in TMEM).  See tc.cuh / mlp_tc_kernels.cuh.
#include <algorithm>
#include <cstdlib>
#include <cstdio>
#include "common.cuh"
#include "philox.cuh"
#include <cuda_fp16.h>
#include <cuda.h>  // CUtensorMap (types only: the encoder is reached through cudaGetDriverEntryPoint)
#include "feistel.cuh"
#include "ppo_math.cuh"
#include "adam.cuh"
#include "tc.cuh"

namespace tsm {
namespace mlptc {

using namespace tsm::ppo;

#include "mlp_tc_kernels.cuh"
#include "mlp_tc_epoch4.cuh"

// ---------------------------------------------------------------------------------------------
// self test: one GEMM D[128][N] = A * B from caller-prepared shared-memory images and caller-chosen
// descriptor fields; used by tests/test_gpu_tc.py to pin the descriptor semantics this file relies on
struct SelfArgs {
    const float* a_img; const float* b_img; int a_floats, b_floats;
    uint32_t a_lbo, a_sbo, a_step, b_lbo, b_sbo, b_step;
    int a_mn, b_mn, N, ksteps, kind, a_bf16, b_bf16;
    float* d_out;
};

__global__ void __launch_bounds__(128) selftest_kernel(SelfArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    float* sA = reinterpret_cast<float*>(smem);
    float* sB = sA + ((a.a_floats + 255) / 256) * 256;
    for (int i = threadIdx.x; i < a.a_floats; i += 128) sA[i] = a.a_img[i];
    for (int i = threadIdx.x; i < a.b_floats; i += 128) sB[i] = a.b_img[i];
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tc::tmem_alloc(&tmem_base_s, 256);
    if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_base_s;
    if (a.kind == 3) {
        // kind 3: FOUR issuing threads (one elected lane per warp) accumulate the same fp16 GEMM 16 times each into ONE TMEM
        // accumulator, concurrently: D = 64 * A * B if accumulating tcgen05.mma from different threads of a CTA do not race
        // (the question behind sharing weight-gradient accumulators between tile groups)
        __shared__ uint64_t bars[4];
        if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) tc::mbar_init(&bars[i], 1); tc::fence_barrier_init(); }
        for (int c = 0; c < a.N; c += 8) {
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            tc::tmem_st8(tmem_base + ((uint32_t)(warp * 32) << 16) + c, z);
        }
        tc::tmem_st_wait();
        tc::fence_before_sync();
        __syncthreads();
        tc::fence_after_sync();
        if (tc::elect_one()) {
            const uint32_t idesc = tc::make_idesc_f16(128, a.N, 0, 0, a.a_mn, a.b_mn);
            for (int rep = 0; rep < 16; ++rep)
                for (int ks = 0; ks < a.ksteps; ++ks) {
                    const uint64_t da = tc::make_desc(tc::smem_u32(sA) + ks * a.a_step, a.a_lbo, a.a_sbo);
                    const uint64_t db = tc::make_desc(tc::smem_u32(sB) + ks * a.b_step, a.b_lbo, a.b_sbo);
                    tc::mma_f16(tmem_base, da, db, idesc, 1);
                }
            tc::commit(&bars[warp]);
        }
        __syncwarp();
        for (int i = 0; i < 4; ++i) tc::mbar_wait(&bars[i], 0);
        tc::fence_after_sync();
    } else {
    if (threadIdx.x == 0) {
        const uint32_t idesc = a.kind == 0 ? tc::make_idesc_tf32(128, a.N, a.a_mn, a.b_mn)
                                           : tc::make_idesc_f16(128, a.N, a.a_bf16, a.b_bf16, a.a_mn, a.b_mn);
        for (int ks = 0; ks < a.ksteps; ++ks) {
            const uint64_t da = tc::make_desc(tc::smem_u32(sA) + ks * a.a_step, a.a_lbo, a.a_sbo);
            const uint64_t db = tc::make_desc(tc::smem_u32(sB) + ks * a.b_step, a.b_lbo, a.b_sbo);
            if (a.kind == 0) tc::mma_tf32(tmem_base, da, db, idesc, ks > 0);
            else tc::mma_f16(tmem_base, da, db, idesc, ks > 0);
        }
        tc::commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    }
    for (int c = 0; c < a.N; c += 16) {
        float v[16];
        tc::tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + c, v);
#pragma unroll
        for (int i = 0; i < 16; ++i) a.d_out[(size_t)threadIdx.x * a.N + c + i] = v[i];
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, 256);
}

}  // namespace mlptc
}  // namespace tsm

using namespace tsm;

extern "C" int tsm_tc_selftest(const float* a_img, int a_floats, const float* b_img, int b_floats, const uint32_t* h_desc /*[6]*/,
                               int a_mn_major, int b_mn_major, int n, int ksteps, int kind, int a_bf16, int b_bf16, float* d_out,
                               tsm_stream_t stream) {
    TSM_REQUIRE(a_img && b_img && h_desc && d_out, "pointers != NULL");
    TSM_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0 && ksteps >= 1, "16 <= n <= 256, n % 16 == 0, ksteps >= 1");
    mlptc::SelfArgs a{};
    a.a_img = a_img; a.b_img = b_img; a.a_floats = a_floats; a.b_floats = b_floats;
    a.a_lbo = h_desc[0]; a.a_sbo = h_desc[1]; a.a_step = h_desc[2]; a.b_lbo = h_desc[3]; a.b_sbo = h_desc[4]; a.b_step = h_desc[5];
    a.a_mn = a_mn_major; a.b_mn = b_mn_major; a.N = n; a.ksteps = ksteps; a.d_out = d_out; a.kind = kind; a.a_bf16 = a_bf16; a.b_bf16 = b_bf16;
    const size_t smem = (size_t)(((a_floats + 255) / 256) * 256 + b_floats) * 4 + 1024;
    TSM_REQUIRE(smem <= 200 * 1024, "images too large");
    TSM_CUDA(cudaFuncSetAttribute(mlptc::selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mlptc::selftest_kernel<<<1, 128, smem, as_stream(stream)>>>(a);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

// ---------------------------------------------------------------------------------------------
// batch records (include/tsm_b200.h: TSM_PPO_RECORD_BYTES): one thread per storage slot, streaming
namespace tsm {
namespace mlptc {
// Block = 256 consecutive slots (or 256 (t, row) pairs of a row list).  The observations of consecutive slots are one contiguous
// range and so are their records up to the 32-byte tail of the pitch, so both sides go through shared memory: the block loads the
// range with coalesced 16-byte accesses, every thread builds its record from its row, and the block writes the records out with
// consecutive lanes on consecutive 16-byte pieces (a thread per record would touch 32 lines per instruction with 16 bytes each).
__global__ void __launch_bounds__(256) pack_records_kernel(const float* __restrict__ obs, const int32_t* __restrict__ act,
                                                          const float* __restrict__ adv, const float* __restrict__ ret,
                                                          const float* __restrict__ logp_old, const float* __restrict__ v_old,
                                                          int64_t n_slots, int in_dim, uint4* __restrict__ records,
                                                          const int64_t* __restrict__ rows, int64_t n_sel, int64_t n_rows) {
    __shared__ __align__(16) float s_buf[256 * TC_XF];  // 32 KB: the block's observations, then its records (256 x 96 bytes)
    const int64_t s_first = (int64_t)blockIdx.x * 256;
    const int64_t s_lin = s_first + threadIdx.x;
    const int n_here = (int)min((int64_t)256, n_slots - s_first);
    const bool valid = threadIdx.x < n_here;
    int64_t s = s_lin;
    if (rows && valid) {  // only the sub-buffers in `rows` (a policy group): thread (t, i) packs slot t * n_rows + rows[i]
        const int64_t t = s_lin / n_sel;
        s = t * n_rows + rows[s_lin - t * n_sel];
    }
    float x[TC_XF];
#pragma unroll
    for (int k = 0; k < TC_XF; ++k) x[k] = 0.f;
    const bool staged = rows == nullptr && (((uintptr_t)(obs + s_first * in_dim)) & 15) == 0;
    if (staged) {
        const int n_f = n_here * in_dim, n4 = n_f >> 2;
        const float4* src = reinterpret_cast<const float4*>(obs + s_first * in_dim);
        for (int i = threadIdx.x; i < n4; i += 256) reinterpret_cast<float4*>(s_buf)[i] = __ldcs(src + i);
        for (int i = (n4 << 2) + threadIdx.x; i < n_f; i += 256) s_buf[i] = __ldcs(obs + s_first * in_dim + i);
        __syncthreads();
        if (valid) {
#pragma unroll
            for (int k = 0; k < TC_XF; ++k)
                if (k < in_dim) x[k] = s_buf[threadIdx.x * in_dim + k];
        }
        __syncthreads();  // the staging area is reused for the records
    } else if (valid) {
        const float* p = obs + s * in_dim;
        if ((in_dim & 1) == 0 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) {
#pragma unroll
            for (int k = 0; k < TC_XF / 2; ++k)
                if (2 * k < in_dim) {
                    const float2 v = __ldcs(reinterpret_cast<const float2*>(p + 2 * k));
                    x[2 * k] = v.x;
                    x[2 * k + 1] = v.y;
                }
        } else {
#pragma unroll
            for (int k = 0; k < TC_XF; ++k)
                if (k < in_dim) x[k] = __ldcs(p + k);
        }
    }
    x[TC_XF - 1] = 1.0f;
    uint4 rec[6];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        rec[c].x = pack_h2(x[8 * c + 0], x[8 * c + 1]); rec[c].y = pack_h2(x[8 * c + 2], x[8 * c + 3]);
        rec[c].z = pack_h2(x[8 * c + 4], x[8 * c + 5]); rec[c].w = pack_h2(x[8 * c + 6], x[8 * c + 7]);
    }
    rec[4] = make_uint4(0, 0, 0, 0);
    rec[5] = make_uint4(0, 0, 0, 0);
    if (valid) {
        rec[4].x = __float_as_uint(adv ? adv[s] : 0.f);
        rec[4].y = __float_as_uint(ret ? ret[s] : 0.f);
        rec[4].z = __float_as_uint(logp_old ? logp_old[s] : 0.f);
        rec[4].w = (uint32_t)(act ? act[s] : 0);
        rec[5].x = __float_as_uint(v_old ? v_old[s] : 0.f);
    }
    if (rows == nullptr) {  // consecutive slots: records leave through shared memory, 16-byte pieces in order
        uint4* s_rec = reinterpret_cast<uint4*>(s_buf);  // [256][6]
#pragma unroll
        for (int c = 0; c < 6; ++c) s_rec[threadIdx.x * 6 + c] = rec[c];
        __syncthreads();
        uint4* out = records + s_first * (REC_PITCH / 16);
        for (int i = threadIdx.x; i < n_here * 6; i += 256) {
            const int r = i / 6, c = i - r * 6;
            out[r * (REC_PITCH / 16) + c] = s_rec[i];
        }
    } else if (valid) {
        uint4* out = records + s * (REC_PITCH / 16);
#pragma unroll
        for (int c = 0; c < 6; ++c) out[c] = rec[c];
    }
}
}  // namespace mlptc
}  // namespace tsm

extern "C" int tsm_ppo_pack_records(const tsm_net* net, const tsm_ppo_batch* batch, void* records, tsm_stream_t stream) {
    TSM_REQUIRE(net && batch && records, "net / batch / records != NULL");
    TSM_REQUIRE(net->in_dim >= 1 && net->in_dim <= 31, "1 <= in_dim <= 31");
    TSM_REQUIRE(batch->obs != nullptr && batch->size > 0 && batch->n_rows > 0, "batch.obs / size / n_rows");
    TSM_REQUIRE((reinterpret_cast<uintptr_t>(records) & 15) == 0, "records must be 16-byte aligned");
    const int64_t n = batch->size * batch->n_rows;
    mlptc::pack_records_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
        batch->obs, batch->act, batch->adv, batch->ret, batch->logp_old, batch->v_old, n, net->in_dim, reinterpret_cast<uint4*>(records),
        nullptr, 0, batch->n_rows);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_ppo_pack_records_rows(const tsm_net* net, const tsm_ppo_batch* batch, void* records, const int64_t* rows, int64_t n_sel,
                                         tsm_stream_t stream) {
    TSM_REQUIRE(net && batch && records && rows, "net / batch / records / rows != NULL");
    TSM_REQUIRE(net->in_dim >= 1 && net->in_dim <= 31, "1 <= in_dim <= 31");
    TSM_REQUIRE(batch->obs != nullptr && batch->size > 0 && batch->n_rows > 0, "batch.obs / size / n_rows");
    TSM_REQUIRE((reinterpret_cast<uintptr_t>(records) & 15) == 0, "records must be 16-byte aligned");
    TSM_REQUIRE(n_sel >= 0 && n_sel <= batch->n_rows, "0 <= n_sel <= n_rows");
    if (n_sel == 0) return TSM_OK;
    const int64_t n = batch->size * n_sel;
    mlptc::pack_records_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
        batch->obs, batch->act, batch->adv, batch->ret, batch->logp_old, batch->v_old, n, net->in_dim, reinterpret_cast<uint4*>(records),
        rows, n_sel, batch->n_rows);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

// ---------------------------------------------------------------------------------------------
// engine 1 entry points (called from the dispatchers in mlp.cu)
namespace tsm {

static int tc_check_net(const tsm_net* n, const char* fn) {
    if (n->hidden != 64 || n->in_dim < 1 || n->in_dim > 31 || n->out_dim < 1 || n->out_dim > 8)
        return fail(TSM_ERR_UNSUPPORTED, "%s: the tcgen05 engine needs hidden == 64, in_dim <= 31, out_dim <= 8", fn);
    return TSM_OK;
}

int tc_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter, uint64_t seed,
                   uint64_t counter, float* out, int32_t* act, float* logp, float* value, float* entropy, int64_t rng_row0, cudaStream_t s) {
    int rc = tc_check_net(net, __func__);
    if (rc) return rc;
    mlptc::TcFwdArgs a{};
    a.image = reinterpret_cast<const unsigned char*>(net->tc_image); a.x = x; a.rows = rows; a.n = n;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim; a.n_act = net->n_act; a.value_index = net->value_index;
    a.mode = mode; a.scatter = scatter; a.seed = seed; a.counter = counter; a.rng_row0 = rng_row0; a.probs = net->probs_param;
    a.out = out; a.act = act; a.logp = logp; a.value = value; a.entropy = entropy;
    const size_t smem = mlptc::SM_FWD_TOTAL + 128;  // ~69 KB: three CTAs per SM (3 x 128 TMEM columns)
    static bool attr = false;
    if (!attr) {
        TSM_CUDA(cudaFuncSetAttribute(mlptc::tc_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    const int64_t tiles = ceil_div(n, mlptc::TC_ROWS);
    static const int ctas = getenv("TSM_TC_FWD_CTAS") ? atoi(getenv("TSM_TC_FWD_CTAS")) : 3;  // experiments: CTAs per SM the grid is sized for
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)sm_count() * ctas);
    TSM_CUDA(launch_pdl(g_pdl_hint, mlptc::tc_forward_kernel, grid, mlptc::TC_THREADS, smem, s, a));
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

// tensor map of the record array as a [n_slots][24 x u32] matrix with a {24, 1} box: the operand of the gather4 copies
static int record_map(const void* records, int64_t n_slots, CUtensorMap* out, const char* fn) {
    typedef CUresult (*Encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static Encode encode = nullptr;
    if (!encode) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        TSM_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
        if (!f || q != cudaDriverEntryPointSuccess) return fail(TSM_ERR_UNSUPPORTED, "%s: the driver has no cuTensorMapEncodeTiled", fn);
        encode = reinterpret_cast<Encode>(f);
    }
    static const void* c_rec = nullptr;  // the map only depends on (records, n_slots): keep the last one
    static int64_t c_n = 0;
    static CUtensorMap c_map;
    if (records != c_rec || n_slots != c_n) {
        if (n_slots <= 0 || n_slots >= (int64_t)1 << 31) return fail(TSM_ERR_UNSUPPORTED, "%s: engine 1 needs size * n_rows < 2^31", fn);
        if (reinterpret_cast<uintptr_t>(records) & 127) return fail(TSM_ERR_INVALID, "%s: records must be 128-byte aligned", fn);
        const cuuint64_t dim[2] = {mlptc::REC_BYTES / 4, (cuuint64_t)n_slots}, stride[1] = {mlptc::REC_PITCH};
        const cuuint32_t box[2] = {mlptc::REC_BYTES / 4, 1}, estr[2] = {1, 1};
        const CUresult r = encode(&c_map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(records), dim, stride, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(TSM_ERR_CUDA, "%s: cuTensorMapEncodeTiled failed (%d)", fn, (int)r);
        c_rec = records; c_n = n_slots;
    }
    *out = c_map;
    return TSM_OK;
}

int tc_ppo_minibatch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t mb_start, int64_t mb_len,
                     const float* mb_stats, double* loss_sums, int pdl, cudaStream_t s) {
    int rc = tc_check_net(net, __func__);
    if (rc) return rc;
    mlptc::TcPpoArgs a{};
    a.image = reinterpret_cast<const unsigned char*>(net->tc_image); a.grads = net->grads;
    if (!batch->records) return fail(TSM_ERR_INVALID, "%s: engine 1 needs batch.records (tsm_ppo_pack_records)", __func__);
    a.records = reinterpret_cast<const unsigned char*>(batch->records); a.slots = ppo::slots_of(batch);
    a.view = ppo::make_batch_view(batch->indices, batch->size, batch->n_rows, batch->start, batch->len, batch->perm_key, batch->perm_n);
    a.mb_start = mb_start; a.mb_len = mb_len; a.mb_stats = mb_stats;
    a.cfg.eps_clip = cfg->eps_clip; a.cfg.dual_clip = cfg->dual_clip; a.cfg.vf_coef = cfg->vf_coef; a.cfg.ent_coef = cfg->ent_coef;
    a.cfg.value_clip = cfg->value_clip; a.cfg.adv_norm = cfg->adv_norm;
    a.cfg.plain_pg = cfg->plain_pg; a.cfg.value_heads = cfg->value_heads; a.cfg.probs_param = net->probs_param;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim; a.n_act = net->n_act; a.value_index = net->value_index;
    float sc = 1.0f;  // S = 2^ceil(log2(mb_len)), capped, so that S / mb_len is in (0.5, 1]
    while (sc < (float)mb_len && sc < 1048576.0f) sc *= 2.0f;
    a.gscale = sc; a.inv_gscale = 1.0f / sc;
    a.loss_sums = loss_sums;
    static const int dbg = getenv("TSM_TC_DEBUG") ? atoi(getenv("TSM_TC_DEBUG")) : 0;
    a.debug = dbg;
    const size_t smem = mlptc::SM_TOTAL;  // ~113 KB: two CTAs per SM (2 x 256 TMEM columns); 1 KB per CTA is reserved by the driver
    static_assert(2 * (mlptc::SM_TOTAL + 1024) <= 228 * 1024 && 3 * (mlptc::SM_TOTAL + 1024) > 228 * 1024,
                  "smem map must give exactly two CTAs per SM");
    static bool attr = false;
    if (!attr) {
        TSM_CUDA(cudaFuncSetAttribute(mlptc::tc_ppo_minibatch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    const int64_t tiles = ceil_div(mb_len, mlptc::TC_ROWS);
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)sm_count() * 2);
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(grid); lc.blockDim = dim3(mlptc::TC_THREADS); lc.dynamicSmemBytes = smem; lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl ? 1 : 0;
    CUtensorMap rec_map;
    rc = record_map(batch->records, batch->size * batch->n_rows, &rec_map, __func__);
    if (rc) return rc;
    TSM_CUDA(cudaLaunchKernelEx(&lc, mlptc::tc_ppo_minibatch_kernel, a, rec_map));
#ifdef TSM_TC_TRACE
    if (getenv("TSM_TC_TRACE_DUMP")) {
        long long h[4][64];
        cudaDeviceSynchronize();
        cudaMemcpyFromSymbol(h, mlptc::g_trace, sizeof(h));
        for (int row = 0; row < 4; ++row) {
            printf("trace row %d:", row);
            for (int k = 1; k < 24; ++k) printf(" %lld", h[row][k] - h[row][k - 1]);
            printf("  total %lld\n", h[row][23] - h[row][0]);
        }
        for (int row = 0; row < 4; row += 2)
            printf("kernel row %d: begin->setupA %lld ->prefetch issued %lld ->setup done %lld | tiles %lld | staging %lld reduce+sync %lld dealloc %lld atomics %lld | total %lld\n",
                   row, h[row][31] - h[row][30], h[row][32] - h[row][31], h[row][33] - h[row][32], h[row][34] - h[row][33], h[row][35] - h[row][34],
                   h[row][36] - h[row][35], h[row][37] - h[row][36], h[row][38] - h[row][37], h[row][38] - h[row][30]);
    }
#endif
    return TSM_OK;
}


// One cooperative launch for a whole epoch of uniform minibatches; returns 1 (not an error) when the launch cannot be used
// (no cooperative launch support, grid not co-resident) so that the caller falls back to one launch pair per minibatch.
int tc_ppo_epoch_persistent(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t start, int64_t mb_size,
                            int64_t n_mb, int64_t total, const float* mb_stats, double* loss_sums, const opt::AdamArgs& adam,
                            cudaStream_t s) {
    int rc = tc_check_net(net, __func__);
    if (rc) return rc;
    if (!batch->records) return fail(TSM_ERR_INVALID, "%s: engine 1 needs batch.records (tsm_ppo_pack_records)", __func__);
    if (reinterpret_cast<uintptr_t>(net->grads) & 15) return fail(TSM_ERR_INVALID, "%s: grads must be 16-byte aligned", __func__);
    if (adam.n > (int64_t)mlptc::EP_G_STRIDE) return 1;
    if (mb_size >= (int64_t)1 << 30 || total - (n_mb - 1) * mb_size >= (int64_t)1 << 30) return 1;  // in-minibatch row numbers are 32-bit in the kernel
    static int coop = -1, occ = 0, hv = mlptc::EP_HV, groups = 4;
    static size_t smem = mlptc::SM_EP_TOTAL;
    if (coop < 0) {
        int dev = 0;
        TSM_CUDA(cudaGetDevice(&dev));
        TSM_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        if (const char* h = getenv("TSM_TC_EPOCH_HALVES")) hv = atoi(h) == 1 ? 1 : 2;  // 1: one thread per tile row (round-1 shape)
        int occ1 = 0, occ2 = 0;
        TSM_CUDA(cudaFuncSetAttribute(mlptc::tc_ppo_epoch_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        TSM_CUDA(cudaFuncSetAttribute(mlptc::tc_ppo_epoch_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        TSM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, mlptc::tc_ppo_epoch_kernel<1>, mlptc::EP_THREADS, smem));
        TSM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, mlptc::tc_ppo_epoch_kernel<2>, 2 * mlptc::EP_THREADS, smem));
        occ = hv == 1 ? occ1 : occ2;
        // four tile groups per SM (mlp_tc_epoch4.cuh) unless TSM_TC_EPOCH_GROUPS=2 or TSM_TC_EPOCH_HALVES asks for the two-group kernels
        if (const char* g = getenv("TSM_TC_EPOCH_GROUPS")) groups = atoi(g) == 2 ? 2 : 4;
        else if (getenv("TSM_TC_EPOCH_HALVES")) groups = 2;
        if (groups == 4) {
            int occ4 = 0;
            TSM_CUDA(cudaFuncSetAttribute(mlptc::tc_ppo_epoch4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mlptc::G4_TOTAL));
            TSM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ4, mlptc::tc_ppo_epoch4_kernel, mlptc::G4_THREADS, (size_t)mlptc::G4_TOTAL));
            if (occ4 >= 1) { occ = occ4; smem = mlptc::G4_TOTAL; } else groups = 2;
        }
        if (occ > 1) occ = 1;  // one CTA owns the SM's 512 TMEM columns
    }
    void* const kernel = groups == 4 ? reinterpret_cast<void*>(mlptc::tc_ppo_epoch4_kernel)
                         : hv == 1   ? reinterpret_cast<void*>(mlptc::tc_ppo_epoch_kernel<1>)
                                     : reinterpret_cast<void*>(mlptc::tc_ppo_epoch_kernel<2>);
    const int cta_threads = groups == 4 ? mlptc::G4_THREADS : hv * mlptc::EP_THREADS;
    static const bool verbose = getenv("TSM_TC_EPOCH_VERBOSE") != nullptr;
    if (verbose) fprintf(stderr, "tsm_ppo_epoch: %d tile groups per SM, %d threads per CTA, %zu bytes of shared memory, occupancy %d\n", groups, cta_threads, smem, occ);
    if (!coop || occ < 1) return 1;
    mlptc::TcEpochArgs e{};
    mlptc::TcPpoArgs& a = e.p;
    a.image = reinterpret_cast<const unsigned char*>(net->tc_image); a.grads = net->grads;
    a.records = reinterpret_cast<const unsigned char*>(batch->records); a.slots = ppo::slots_of(batch);
    a.view = ppo::make_batch_view(batch->indices, batch->size, batch->n_rows, batch->start, batch->len, batch->perm_key, batch->perm_n);
    a.mb_start = start; a.mb_len = mb_size; a.mb_stats = nullptr;
    a.cfg.eps_clip = cfg->eps_clip; a.cfg.dual_clip = cfg->dual_clip; a.cfg.vf_coef = cfg->vf_coef; a.cfg.ent_coef = cfg->ent_coef;
    a.cfg.value_clip = cfg->value_clip; a.cfg.adv_norm = cfg->adv_norm;
    a.cfg.plain_pg = cfg->plain_pg; a.cfg.value_heads = cfg->value_heads; a.cfg.probs_param = net->probs_param;
    a.in_dim = net->in_dim; a.out_dim = net->out_dim; a.n_act = net->n_act; a.value_index = net->value_index;
    a.gscale = 1.f; a.inv_gscale = 1.f; a.loss_sums = nullptr;
    a.debug = getenv("TSM_TC_DEBUG") ? atoi(getenv("TSM_TC_DEBUG")) : 0;
    // workspace behind the image (TSM_TC_IMAGE_BYTES): two more gradient buffers
    static_assert(TSM_TC_IMAGE_BYTES >= opt::IMG_BYTES + 2 * mlptc::EP_G_STRIDE * 4 + 256 + mlptc::EP_Q_WORDS * 8 + mlptc::EP_G_STRIDE * 4,
                  "TSM_TC_IMAGE_BYTES too small for the epoch kernel's workspace");
    e.grads1 = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(net->tc_image) + opt::IMG_BYTES);
    e.grads2 = e.grads1 + mlptc::EP_G_STRIDE;
    e.sync_words = reinterpret_cast<uint32_t*>(e.grads2 + mlptc::EP_G_STRIDE);
    e.q_table = reinterpret_cast<uint2*>(e.sync_words + 64);
    e.g_total = reinterpret_cast<float*>(e.q_table + mlptc::EP_Q_WORDS);
    e.n_mb = n_mb; e.mb_size = mb_size; e.total = total; e.mb_stats = mb_stats; e.loss_sums = loss_sums; e.adam = adam;
    const int64_t tiles = ceil_div(n_mb > 1 ? mb_size : total, mlptc::TC_ROWS);
    // (four groups: one CTA per tile until the SMs run out, so that small minibatches spread over the SMs)
    unsigned grid = (unsigned)std::min<int64_t>(groups == 4 ? tiles : ceil_div(tiles, (int64_t)groups), (int64_t)sm_count() * occ);
    if (const char* gmax = getenv("TSM_TC_EPOCH_GRID")) grid = std::max(1u, std::min(grid, (unsigned)atoi(gmax)));  // experiments: fewer CTAs
    CUtensorMap rec_map;
    rc = record_map(batch->records, batch->size * batch->n_rows, &rec_map, __func__);
    if (rc) return rc;
    void* args[] = {&e, &rec_map};
    const cudaError_t err = cudaLaunchCooperativeKernel(kernel, dim3(grid), dim3(cta_threads), args, smem, s);
    if (err == cudaErrorCooperativeLaunchTooLarge) {
        cudaGetLastError();
        return 1;
    }
    TSM_CUDA(err);
#ifdef TSM_TC_TRACE
    if (getenv("TSM_TC_TRACE_DUMP")) {
        long long h[4][64];
        cudaDeviceSynchronize();
        cudaMemcpyFromSymbol(h, mlptc::g_trace, sizeof(h));
        for (int row = 0; row < 4; ++row) {
            printf("epoch tile row %d:", row);
            for (int k = 1; k < 24; ++k) printf(" %lld", h[row][k] - h[row][k - 1]);
            printf("  total %lld\n", h[row][23] - h[row][0]);
        }
        {
            cudaFuncAttributes fa;
            cudaFuncGetAttributes(&fa, kernel);
            printf("grid %u x %d threads, occ %d, smem %zu, regs %d\n", grid, cta_threads, occ, smem, fa.numRegs);
        }
        printf("half-1 prefetch (tile 0): pos/slot %lld gather issue %lld advance %lld slot load issue %lld | start at +%lld of the tile\n", h[3][51] - h[3][50],
               h[3][52] - h[3][51], h[3][53] - h[3][52], h[3][54] - h[3][53], h[3][50] - h[0][0]);
        for (int row = 0; row < 3; ++row)
            printf("tile row %d: store_rec %lld prefetch-issue %lld to-forward %lld\n", row, h[row][24] - h[row][0], h[row][25] - h[row][24],
                   h[row][1] - h[row][25]);
        printf("abs offsets from slot40 (CTA0): tile0 start %lld end %lld | tile1 start %lld end %lld | loop end %lld\n", h[0][0] - h[0][40],
               h[0][23] - h[0][40], h[1][0] - h[0][40], h[1][23] - h[0][40], h[0][41] - h[0][40]);
        if (groups == 4) {
            unsigned long long hc[160][8];
            cudaMemcpyFromSymbol(hc, mlptc::g_trace_cta, sizeof(hc));
            unsigned long long t0 = ~0ull;
            for (unsigned c = 0; c < grid; ++c) t0 = hc[c][0] < t0 ? hc[c][0] : t0;
            printf("minibatch 2, ns after the first CTA's start: CTA: start tiles-done flush-done arrive passed end\n");
            for (unsigned c = 0; c < grid; c += (grid > 24 ? grid / 24 : 1))
                printf("  CTA %3u: %5llu %5llu %5llu %5llu %5llu %5llu\n", c, hc[c][0] - t0, hc[c][1] - t0, hc[c][2] - t0, hc[c][3] - t0, hc[c][4] - t0, hc[c][5] - t0);
            unsigned long long mx[6] = {0, 0, 0, 0, 0, 0};
            for (unsigned c = 0; c < grid; ++c) for (int k = 0; k < 6; ++k) mx[k] = hc[c][k] - t0 > mx[k] ? hc[c][k] - t0 : mx[k];
            printf("  max    : %5llu %5llu %5llu %5llu %5llu %5llu\n", mx[0], mx[1], mx[2], mx[3], mx[4], mx[5]);
            unsigned long long hg[4][8];
            cudaMemcpyFromSymbol(hg, mlptc::g_trace_grp, sizeof(hg));
            printf("CTA 0, groups (ns after the first CTA's start): tile start, records taken, last product issued, products done, all groups done, staged, flushed\n");
            for (int g = 0; g < 4; ++g)
                printf("  group %d: %5llu %5llu %5llu %5llu %5llu %5llu %5llu\n", g, hg[g][0] - t0, hg[g][1] - t0, hg[g][2] - t0, hg[g][3] - t0, hg[g][4] - t0,
                       hg[g][5] - t0, hg[g][6] - t0);
            printf("CTA 0 group 0, tile 0 (cycles): tile start -> records landed %lld -> taken %lld -> bookkeeping %lld\n", h[0][26] - h[0][0], h[0][27] - h[0][26], h[0][1] - h[0][27]);
            printf("CTA 0 group 0 (cycles): last product issued -> flush prologue done %lld -> products done %lld\n", h[0][49] - h[0][41], h[0][50] - h[0][49]);
            printf("four-group kernel, CTA 0, cycles per minibatch:");
            for (int k = 1; k < 16 && h[1][k] > 0; ++k) printf(" %lld", h[1][k] - h[1][k - 1]);
            printf("\n");
        }
        for (int row = 0; row < 4; row += 2)
            printf("epoch row %d (minibatch 2): tiles %lld flush %lld grid-barrier %lld rank-exchange %lld norm %lld adam %lld clear+publish %lld | total %lld\n",
                   row, h[row][41] - h[row][40], h[row][42] - h[row][41], h[row][43] - h[row][42], h[row][48] - h[row][43], h[row][44] - h[row][48],
                   h[row][45] - h[row][44], h[row][47] - h[row][45], h[row][47] - h[row][40]);
    }
#endif
    return TSM_OK;
}

}  // namespace tsm
