// buffer.cu -- K2: device-resident VectorReplayBuffer (sm_100a).
//
// Replaces the per-env Python/numba state machine of
//   tianshou/data/buffer/buffer_base.py:292-298 (reset), :309-312 (unfinished_index),
//     :355-410 (_update_state_pre_add), :495-531 (sample_indices), :591-635 (__getitem__)
//   tianshou/data/buffer/manager.py:70-75, :85-91, :131-193 (add), :195-229, :306-358
// with one thread per row / per index.  Counters are SoA int64 arrays of length n_bufs;
// payload is time-major ([slot][row]) so a lock-step vector step is one contiguous write.
// Every index that leaves this file is numbered env-major (row * size + slot) like the
// reference.
#include "common.cuh"

namespace tsm {
namespace buf {

constexpr int THREADS = 256;

__global__ void reset_kernel(tsm_buf b, int keep_statistics) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= b.n_bufs) return;
    b.ins[r] = 0;
    b.sz[r] = 0;
    b.ep_start[r] = 0;
    b.last_index[r] = r * b.size;  // manager.py:72
    if (!keep_statistics) {
        b.ep_ret[r] = 0.0;
        b.ep_len[r] = 0;
    }
}

// payload scatter; runs BEFORE the counter update and reads the pre-update insertion index
template <int V>
__global__ void add_obs_kernel(tsm_buf b, const float* __restrict__ obs, const float* __restrict__ obs_next,
                               const int64_t* __restrict__ ids, int64_t n) {
    const int64_t O = b.obs_dim, OV = O / V;
    const int64_t total = n * OV;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = e / OV, d = (e - k * OV) * V;
        const int64_t row = ids ? ids[k] : k;
        const int64_t dst = (b.ins[row] * b.n_bufs + row) * O + d;
        const int64_t src = k * O + d;
        if constexpr (V == 4) {
            if (obs) *reinterpret_cast<float4*>(b.obs + dst) = *reinterpret_cast<const float4*>(obs + src);
            if (obs_next) *reinterpret_cast<float4*>(b.obs_next + dst) = *reinterpret_cast<const float4*>(obs_next + src);
        } else if constexpr (V == 2) {
            if (obs) *reinterpret_cast<float2*>(b.obs + dst) = *reinterpret_cast<const float2*>(obs + src);
            if (obs_next) *reinterpret_cast<float2*>(b.obs_next + dst) = *reinterpret_cast<const float2*>(obs_next + src);
        } else {
            if (obs) b.obs[dst] = obs[src];
            if (obs_next) b.obs_next[dst] = obs_next[src];
        }
    }
}

// counters + scalar payload: buffer_base.py:373-410 and manager.py:162-177, one thread per row
__global__ void add_state_kernel(tsm_buf b, const int32_t* __restrict__ act, const float* __restrict__ rew,
                                 const uint8_t* __restrict__ term, const uint8_t* __restrict__ trunc,
                                 const float* __restrict__ logp, const float* __restrict__ value,
                                 const int64_t* __restrict__ ids, int64_t n, int64_t* __restrict__ idx_out,
                                 double* __restrict__ ep_ret_out, int64_t* __restrict__ ep_len_out,
                                 int64_t* __restrict__ ep_start_out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int64_t row = ids ? ids[k] : k;
    const int64_t off = row * b.size;
    const int64_t cur = b.ins[row];
    const int64_t nsz = min(b.sz[row] + 1, b.size);
    const int64_t nins = (cur + 1) % b.size;
    const uint8_t te = term[k] ? 1 : 0, tr = trunc[k] ? 1 : 0, dn = te | tr;
    const float r = rew[k];
    const double eret = b.ep_ret[row] + (double)r;
    const int64_t elen = b.ep_len[row] + 1;
    const int64_t estart = b.ep_start[row];
    // payload scalars, time-major slot
    const int64_t slot = cur * b.n_bufs + row;
    if (act) b.act[slot] = act[k];
    b.rew[slot] = r;
    b.terminated[slot] = te;
    b.truncated[slot] = tr;
    b.done[slot] = dn;
    if (logp && b.logp) b.logp[slot] = logp[k];
    if (value && b.value) b.value[slot] = value[k];
    // results (0 for unfinished episodes, buffer_base.py:389-403)
    if (idx_out) idx_out[k] = cur + off;
    if (ep_ret_out) ep_ret_out[k] = dn ? eret : 0.0;
    if (ep_len_out) ep_len_out[k] = dn ? elen : 0;
    if (ep_start_out) ep_start_out[k] = estart + off;
    // state
    b.ins[row] = nins;
    b.sz[row] = nsz;
    b.last_index[row] = cur + off;
    b.ep_ret[row] = dn ? 0.0 : eret;
    b.ep_len[row] = dn ? 0 : elen;
    b.ep_start[row] = dn ? nins : estart;  // buffer_base.py:408
}

// manager.py:306-358, O(1) per index (the reference loops over all sub-buffers with masks)
template <bool NEXT>
__global__ void neighbour_kernel(tsm_buf b, const int64_t* __restrict__ idx, int64_t n, int64_t* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int64_t maxsize = b.n_bufs * b.size;
    int64_t i = idx[k] % maxsize;
    if (i < 0) i += maxsize;  // numpy modulo is non-negative
    const int64_t row = i / b.size, start = row * b.size;
    const int64_t cur_len = max((int64_t)1, b.sz[row]);
    const int64_t last = b.last_index[row];
    if (NEXT) {
        const int64_t t = i - start;
        const int end_flag = (b.done[t * b.n_bufs + row] != 0) | (i == last);
        int64_t v = (t + 1 - end_flag) % cur_len;
        if (v < 0) v += cur_len;
        out[k] = v + start;
    } else {
        int64_t sub = (i - start - 1) % cur_len;
        if (sub < 0) sub += cur_len;
        const int end_flag = (b.done[sub * b.n_bufs + row] != 0) | (sub + start == last);
        out[k] = (sub + end_flag) % cur_len + start;
    }
}


// ---------------------------------------------------------------------------------------------
// row-ordered compaction helpers (exclusive scan over rows in blocks of SCAN_ROWS)
constexpr int SCAN_ROWS = 1024;

__device__ __forceinline__ int64_t row_count(const tsm_buf& b, int64_t row, int mode) {
    if (row >= b.n_bufs) return 0;
    const int64_t sz = b.sz[row];
    if (mode == 0) return sz;  // sample_indices(0): every valid slot
    // unfinished_index, buffer_base.py:309-312: last = (ins - 1) % size_cur ; keep if !done[last]
    if (sz == 0) return 0;
    int64_t last = (b.ins[row] - 1) % sz;
    if (last < 0) last += sz;
    return b.done[last * b.n_bufs + row] ? 0 : 1;
}

// exclusive scan of one value per thread across a SCAN_ROWS-thread block; returns block total via *total
__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
    __shared__ int64_t s_w[SCAN_ROWS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int64_t w = s_w[lane];
        int64_t winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int64_t o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        s_w[lane] = winc - w;  // exclusive warp offsets
        if (lane == 31) *total = winc;
    }
    __syncthreads();
    return inc - v + s_w[warp];
}

__global__ void __launch_bounds__(SCAN_ROWS) block_sums_kernel(tsm_buf b, int mode, int64_t* __restrict__ bsum) {
    __shared__ int64_t s_total;
    const int64_t row = (int64_t)blockIdx.x * SCAN_ROWS + threadIdx.x;
    (void)block_exclusive_scan(row_count(b, row, mode), &s_total);
    __syncthreads();
    if (threadIdx.x == 0) bsum[blockIdx.x] = s_total;
}

// single block: exclusive scan of the per-block sums in place; total -> bsum[nblk] and *count_out
__global__ void __launch_bounds__(SCAN_ROWS) scan_block_sums_kernel(int64_t* __restrict__ bsum, int64_t nblk,
                                                                    int64_t* __restrict__ count_out) {
    __shared__ int64_t s_total;
    __shared__ int64_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < nblk; base += SCAN_ROWS) {
        const int64_t i = base + threadIdx.x;
        const int64_t v = i < nblk ? bsum[i] : 0;
        const int64_t ex = block_exclusive_scan(v, &s_total);
        __syncthreads();
        if (i < nblk) bsum[i] = ex + s_carry;
        __syncthreads();
        if (threadIdx.x == 0) s_carry += s_total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        bsum[nblk] = s_carry;
        if (count_out) *count_out = s_carry;
    }
}

// sample_indices(0): row by row, each row in time order [ins..sz) ++ [0..ins)  (buffer_base.py:511-514)
__global__ void __launch_bounds__(SCAN_ROWS) indices_all_kernel(tsm_buf b, const int64_t* __restrict__ bsum,
                                                                int64_t* __restrict__ out) {
    __shared__ int64_t s_total;
    __shared__ int64_t s_off[SCAN_ROWS];
    const int64_t row0 = (int64_t)blockIdx.x * SCAN_ROWS;
    const int64_t row = row0 + threadIdx.x;
    const int64_t cnt = row_count(b, row, 0);
    s_off[threadIdx.x] = block_exclusive_scan(cnt, &s_total) + bsum[blockIdx.x];
    __syncthreads();
    // each warp writes the runs of its 32 rows cooperatively so stores are contiguous
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int q = 0; q < 32; ++q) {
        const int64_t rr = row0 + warp * 32 + q;
        if (rr >= b.n_bufs) break;
        const int64_t sz = b.sz[rr], ins = b.ins[rr];
        const int64_t first = max(sz - ins, (int64_t)0);  // len(arange(ins, sz))
        const int64_t o = s_off[warp * 32 + q], base = rr * b.size;
        for (int64_t j = lane; j < sz; j += 32) out[o + j] = base + (j < first ? ins + j : j - first);
    }
}

__global__ void __launch_bounds__(SCAN_ROWS) unfinished_kernel(tsm_buf b, const int64_t* __restrict__ bsum,
                                                               int64_t* __restrict__ out) {
    __shared__ int64_t s_total;
    const int64_t row = (int64_t)blockIdx.x * SCAN_ROWS + threadIdx.x;
    const int64_t cnt = row_count(b, row, 1);
    const int64_t o = block_exclusive_scan(cnt, &s_total) + bsum[blockIdx.x];
    if (cnt) {
        const int64_t sz = b.sz[row];
        int64_t last = (b.ins[row] - 1) % sz;
        if (last < 0) last += sz;
        out[o] = last + row * b.size;
    }
}

// ---------------------------------------------------------------------------------------------
// gather by env-major index (buffer_base.py:591-635), payload time-major
template <int V>
__global__ void gather_obs_kernel(tsm_buf b, const int64_t* __restrict__ idx, int64_t n, float* __restrict__ obs_out,
                                  float* __restrict__ obs_next_out) {
    const int64_t O = b.obs_dim, OV = O / V, total = n * OV;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = e / OV, d = (e - k * OV) * V;
        const int64_t i = idx[k];
        const int64_t row = i / b.size, t = i - row * b.size;
        const int64_t src = (t * b.n_bufs + row) * O + d, dst = k * O + d;
        if constexpr (V == 4) {
            if (obs_out) *reinterpret_cast<float4*>(obs_out + dst) = *reinterpret_cast<const float4*>(b.obs + src);
            if (obs_next_out) *reinterpret_cast<float4*>(obs_next_out + dst) = *reinterpret_cast<const float4*>(b.obs_next + src);
        } else if constexpr (V == 2) {
            if (obs_out) *reinterpret_cast<float2*>(obs_out + dst) = *reinterpret_cast<const float2*>(b.obs + src);
            if (obs_next_out) *reinterpret_cast<float2*>(obs_next_out + dst) = *reinterpret_cast<const float2*>(b.obs_next + src);
        } else {
            if (obs_out) obs_out[dst] = b.obs[src];
            if (obs_next_out) obs_next_out[dst] = b.obs_next[src];
        }
    }
}

constexpr int MAX_EXTRA = 8;
struct ExtraPtrs {
    const float* in[MAX_EXTRA];
    float* out[MAX_EXTRA];
    int n;
};

__global__ void gather_scalar_kernel(tsm_buf b, const int64_t* __restrict__ idx, int64_t n, int32_t* __restrict__ act_out,
                                     float* __restrict__ rew_out, uint8_t* __restrict__ term_out,
                                     uint8_t* __restrict__ trunc_out, uint8_t* __restrict__ done_out,
                                     float* __restrict__ logp_out, float* __restrict__ value_out, ExtraPtrs ex) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int64_t i = idx[k];
    const int64_t row = i / b.size, t = i - row * b.size;
    const int64_t s = t * b.n_bufs + row;
    if (act_out) act_out[k] = b.act[s];
    if (rew_out) rew_out[k] = b.rew[s];
    if (term_out) term_out[k] = b.terminated[s];
    if (trunc_out) trunc_out[k] = b.truncated[s];
    if (done_out) done_out[k] = b.done[s];
    if (logp_out) logp_out[k] = b.logp[s];
    if (value_out) value_out[k] = b.value[s];
#pragma unroll
    for (int q = 0; q < MAX_EXTRA; ++q)
        if (q < ex.n) ex.out[q][k] = ex.in[q][s];
}

// ---------------------------------------------------------------------------------------------
// episode statistics of one vector step: rows whose ep_len > 0 finished an episode
// (collector.py:983-1024 loops over done envs on the host; CollectStats :116-254).
// acc[0..8] = {count, sum_ret, sum_ret^2, min_ret, max_ret, sum_len, sum_len^2, min_len, max_len}
__device__ __forceinline__ void atomic_min_d(double* a, double v) {
    unsigned long long* p = reinterpret_cast<unsigned long long*>(a);
    unsigned long long old = *p;
    while (__longlong_as_double((long long)old) > v) {
        const unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v));
        if (prev == old) break;
        old = prev;
    }
}
__device__ __forceinline__ void atomic_max_d(double* a, double v) {
    unsigned long long* p = reinterpret_cast<unsigned long long*>(a);
    unsigned long long old = *p;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v));
        if (prev == old) break;
        old = prev;
    }
}

__global__ void __launch_bounds__(THREADS) episode_stats_kernel(const double* __restrict__ ep_ret, const int64_t* __restrict__ ep_len,
                                                               int64_t n, double* __restrict__ acc, double* __restrict__ list_ret,
                                                               int64_t* __restrict__ list_len, int64_t* __restrict__ list_count,
                                                               int64_t cap) {
    double c = 0, s = 0, s2 = 0, mn = INFINITY, mx = -INFINITY, l = 0, l2 = 0, lmn = INFINITY, lmx = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * THREADS) {
        const int64_t len = ep_len[i];
        if (len > 0) {
            const double r = ep_ret[i], dl = (double)len;
            c += 1; s += r; s2 += r * r; mn = fmin(mn, r); mx = fmax(mx, r);
            l += dl; l2 += dl * dl; lmn = fmin(lmn, dl); lmx = fmax(lmx, dl);
            if (list_count) {
                const int64_t at = (int64_t)atomicAdd(reinterpret_cast<unsigned long long*>(list_count), 1ull);
                if (at < cap) { list_ret[at] = r; list_len[at] = len; }
            }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        c += __shfl_down_sync(0xffffffffu, c, d); s += __shfl_down_sync(0xffffffffu, s, d); s2 += __shfl_down_sync(0xffffffffu, s2, d);
        l += __shfl_down_sync(0xffffffffu, l, d); l2 += __shfl_down_sync(0xffffffffu, l2, d);
        mn = fmin(mn, __shfl_down_sync(0xffffffffu, mn, d)); mx = fmax(mx, __shfl_down_sync(0xffffffffu, mx, d));
        lmn = fmin(lmn, __shfl_down_sync(0xffffffffu, lmn, d)); lmx = fmax(lmx, __shfl_down_sync(0xffffffffu, lmx, d));
    }
    if ((threadIdx.x & 31) == 0 && c > 0) {
        atomicAdd(acc + 0, c); atomicAdd(acc + 1, s); atomicAdd(acc + 2, s2); atomic_min_d(acc + 3, mn); atomic_max_d(acc + 4, mx);
        atomicAdd(acc + 5, l); atomicAdd(acc + 6, l2); atomic_min_d(acc + 7, lmn); atomic_max_d(acc + 8, lmx);
    }
}

}  // namespace buf
}  // namespace tsm

using namespace tsm;
using namespace tsm::buf;

static int check_buf(const tsm_buf* b, const char* fn) {
    if (!b) return fail(TSM_ERR_INVALID, "%s: buf is NULL", fn);
    if (b->n_bufs <= 0 || b->size <= 0 || b->obs_dim < 0)
        return fail(TSM_ERR_INVALID, "%s: n_bufs, size must be > 0 and obs_dim >= 0", fn);
    if (!b->ins || !b->sz || !b->last_index || !b->ep_len || !b->ep_start || !b->ep_ret)
        return fail(TSM_ERR_INVALID, "%s: counter arrays must not be NULL", fn);
    return TSM_OK;
}

static inline unsigned grid_for(int64_t work, int threads) {
    int64_t g = ceil_div(work, threads);
    int64_t cap = (int64_t)sm_count() * 32;  // grid-stride kernels: a few waves of the 148 SMs
    return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

extern "C" int tsm_buf_reset(const tsm_buf* b, int keep_statistics, tsm_stream_t stream) {
    int rc = check_buf(b, __func__);
    if (rc) return rc;
    reset_kernel<<<(unsigned)ceil_div(b->n_bufs, THREADS), THREADS, 0, as_stream(stream)>>>(*b, keep_statistics);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_buf_add(const tsm_buf* b, const float* obs, const int32_t* act, const float* rew,
                           const uint8_t* terminated, const uint8_t* truncated, const float* obs_next,
                           const float* logp, const float* value, const int64_t* buffer_ids, int64_t n,
                           int64_t* idx_out, double* ep_ret_out, int64_t* ep_len_out, int64_t* ep_start_out,
                           tsm_stream_t stream) {
    int rc = check_buf(b, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n >= 0 && n <= b->n_bufs, "0 <= n <= n_bufs");
    TSM_REQUIRE(buffer_ids != nullptr || n == b->n_bufs, "buffer_ids == NULL needs n == n_bufs");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(rew && terminated && truncated, "rew/terminated/truncated != NULL");
    TSM_REQUIRE(b->rew && b->terminated && b->truncated && b->done, "buffer payload arrays != NULL");
    TSM_REQUIRE(!act || b->act, "buffer act array");
    TSM_REQUIRE(!obs || b->obs, "buffer obs array");
    TSM_REQUIRE(!obs_next || b->obs_next, "buffer obs_next array");
    cudaStream_t s = as_stream(stream);
    if ((obs || obs_next) && b->obs_dim > 0) {
        const int64_t O = b->obs_dim;
        auto al = [](const void* p, int a) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) % a) == 0; };
        if (O % 4 == 0 && al(obs, 16) && al(obs_next, 16) && al(b->obs, 16) && al(b->obs_next, 16))
            add_obs_kernel<4><<<grid_for(n * (O / 4), THREADS), THREADS, 0, s>>>(*b, obs, obs_next, buffer_ids, n);
        else if (O % 2 == 0 && al(obs, 8) && al(obs_next, 8) && al(b->obs, 8) && al(b->obs_next, 8))
            add_obs_kernel<2><<<grid_for(n * (O / 2), THREADS), THREADS, 0, s>>>(*b, obs, obs_next, buffer_ids, n);
        else
            add_obs_kernel<1><<<grid_for(n * O, THREADS), THREADS, 0, s>>>(*b, obs, obs_next, buffer_ids, n);
        TSM_LAUNCH_CHECK();
    }
    add_state_kernel<<<(unsigned)ceil_div(n, THREADS), THREADS, 0, s>>>(*b, act, rew, terminated, truncated, logp, value,
                                                                      buffer_ids, n, idx_out, ep_ret_out, ep_len_out,
                                                                      ep_start_out);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" size_t tsm_buf_indices_workspace_bytes(int64_t n_bufs) {
    if (n_bufs < 0) n_bufs = 0;
    return (size_t)(ceil_div(n_bufs, SCAN_ROWS) + 2) * sizeof(int64_t);
}

static int compact_rows(const tsm_buf* b, int mode, int64_t* out, int64_t* count_out, void* workspace,
                        size_t workspace_bytes, cudaStream_t s, const char* fn) {
    int rc = check_buf(b, fn);
    if (rc) return rc;
    if (!out) return fail(TSM_ERR_INVALID, "%s: out is NULL", fn);
    if (!workspace || workspace_bytes < tsm_buf_indices_workspace_bytes(b->n_bufs) ||
        (reinterpret_cast<uintptr_t>(workspace) % 8) != 0)
        return fail(TSM_ERR_WORKSPACE, "%s: workspace too small or misaligned (need %lld bytes)", fn,
                    (long long)tsm_buf_indices_workspace_bytes(b->n_bufs));
    if (!b->done) return fail(TSM_ERR_INVALID, "%s: buffer done array is NULL", fn);
    const int64_t nblk = ceil_div(b->n_bufs, SCAN_ROWS);
    int64_t* bsum = reinterpret_cast<int64_t*>(workspace);
    block_sums_kernel<<<(unsigned)nblk, SCAN_ROWS, 0, s>>>(*b, mode, bsum);
    scan_block_sums_kernel<<<1, SCAN_ROWS, 0, s>>>(bsum, nblk, count_out);
    if (mode == 0) indices_all_kernel<<<(unsigned)nblk, SCAN_ROWS, 0, s>>>(*b, bsum, out);
    else unfinished_kernel<<<(unsigned)nblk, SCAN_ROWS, 0, s>>>(*b, bsum, out);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_buf_indices_all(const tsm_buf* b, int64_t* out, int64_t* count_out, void* workspace,
                                   size_t workspace_bytes, tsm_stream_t stream) {
    return compact_rows(b, 0, out, count_out, workspace, workspace_bytes, as_stream(stream), __func__);
}

extern "C" int tsm_buf_unfinished(const tsm_buf* b, int64_t* out, int64_t* count_out, void* workspace,
                                  size_t workspace_bytes, tsm_stream_t stream) {
    return compact_rows(b, 1, out, count_out, workspace, workspace_bytes, as_stream(stream), __func__);
}

static int neighbour(const tsm_buf* b, const int64_t* idx, int64_t n, int64_t* out, cudaStream_t s, bool next,
                     const char* fn) {
    int rc = check_buf(b, fn);
    if (rc) return rc;
    if (n < 0 || (n > 0 && (!idx || !out))) return fail(TSM_ERR_INVALID, "%s: idx/out NULL or n < 0", fn);
    if (n == 0) return TSM_OK;
    if (!b->done) return fail(TSM_ERR_INVALID, "%s: buffer done array is NULL", fn);
    const unsigned g = (unsigned)ceil_div(n, THREADS);
    if (next) neighbour_kernel<true><<<g, THREADS, 0, s>>>(*b, idx, n, out);
    else neighbour_kernel<false><<<g, THREADS, 0, s>>>(*b, idx, n, out);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_buf_prev(const tsm_buf* b, const int64_t* idx, int64_t n, int64_t* out, tsm_stream_t stream) {
    return neighbour(b, idx, n, out, as_stream(stream), false, __func__);
}
extern "C" int tsm_buf_next(const tsm_buf* b, const int64_t* idx, int64_t n, int64_t* out, tsm_stream_t stream) {
    return neighbour(b, idx, n, out, as_stream(stream), true, __func__);
}

extern "C" int tsm_buf_gather(const tsm_buf* b, const int64_t* idx, int64_t n, float* obs_out, float* obs_next_out,
                              int32_t* act_out, float* rew_out, uint8_t* term_out, uint8_t* trunc_out,
                              uint8_t* done_out, float* logp_out, float* value_out, const float* const* h_extra_in,
                              float* const* h_extra_out, int n_extra, tsm_stream_t stream) {
    int rc = check_buf(b, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n >= 0, "n >= 0");
    TSM_REQUIRE(n_extra >= 0 && n_extra <= MAX_EXTRA, "0 <= n_extra <= 8");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(idx != nullptr, "idx != NULL");
    TSM_REQUIRE(!logp_out || b->logp, "buffer has no logp array");
    TSM_REQUIRE(!value_out || b->value, "buffer has no value array");
    cudaStream_t s = as_stream(stream);
    if ((obs_out || obs_next_out) && b->obs_dim > 0) {
        TSM_REQUIRE(!obs_out || b->obs, "buffer obs array");
        TSM_REQUIRE(!obs_next_out || b->obs_next, "buffer obs_next array");
        const int64_t O = b->obs_dim;
        auto al = [](const void* p, int a) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) % a) == 0; };
        if (O % 4 == 0 && al(obs_out, 16) && al(obs_next_out, 16) && al(b->obs, 16) && al(b->obs_next, 16))
            gather_obs_kernel<4><<<grid_for(n * (O / 4), THREADS), THREADS, 0, s>>>(*b, idx, n, obs_out, obs_next_out);
        else if (O % 2 == 0 && al(obs_out, 8) && al(obs_next_out, 8) && al(b->obs, 8) && al(b->obs_next, 8))
            gather_obs_kernel<2><<<grid_for(n * (O / 2), THREADS), THREADS, 0, s>>>(*b, idx, n, obs_out, obs_next_out);
        else
            gather_obs_kernel<1><<<grid_for(n * O, THREADS), THREADS, 0, s>>>(*b, idx, n, obs_out, obs_next_out);
        TSM_LAUNCH_CHECK();
    }
    ExtraPtrs ex{};
    ex.n = n_extra;
    for (int q = 0; q < n_extra; ++q) {
        TSM_REQUIRE(h_extra_in && h_extra_out && h_extra_in[q] && h_extra_out[q], "extra pointers != NULL");
        ex.in[q] = h_extra_in[q];
        ex.out[q] = h_extra_out[q];
    }
    if (act_out || rew_out || term_out || trunc_out || done_out || logp_out || value_out || n_extra) {
        gather_scalar_kernel<<<(unsigned)ceil_div(n, THREADS), THREADS, 0, s>>>(*b, idx, n, act_out, rew_out, term_out,
                                                                              trunc_out, done_out, logp_out, value_out, ex);
        TSM_LAUNCH_CHECK();
    }
    return TSM_OK;
}

extern "C" int tsm_episode_stats(const double* ep_ret, const int64_t* ep_len, int64_t n, double* acc, double* list_ret,
                                 int64_t* list_len, int64_t* list_count, int64_t list_capacity, tsm_stream_t stream) {
    TSM_REQUIRE(n >= 0, "n >= 0");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(ep_ret && ep_len && acc, "ep_ret / ep_len / acc != NULL");
    TSM_REQUIRE(!list_count || (list_ret && list_len && list_capacity >= 0), "episode list arrays");
    episode_stats_kernel<<<grid_for(n, THREADS), THREADS, 0, as_stream(stream)>>>(ep_ret, ep_len, n, acc, list_ret, list_len,
                                                                                 list_count, list_capacity);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
