// gae.cu -- K3: GAE / discounted returns as a segmented reverse scan (sm_100a).
//
// Replaces tianshou/algorithm/algorithm_base.py:651-717 (compute_episodic_return),
// :631-649 (value_mask), :1079-1134 (_gae), :1137-1151 (episode_mc_return_to_go).
//
// Two layouts:
//  * tsm_gae_flat : the reference's flat 1-D batch order.  One pass over HBM: every CTA scans
//    a tile of TILE elements (blocked, ITEMS contiguous elements per thread), composes the
//    affine maps g -> delta + c*g with warp shuffles, and obtains its carry from the tile to
//    its right by decoupled look-back.  Because c == 0 at every episode end, a tile that
//    contains any end flag publishes its inclusive value immediately, so the look-back depth
//    is 1 for MPE-like data and the scan never serialises.  Tile states are self-validating values (no flag, fence
//    or atomic between publisher and reader).
//  * tsm_gae_tm   : the device buffer's time-major [t][row] layout.  One thread per row walks
//    t backwards exactly like the reference's serial loop; rows are adjacent across lanes so
//    every load/store is a fully coalesced 128-byte line.
// Both carry the recurrence in float64 (numba computes in float64) and round once on store.
#include <algorithm>
#include "common.cuh"

namespace tsm {
namespace gae {

// 128 threads x 8 elements per tile and ten CTAs per SM (48 registers, 48 bytes of spills): the kernel is one dependent chain
// per tile (loads -> scan -> neighbour state -> stores), so what matters is how many tiles an SM holds.  Measured at 134 M
// elements: 256 threads / 3 CTAs per SM 4.07 TB/s, 256 / 4 4.74, 128 / 8 4.92, 128 / 10 5.08 (0.78 of the measured HBM peak),
// 128 / 12 3.99 (spills), 64 / 20 4.87.
#ifndef TSM_GAE_THREADS
#define TSM_GAE_THREADS 128
#endif
#ifndef TSM_GAE_ITEMS
#define TSM_GAE_ITEMS 8
#endif
constexpr int THREADS = TSM_GAE_THREADS;
constexpr int ITEMS = TSM_GAE_ITEMS;  // contiguous elements per thread
constexpr int TILE = THREADS * ITEMS;
constexpr int WARPS = THREADS / 32;

// Tile state for decoupled look-back.  Every field is SELF-VALIDATING: the workspace is filled with 0xFF bytes (an
// all-ones NaN no arithmetic produces) and a field is valid as soon as it holds anything else, so publishing is a plain
// 8-byte store and reading is a plain 8-byte load -- no status word, no fence, no atomic on the tile's critical path
// (the status + __threadfence + atomicExch version spent ~1.5 us per tile in dependent L2 round trips).
struct __align__(32) TileState {
    double a;     // aggregate multiplier
    double b;     // aggregate offset
    double incl;  // g at the tile's first element
    double pad;
};
constexpr unsigned long long TS_EMPTY = 0xffffffffffffffffull;
__device__ __forceinline__ void ts_publish(double* p, double v) {
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    if (u == TS_EMPTY) u = 0x7ff8000000000000ull;  // (a NaN input with this payload: publish the canonical NaN instead)
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(u) : "memory");
}
__device__ __forceinline__ bool ts_read(const double* p, double& v) {
    unsigned long long u;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(u) : "l"(p) : "memory");
    v = __longlong_as_double((long long)u);
    return u != TS_EMPTY;
}

struct Map {
    double a, b;  // g_out = b + a * g_in
};
__device__ __forceinline__ Map compose(const Map& l, const Map& r) {
    // apply r first (it is to the right), then l
    return Map{l.a * r.a, fma(l.a, r.b, l.b)};
}

template <typename T>
__device__ __forceinline__ void load_items(const T* __restrict__ p, int64_t base, int64_t n, T (&out)[ITEMS], T fill) {
    // base is a multiple of ITEMS; fast path when the whole chunk is in range and aligned
    constexpr int BYTES = sizeof(T) * ITEMS;
    if (p == nullptr) {
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) out[e] = fill;
        return;
    }
    if (base + ITEMS <= n && (reinterpret_cast<uintptr_t>(p + base) % (BYTES >= 16 ? 16 : BYTES)) == 0) {
        if constexpr (BYTES >= 16) {
            const uint4* q = reinterpret_cast<const uint4*>(p + base);
            uint4* o = reinterpret_cast<uint4*>(out);
#pragma unroll
            for (int k = 0; k < BYTES / 16; ++k) o[k] = q[k];
        } else if constexpr (BYTES == 8) {  // uint8 flags, ITEMS == 8
            *reinterpret_cast<uint2*>(out) = *reinterpret_cast<const uint2*>(p + base);
        } else {  // 4 bytes
            *reinterpret_cast<uint32_t*>(out) = *reinterpret_cast<const uint32_t*>(p + base);
        }
    } else {
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) out[e] = (base + e < n) ? p[base + e] : fill;
    }
}

template <typename T>
__device__ __forceinline__ void store_items(T* __restrict__ p, int64_t base, int64_t n, const T (&v)[ITEMS]) {
    constexpr int BYTES = sizeof(T) * ITEMS;
    if (base + ITEMS <= n && (reinterpret_cast<uintptr_t>(p + base) % 16) == 0) {
        uint4* q = reinterpret_cast<uint4*>(p + base);
        const uint4* o = reinterpret_cast<const uint4*>(v);
#pragma unroll
        for (int k = 0; k < BYTES / 16; ++k) q[k] = o[k];
    } else {
#pragma unroll
        for (int e = 0; e < ITEMS; ++e)
            if (base + e < n) p[base + e] = v[e];
    }
}

// T in {float,double}: dtype of v_s, v_next, rew and of both outputs (float64 = the numpy path
// of the reference, float32 = device-resident rollouts); the carry is float64 either way
#ifndef TSM_GAE_MIN_BLOCKS
#define TSM_GAE_MIN_BLOCKS 10
#endif
template <typename T>
__global__ void __launch_bounds__(THREADS, (sizeof(T) == 8 ? (TSM_GAE_MIN_BLOCKS * 3) / 5 : TSM_GAE_MIN_BLOCKS)) gae_flat_kernel(
    const T* __restrict__ v_s, const T* __restrict__ v_next, const T* __restrict__ rew,
    const uint8_t* __restrict__ term, const uint8_t* __restrict__ trunc, const uint8_t* __restrict__ force_end,
    int64_t n, int64_t seg_len, double gamma, double gl, T* __restrict__ adv_out, T* __restrict__ ret_out,
    TileState* __restrict__ tiles, int num_tiles) {
    __shared__ Map s_warp[WARPS];
    __shared__ double s_carry;

    // Tiles run right to left in block order: a tile only ever waits on tiles of lower block index, and the hardware hands
    // out the blocks of a one-dimensional grid in increasing index order, so those have started (the assumption CUB's
    // decoupled look-back makes as well).  Claiming the tile with an atomic counter instead put an L2 round trip in front
    // of every tile's loads (4.71 against 4.92 TB/s), and persistent CTAs claiming one tile ahead were slower still.
    const int tile = num_tiles - 1 - (int)blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // the right neighbour started earlier: probe its state now, so that the round trip overlaps this tile's loads
    double nb_incl = 0.0, nb_a = 0.0, nb_b = 0.0;
    bool nb_has_incl = false, nb_has_agg = false;
    if (threadIdx.x == 0 && tile + 1 < num_tiles) {
        const TileState* t = &tiles[tile + 1];
        nb_has_incl = ts_read(&t->incl, nb_incl);
        const bool ha = ts_read(&t->a, nb_a), hb = ts_read(&t->b, nb_b);
        nb_has_agg = ha && hb;
    }
    const int64_t base = (int64_t)tile * TILE + (int64_t)threadIdx.x * ITEMS;

    T vs[ITEMS], vn[ITEMS], rw[ITEMS];
    uint8_t te[ITEMS], tr[ITEMS], fe[ITEMS];
    load_items<T>(v_s, base, n, vs, (T)0);
    load_items<T>(v_next, base, n, vn, (T)0);
    load_items<T>(rew, base, n, rw, (T)0);
    load_items<uint8_t>(term, base, n, te, (uint8_t)0);
    load_items<uint8_t>(trunc, base, n, tr, (uint8_t)0);
    load_items<uint8_t>(force_end, base, n, fe, (uint8_t)0);

    // per element: delta (float64, unfused like numba) and c in {0, gamma*lambda}
    double delta[ITEMS];
    bool endf[ITEMS];
    // position of element `base` inside its segment: one 64-bit modulo per thread, then a running counter
    int64_t seg_pos = seg_len > 0 ? base % seg_len : 0;
#pragma unroll
    for (int e = 0; e < ITEMS; ++e) {
        const int64_t i = base + e;
        double vnm = te[e] ? 0.0 : (double)vn[e];                                   // value_mask, :709
        delta[e] = __dadd_rn(__dadd_rn((double)rw[e], __dmul_rn(vnm, gamma)), -(double)vs[e]);  // :1128
        bool en = (te[e] | tr[e] | fe[e]) != 0;
        if (seg_len > 0) {
            if (++seg_pos == seg_len) { en = true; seg_pos = 0; }
        }
        if (i >= n) { en = true; delta[e] = 0.0; }
        endf[e] = en;
    }
    // pass 1: thread aggregate  g_first = B + A * g_in
    Map m{1.0, 0.0};
#pragma unroll
    for (int e = ITEMS - 1; e >= 0; --e) {
        double c = endf[e] ? 0.0 : gl;
        m.b = __dadd_rn(delta[e], __dmul_rn(c, m.b));  // same rounding as the serial loop
        m.a = c * m.a;
    }
    // warp inclusive suffix scan of maps (Hillis-Steele with shfl_down)
    Map incl = m;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Map r{__shfl_down_sync(0xffffffffu, incl.a, d), __shfl_down_sync(0xffffffffu, incl.b, d)};
        if (lane + d < 32) incl = compose(incl, r);
    }
    Map excl{__shfl_down_sync(0xffffffffu, incl.a, 1), __shfl_down_sync(0xffffffffu, incl.b, 1)};
    if (lane == 31) excl = Map{1.0, 0.0};
    if (lane == 0) s_warp[warp] = incl;
    __syncthreads();

    if (threadIdx.x == 0) {
        Map blk{1.0, 0.0};
#pragma unroll
        for (int w = WARPS - 1; w >= 0; --w) blk = compose(s_warp[w], blk);
        TileState* me = &tiles[tile];
        double carry = 0.0;
        if (tile == num_tiles - 1 || blk.a == 0.0) {
            ts_publish(&me->incl, blk.b);  // nothing to the right reaches the tile's first element: its value is just b
        } else {
            ts_publish(&me->a, blk.a);
            ts_publish(&me->b, blk.b);
        }
        if (tile != num_tiles - 1) {
            // decoupled look-back over the tiles to the right, starting from the early probe
            Map run{1.0, 0.0};
            int j = tile + 1;
            while (true) {
                if (j >= num_tiles) { carry = run.b; break; }
                double ti, ta, tb;
                bool hi, hg;
                if (j == tile + 1 && (nb_has_incl || nb_has_agg)) {
                    ti = nb_incl; ta = nb_a; tb = nb_b; hi = nb_has_incl; hg = nb_has_agg;
                } else {
                    const TileState* t = &tiles[j];
                    while (true) {
                        hi = ts_read(&t->incl, ti);
                        if (hi) break;
                        const bool ha = ts_read(&t->a, ta), hb = ts_read(&t->b, tb);
                        hg = ha && hb;
                        if (hg) break;
                        __nanosleep(20);
                    }
                }
                if (hi) { carry = fma(run.a, ti, run.b); break; }
                run = compose(run, Map{ta, tb});
                if (run.a == 0.0) { carry = run.b; break; }
                ++j;
            }
            if (blk.a != 0.0) ts_publish(&me->incl, fma(blk.a, carry, blk.b));
        }
        s_carry = carry;
    }
    __syncthreads();

    // carry into this warp = maps of the warps to its right applied to the tile carry
    double g = s_carry;
    for (int w = WARPS - 1; w > warp; --w) g = fma(s_warp[w].a, g, s_warp[w].b);
    g = fma(excl.a, g, excl.b);  // carry into this thread

    // pass 2: the reference's serial recurrence with the true carry (:1131-1133)
    T adv[ITEMS], ret[ITEMS];
#pragma unroll
    for (int e = ITEMS - 1; e >= 0; --e) {
        double c = endf[e] ? 0.0 : gl;
        g = __dadd_rn(delta[e], __dmul_rn(c, g));
        adv[e] = (T)g;
        ret[e] = (T)__dadd_rn(g, (double)vs[e]);  // :715
    }
    if (adv_out) store_items<T>(adv_out, base, n, adv);
    if (ret_out) store_items<T>(ret_out, base, n, ret);
}

// ---------------------------------------------------------------------------------------------
// time-major: one thread per row, serial over t (bit-exact float64 recurrence of the reference)
constexpr int TM_THREADS = 128;
constexpr int TM_UNROLL = 8;

// SCALED (return_scaling, a2c.py:132-146; return_standardization, reinforce.py:297-312): values enter multiplied by v_scale =
// sqrt(ret_rms.var + eps), returns leave as (return - ret_shift) / v_scale, and the sum / sum of squares of the UNNORMALISED returns go to ret_sums[0..1] (what RunningMeanStd.update needs).
template <bool SCALED>
__global__ void __launch_bounds__(TM_THREADS) gae_tm_kernel(
    const float* __restrict__ v_s, const float* __restrict__ v_next, const float* __restrict__ rew,
    const uint8_t* __restrict__ term, const uint8_t* __restrict__ trunc, int64_t size, int64_t R, int64_t start,
    int64_t len, double gamma, double gl, float* __restrict__ adv_out, float* __restrict__ ret_out, double v_scale,
    double ret_shift, double* __restrict__ ret_sums) {
    const int64_t r_raw = (int64_t)blockIdx.x * TM_THREADS + threadIdx.x;
    const bool live = r_raw < R;
    if (!SCALED && !live) return;
    const int64_t r = live ? r_raw : 0;
    double g = 0.0, acc_s = 0.0, acc_q = 0.0;
    int64_t j = live ? len - 1 : -1;
    int64_t tj = (start + j) % size;  // ring slot of element j; walks backwards with one wrap
    // main loop: TM_UNROLL independent loads in flight per array, then the dependent chain
    while (j >= 0) {
        float a_vs[TM_UNROLL], a_vn[TM_UNROLL], a_rw[TM_UNROLL];
        uint8_t a_te[TM_UNROLL], a_tr[TM_UNROLL];
        int64_t off[TM_UNROLL];
#pragma unroll
        for (int u = 0; u < TM_UNROLL; ++u) {
            int64_t jj = j - u;
            bool ok = jj >= 0;  // ok implies u <= j < size, so a single wrap is enough
            int64_t t = tj - u;
            if (t < 0) t += size;
            off[u] = ok ? t * R + r : r;
            a_vs[u] = (ok && v_s) ? v_s[off[u]] : 0.f;
            a_vn[u] = (ok && v_next) ? v_next[off[u]] : 0.f;
            a_rw[u] = ok ? rew[off[u]] : 0.f;
            a_te[u] = ok ? term[off[u]] : (uint8_t)0;
            a_tr[u] = ok ? trunc[off[u]] : (uint8_t)0;
        }
#pragma unroll
        for (int u = 0; u < TM_UNROLL; ++u) {
            int64_t jj = j - u;
            if (jj < 0) break;
            double vs_d = (double)a_vs[u], vn_d = (double)a_vn[u];
            if (SCALED) { vs_d = __dmul_rn(vs_d, v_scale); vn_d = __dmul_rn(vn_d, v_scale); }
            double vnm = a_te[u] ? 0.0 : vn_d;
            double delta = __dadd_rn(__dadd_rn((double)a_rw[u], __dmul_rn(vnm, gamma)), -vs_d);
            bool en = (a_te[u] | a_tr[u]) != 0 || jj == len - 1;
            double c = en ? 0.0 : gl;
            g = __dadd_rn(delta, __dmul_rn(c, g));
            if (adv_out) adv_out[off[u]] = (float)g;
            const double ru = __dadd_rn(g, vs_d);
            if (SCALED) {
                acc_s += ru;
                acc_q = fma(ru, ru, acc_q);
                if (ret_out) ret_out[off[u]] = (float)((ru - ret_shift) / v_scale);
            } else {
                if (ret_out) ret_out[off[u]] = (float)ru;
            }
        }
        j -= TM_UNROLL;
        tj -= TM_UNROLL;
        if (tj < 0) tj = ((tj % size) + size) % size;
    }
    if (SCALED && ret_sums) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            acc_s += __shfl_down_sync(0xffffffffu, acc_s, d);
            acc_q += __shfl_down_sync(0xffffffffu, acc_q, d);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(ret_sums, acc_s);
            atomicAdd(ret_sums + 1, acc_q);
        }
    }
}

}  // namespace gae
}  // namespace tsm

using namespace tsm;
using namespace tsm::gae;

// ---------------------------------------------------------------------------------------------
// critic(obs_next) without a second full forward pass (a2c.py:120-127 evaluates the critic on obs and on obs_next).
// In a rollout obs_next[t] IS obs[t + 1] unless the row's episode ended at t (the env was reset) or t is the last
// collected slot, so v(obs_next[t]) = value[t + 1] there -- the value the policy forward already produced at collect time.
// This kernel writes those shifted values and appends the (slot, row) pairs that do need a critic evaluation of obs_next
// to need_list (order irrelevant: the forward kernel scatters by index).
namespace tsm {
namespace gae {
__global__ void __launch_bounds__(256) values_next_kernel(const float* __restrict__ value, const uint8_t* __restrict__ done, int64_t size,
                                                         int64_t n_rows, int64_t start, int64_t len, float* __restrict__ v_next,
                                                         int64_t* __restrict__ need_list, unsigned long long* __restrict__ need_count) {
    const int64_t total = len * n_rows;
    const int lane = threadIdx.x & 31;
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < total; i0 += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = i0 + lane;
        bool need = false;
        int64_t at = 0;
        if (i < total) {
            const int64_t j = i / n_rows, r = i - j * n_rows;     // j-th slot in time order
            const int64_t p = (start + j) % size;                 // its ring slot
            at = p * n_rows + r;
            need = done[at] != 0 || j == len - 1;
            v_next[at] = need ? 0.f : value[((p + 1) % size) * n_rows + r];
        }
        const unsigned m = __ballot_sync(0xffffffffu, need);
        if (m) {
            unsigned long long base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(need_count, (unsigned long long)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (need) need_list[base + __popc(m & ((1u << lane) - 1u))] = at;
        }
    }
}
}  // namespace gae
}  // namespace tsm

extern "C" int tsm_values_next_prepare(const float* value, const uint8_t* done, int64_t size, int64_t n_rows, int64_t start, int64_t len,
                                       float* v_next, int64_t* need_list, int64_t* need_count, tsm_stream_t stream) {
    using namespace tsm;
    TSM_REQUIRE(value && done && v_next && need_list && need_count, "value / done / v_next / need_list / need_count != NULL");
    TSM_REQUIRE(size > 0 && n_rows > 0 && start >= 0 && start < size && len >= 0 && len <= size, "size, n_rows > 0; 0 <= start < size; 0 <= len <= size");
    cudaStream_t s = as_stream(stream);
    TSM_CUDA(cudaMemsetAsync(need_count, 0, sizeof(int64_t), s));
    if (len == 0) return TSM_OK;
    const int64_t total = len * n_rows;
    int64_t grid = ceil_div(total, (int64_t)256);
    const int64_t cap = (int64_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    gae::values_next_kernel<<<(unsigned)grid, 256, 0, s>>>(value, done, size, n_rows, start, len, v_next, need_list,
                                                          reinterpret_cast<unsigned long long*>(need_count));
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" size_t tsm_gae_flat_workspace_bytes(int64_t n) {
    if (n < 0) n = 0;
    int64_t tiles = ceil_div(n, TILE);
    return 64 + (size_t)tiles * sizeof(TileState);
}

extern "C" int tsm_gae_flat(const void* v_s, const void* v_next, const void* rew, int io_is_f64,
                            const uint8_t* terminated, const uint8_t* truncated, const uint8_t* force_end,
                            int64_t n, int64_t seg_len, double gamma, double lam, void* adv_out, void* ret_out,
                            void* workspace, size_t workspace_bytes, tsm_stream_t stream) {
    TSM_REQUIRE(n >= 0, "n >= 0");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(rew != nullptr, "rew != NULL");
    TSM_REQUIRE(terminated != nullptr && truncated != nullptr, "terminated/truncated != NULL");
    TSM_REQUIRE(adv_out != nullptr || ret_out != nullptr, "an output pointer");
    TSM_REQUIRE(n / TILE < (int64_t)INT32_MAX, "n too large");
    if (workspace == nullptr || workspace_bytes < tsm_gae_flat_workspace_bytes(n))
        return fail(TSM_ERR_WORKSPACE, "%s: workspace too small (need %lld bytes)", __func__,
                    (long long)tsm_gae_flat_workspace_bytes(n));
    TSM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) % 32) == 0, "workspace 32-byte aligned");
    cudaStream_t s = as_stream(stream);
    const int num_tiles = (int)ceil_div(n, TILE);
    TSM_CUDA(cudaMemsetAsync(reinterpret_cast<char*>(workspace) + 64, 0xFF, tsm_gae_flat_workspace_bytes(n) - 64, s));  // tile states: empty
    TileState* tiles = reinterpret_cast<TileState*>(reinterpret_cast<char*>(workspace) + 64);
    const double gl = gamma * lam;
    if (io_is_f64)
        gae_flat_kernel<double><<<num_tiles, THREADS, 0, s>>>(
            (const double*)v_s, (const double*)v_next, (const double*)rew, terminated, truncated, force_end, n, seg_len,
            gamma, gl, (double*)adv_out, (double*)ret_out, tiles, num_tiles);
    else
        gae_flat_kernel<float><<<num_tiles, THREADS, 0, s>>>(
            (const float*)v_s, (const float*)v_next, (const float*)rew, terminated, truncated, force_end, n, seg_len,
            gamma, gl, (float*)adv_out, (float*)ret_out, tiles, num_tiles);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

// ---------------------------------------------------------------------------------------------
// tsm_gae_flat_static: the same kernel with its scratch (64 reserved bytes + tile states) in a file-scope device array instead
// of a caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving loops).  No counterpart in
// the reference.  Every GPU that loads the library holds the array until
// exit.  All calls share it: calls that overlap on one GPU (two streams) race; it is re-initialised on the caller's
// stream at the start of every call (tile states = 0xFF), so nothing is assumed about its previous contents.
constexpr int64_t STATIC_MAX_N = TSM_GAE_FLAT_STATIC_MAX_N;
constexpr size_t STATIC_BYTES = 64 + (size_t)((STATIC_MAX_N + TILE - 1) / TILE) * sizeof(TileState);
static __device__ __align__(256) unsigned char g_gae_flat_scratch[STATIC_BYTES];

// In host code the array's name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
static int gae_flat_static_scratch(void** out, const char* fn) {
    constexpr int MAX_DEV = 64;
    static void* ptrs[MAX_DEV] = {nullptr};
    int dev = 0;
    TSM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return fail(TSM_ERR_UNSUPPORTED, "%s: device ordinal %d out of range", fn, dev);
    if (!ptrs[dev]) {
        void* p = nullptr;
        TSM_CUDA(cudaGetSymbolAddress(&p, g_gae_flat_scratch));
        if (!p) return fail(TSM_ERR_CUDA, "%s: cudaGetSymbolAddress returned NULL", fn);
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return TSM_OK;
}

extern "C" int64_t tsm_gae_flat_static_max_n(void) { return STATIC_MAX_N; }

extern "C" int tsm_gae_flat_static(const void* v_s, const void* v_next, const void* rew, int io_is_f64,
                                   const uint8_t* terminated, const uint8_t* truncated, const uint8_t* force_end,
                                   int64_t n, int64_t seg_len, double gamma, double lam, void* adv_out, void* ret_out,
                                   tsm_stream_t stream) {
    TSM_REQUIRE(n >= 0, "n >= 0");
    if (n > STATIC_MAX_N)
        return fail(TSM_ERR_INVALID, "%s: n = %lld exceeds the static scratch (TSM_GAE_FLAT_STATIC_MAX_N = %lld elements); use tsm_gae_flat",
                    __func__, (long long)n, (long long)STATIC_MAX_N);
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(rew != nullptr, "rew != NULL");
    TSM_REQUIRE(terminated != nullptr && truncated != nullptr, "terminated/truncated != NULL");
    TSM_REQUIRE(adv_out != nullptr || ret_out != nullptr, "an output pointer");
    void* scratch = nullptr;
    const int rc = gae_flat_static_scratch(&scratch, __func__);
    if (rc) return rc;
    cudaStream_t s = as_stream(stream);
    const int num_tiles = (int)ceil_div(n, TILE);
    const size_t used = 64 + (size_t)num_tiles * sizeof(TileState);
    TSM_CUDA(cudaMemsetAsync(reinterpret_cast<char*>(scratch) + 64, 0xFF, used - 64, s));       // tile states: empty
    TileState* tiles = reinterpret_cast<TileState*>(reinterpret_cast<char*>(scratch) + 64);
    const double gl = gamma * lam;
    if (io_is_f64)
        gae_flat_kernel<double><<<num_tiles, THREADS, 0, s>>>(
            (const double*)v_s, (const double*)v_next, (const double*)rew, terminated, truncated, force_end, n, seg_len,
            gamma, gl, (double*)adv_out, (double*)ret_out, tiles, num_tiles);
    else
        gae_flat_kernel<float><<<num_tiles, THREADS, 0, s>>>(
            (const float*)v_s, (const float*)v_next, (const float*)rew, terminated, truncated, force_end, n, seg_len,
            gamma, gl, (float*)adv_out, (float*)ret_out, tiles, num_tiles);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_gae_tm(const float* v_s, const float* v_next, const float* rew, const uint8_t* terminated,
                          const uint8_t* truncated, int64_t size, int64_t n_rows, int64_t start, int64_t len,
                          double gamma, double lam, float* adv_out, float* ret_out, tsm_stream_t stream) {
    TSM_REQUIRE(size > 0 && n_rows >= 0, "size > 0, n_rows >= 0");
    TSM_REQUIRE(len >= 0 && len <= size && start >= 0 && start < size, "0 <= len <= size, 0 <= start < size");
    if (n_rows == 0 || len == 0) return TSM_OK;
    TSM_REQUIRE(rew && terminated && truncated, "rew/terminated/truncated != NULL");
    TSM_REQUIRE(adv_out != nullptr || ret_out != nullptr, "an output pointer");
    const int64_t blocks = ceil_div(n_rows, TM_THREADS);
    TSM_REQUIRE(blocks < (int64_t)INT32_MAX, "n_rows too large");
    gae_tm_kernel<false><<<(unsigned)blocks, TM_THREADS, 0, as_stream(stream)>>>(
        v_s, v_next, rew, terminated, truncated, size, n_rows, start, len, gamma, gamma * lam, adv_out, ret_out, 1.0, 0.0, nullptr);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_gae_tm_scaled(const float* v_s, const float* v_next, const float* rew, const uint8_t* terminated,
                                 const uint8_t* truncated, int64_t size, int64_t n_rows, int64_t start, int64_t len,
                                 double gamma, double lam, double v_scale, double ret_shift, float* adv_out, float* ret_out,
                                 double* ret_sums, tsm_stream_t stream) {
    TSM_REQUIRE(size > 0 && n_rows >= 0, "size > 0, n_rows >= 0");
    TSM_REQUIRE(len >= 0 && len <= size && start >= 0 && start < size, "0 <= len <= size, 0 <= start < size");
    TSM_REQUIRE(v_scale > 0.0, "v_scale > 0");
    if (ret_sums) TSM_CUDA(cudaMemsetAsync(ret_sums, 0, 2 * sizeof(double), as_stream(stream)));
    if (n_rows == 0 || len == 0) return TSM_OK;
    TSM_REQUIRE(rew && terminated && truncated, "rew/terminated/truncated != NULL");
    TSM_REQUIRE(adv_out != nullptr || ret_out != nullptr, "an output pointer");
    const int64_t blocks = ceil_div(n_rows, TM_THREADS);
    TSM_REQUIRE(blocks < (int64_t)INT32_MAX, "n_rows too large");
    gae_tm_kernel<true><<<(unsigned)blocks, TM_THREADS, 0, as_stream(stream)>>>(
        v_s, v_next, rew, terminated, truncated, size, n_rows, start, len, gamma, gamma * lam, adv_out, ret_out, v_scale, ret_shift, ret_sums);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
