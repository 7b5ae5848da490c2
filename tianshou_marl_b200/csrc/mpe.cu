// mpe.cu -- K1: batched MPE simple_spread / simple_tag / simple_adversary world step (sm_100a).
//
// Restates pettingzoo 1.24.2 mpe (core.World.step, simple_env._set_action /
// _execute_world_step, scenarios/simple_spread.py, scenarios/simple_tag.py), which the
// reference drives through tianshou/env/venvs.py:237-322 and its PettingZoo wrappers.
//
// Mapping: G lanes (a sub-warp group, G in {4,8,16,32}) own one environment; lane j is agent j
// (and carries landmark j).  Positions travel between lanes with __shfl_sync, so an env never
// touches shared memory.  State is SoA per field with float2 elements, indexed by row
// (env*A + agent), so consecutive lanes read consecutive 8-byte words: every state load and
// store of a warp is one contiguous segment.
#include <cstdlib>
#include "common.cuh"
#include "philox.cuh"

namespace tsm {
namespace mpe {

constexpr float DT = 0.1f, DAMPING = 0.25f, CONTACT_FORCE = 1e2f, CONTACT_MARGIN = 1e-3f;
constexpr int THREADS = 256;

__device__ __forceinline__ float agent_size(const tsm_mpe& e, int j) {
    if (e.scenario != TSM_MPE_TAG) return 0.15f;  // simple_spread and simple_adversary: every agent 0.15
    return j < e.n_adversaries ? 0.075f : 0.05f;
}
__device__ __forceinline__ float agent_accel(const tsm_mpe& e, int j) {
    if (e.scenario != TSM_MPE_TAG) return 5.0f;  // simple_env._set_action default sensitivity (accel None)
    return j < e.n_adversaries ? 3.0f : 4.0f;
}
__device__ __forceinline__ float agent_max_speed(const tsm_mpe& e, int j) {
    if (e.scenario != TSM_MPE_TAG) return INFINITY;
    return j < e.n_adversaries ? 1.0f : 1.3f;
}
__device__ __forceinline__ float landmark_size(const tsm_mpe& e) { return e.scenario == TSM_MPE_SPREAD ? 0.05f : 0.2f; }

// core.World.get_collision_force: force on `me` from `other`
__device__ __forceinline__ float2 collision_force(float2 me, float2 other, float dist_min) {
    const float dx = me.x - other.x, dy = me.y - other.y;
    const float dist = sqrtf(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
    const float x = -(dist - dist_min) / CONTACT_MARGIN;
    // np.logaddexp(0, x) = max(x,0) + log1p(exp(-|x|))
    const float pen = (fmaxf(x, 0.f) + log1pf(expf(-fabsf(x)))) * CONTACT_MARGIN;
    return make_float2(CONTACT_FORCE * dx / dist * pen, CONTACT_FORCE * dy / dist * pen);
}

__device__ __forceinline__ float dist2d(float2 a, float2 b) {
    const float dx = a.x - b.x, dy = a.y - b.y;
    return sqrtf(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// shuffles are issued with the mask of the caller's own G-lane group, so groups of one warp may
// diverge from each other (e.g. only some envs reset) without violating __shfl_sync's contract
template <int G>
__device__ __forceinline__ unsigned group_mask() {
    if constexpr (G >= 32) {
        return 0xffffffffu;
    } else {
        const unsigned lane = threadIdx.x & 31u;
        return ((1u << G) - 1u) << (lane & ~(unsigned)(G - 1));
    }
}
template <int G>
__device__ __forceinline__ float2 shfl2(float2 v, int src) {
    const unsigned m = group_mask<G>();
    return make_float2(__shfl_sync(m, v.x, src, G), __shfl_sync(m, v.y, src, G));
}
template <int G>
__device__ __forceinline__ float shfl1(float v, int src) {
    return __shfl_sync(group_mask<G>(), v, src, G);
}

// draw the reset position of entity `ent` (agents 0..A-1 then landmarks) of env `env_id`
__device__ __forceinline__ float2 draw_entity(const tsm_mpe& e, int64_t env_id, int64_t episode, int ent) {
    const Philox4 r = philox4x32((uint32_t)env_id, (uint32_t)((uint64_t)env_id >> 32), (uint32_t)episode, (uint32_t)(ent >> 1),
                                 (uint32_t)e.seed, (uint32_t)(e.seed >> 32) ^ STREAM_RESET);
    const float u0 = u01((ent & 1) ? r.z : r.x), u1 = u01((ent & 1) ? r.w : r.y);
    float lo = -1.0f, w = 2.0f;
    if (ent >= e.n_agents && e.scenario == TSM_MPE_TAG) { lo = -0.9f; w = 1.8f; }
    return make_float2(__fadd_rn(lo, __fmul_rn(w, u0)), __fadd_rn(lo, __fmul_rn(w, u1)));
}

// simple_adversary: the goal landmark of episode `episode` of env `env_id` (scenario.reset_world: np_random.choice(landmarks)).
// Not stored: it is a pure function of (seed, env, episode) -- the Philox draw of the entity slot behind the last landmark.
__device__ __forceinline__ int goal_index(const tsm_mpe& e, int64_t env_id, int64_t episode) {
    const int ent = e.n_agents + e.n_landmarks;
    const Philox4 r = philox4x32((uint32_t)env_id, (uint32_t)((uint64_t)env_id >> 32), (uint32_t)episode, (uint32_t)(ent >> 1),
                                 (uint32_t)e.seed, (uint32_t)(e.seed >> 32) ^ STREAM_RESET);
    const float u = u01((ent & 1) ? r.z : r.x);
    const int g = (int)(u * (float)e.n_landmarks);
    return g < e.n_landmarks ? g : e.n_landmarks - 1;
}

// scenario.observation for agent j (lane j); all lanes of the group must call (shuffles inside)
template <int G>
__device__ __forceinline__ void write_obs(const tsm_mpe& e, int j, bool valid, float2 p, float2 v, float2 lm,
                                          float* __restrict__ dst0, float* __restrict__ dst1, int goal = 0) {
    const int A = e.n_agents, L = e.n_landmarks;
    float2* o0 = reinterpret_cast<float2*>(dst0);
    float2* o1 = reinterpret_cast<float2*>(dst1);
    int q = 0;
    auto put = [&](float2 val) {
        if (valid) {
            if (dst0) o0[q] = val;
            if (dst1) o1[q] = val;
        }
        ++q;
    };
    if (e.scenario == TSM_MPE_ADVERSARY) {
        // simple_adversary.observation: good agents [goal - p, landmarks - p, others - p]; the adversary [landmarks - p, others - p]
        // (no velocities, no own position); the adversary's 2 + 2L + 2(A-1) - 2 values are zero-padded to the common obs_dim
        const float2 pg = shfl2<G>(lm, goal);
        if (j >= e.n_adversaries) put(make_float2(pg.x - p.x, pg.y - p.y));
        for (int l = 0; l < L; ++l) {
            const float2 pl = shfl2<G>(lm, l);
            put(make_float2(pl.x - p.x, pl.y - p.y));
        }
        for (int k = 0; k < A; ++k) {
            const float2 pk = shfl2<G>(p, k);
            if (k != j) put(make_float2(pk.x - p.x, pk.y - p.y));
        }
        const int O2 = e.obs_dim / 2;
        while (q < O2) put(make_float2(0.f, 0.f));
        return;
    }
    put(v);
    put(p);
    for (int l = 0; l < L; ++l) {
        const float2 pl = shfl2<G>(lm, l);
        put(make_float2(pl.x - p.x, pl.y - p.y));
    }
    for (int k = 0; k < A; ++k) {
        const float2 pk = shfl2<G>(p, k);
        if (k != j) put(make_float2(pk.x - p.x, pk.y - p.y));
    }
    if (e.scenario == TSM_MPE_SPREAD) {
        for (int k = 0; k < A - 1; ++k) put(make_float2(0.f, 0.f));  // comm of silent agents
    } else {
        for (int k = e.n_adversaries; k < A; ++k) {  // velocities of the good agents other than me
            const float2 vk = shfl2<G>(v, k);
            if (k != j) put(vk);
        }
        const int O2 = e.obs_dim / 2;
        while (q < O2) put(make_float2(0.f, 0.f));  // good agent: zero-pad to the common obs_dim
    }
}

template <int G>
__global__ void __launch_bounds__(THREADS) reset_kernel(tsm_mpe e, const int64_t* __restrict__ env_ids, int64_t n,
                                                       const float* __restrict__ init_pos, float* __restrict__ obs_out) {
    const int64_t gid = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / G;
    const int j = threadIdx.x % G;
    const bool live = gid < n;
    const int A = e.n_agents, L = e.n_landmarks;
    const int64_t env = live ? (env_ids ? env_ids[gid] : gid) : 0;
    float2 p = make_float2(0.f, 0.f), lm = make_float2(0.f, 0.f);
    const float2 v = make_float2(0.f, 0.f);
    if (live) {
        const int64_t ep = e.episode[env];
        if (j < A) {
            p = init_pos ? reinterpret_cast<const float2*>(init_pos)[gid * (A + L) + j] : draw_entity(e, env, ep, j);
            reinterpret_cast<float2*>(e.pos)[env * A + j] = p;
            reinterpret_cast<float2*>(e.vel)[env * A + j] = v;
        }
        if (j < L) {
            lm = init_pos ? reinterpret_cast<const float2*>(init_pos)[gid * (A + L) + A + j] : draw_entity(e, env, ep, A + j);
            reinterpret_cast<float2*>(e.lm)[env * L + j] = lm;
        }
    }
    const int goal = (live && e.scenario == TSM_MPE_ADVERSARY) ? goal_index(e, env, e.episode[env]) : 0;
    write_obs<G>(e, j, live && j < A && obs_out != nullptr, p, v, lm,
                 obs_out ? obs_out + (gid * A + j) * e.obs_dim : nullptr, nullptr, goal);
    if (live && j == 0) {
        e.steps[env] = 0;
        e.episode[env] += 1;
    }
}

// ---- what a lock-step rollout does after the world step, fused into the same launch (tsm_rollout_step) ----
// per row: the ring-buffer state machine of ReplayBufferManager.add (manager.py:131-193, buffer_base.py:355-410; the
// standalone version is buf::add_state_kernel) and the episode statistics of Collector._collect (collector.py:983-1024;
// standalone: buf::episode_stats_kernel).
struct RolloutTail {
    tsm_buf b;
    double* ep_ret_out;   // [R] episode return of rows that finished an episode, else 0 (tsm_buf_add's outputs)
    int64_t* ep_len_out;  // [R]
    double* acc;          // [9] {count, sum ret, sum ret^2, min ret, max ret, sum len, sum len^2, min len, max len}
    double* list_ret;     // optional episode list
    int64_t* list_len;
    int64_t* list_count;
    int64_t list_cap;
};
__device__ __forceinline__ void atomic_min_f64(double* a, double v) {
    unsigned long long* p = reinterpret_cast<unsigned long long*>(a);
    unsigned long long old = *p;
    while (__longlong_as_double((long long)old) > v) {
        const unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v));
        if (prev == old) break;
        old = prev;
    }
}
__device__ __forceinline__ void atomic_max_f64(double* a, double v) {
    unsigned long long* p = reinterpret_cast<unsigned long long*>(a);
    unsigned long long old = *p;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v));
        if (prev == old) break;
        old = prev;
    }
}

// STAGE: observations are laid down in shared memory by their owner lanes and copied out by the whole block as one
// contiguous range (the worlds of a block are consecutive, so their rows are too).  Direct per-lane writes put consecutive
// lanes obs_dim * 4 bytes apart: every 8-byte store is its own 32-byte sector (N = 10: 60 such stores per lane and
// destination, 102 us per 163 840 rows before this change).
template <int G, bool TAIL, bool STAGE>
__global__ void __launch_bounds__(THREADS) step_kernel(tsm_mpe e, const int32_t* __restrict__ act,
                                                      float* __restrict__ obs_next, float* __restrict__ rew,
                                                      uint8_t* __restrict__ term, uint8_t* __restrict__ trunc, int autoreset,
                                                      float* __restrict__ obs_carry, RolloutTail t, int64_t env0, int64_t n_sub,
                                                      const float* __restrict__ act_f) {
    // the launch covers environments [env0, env0 + n_sub) (all arrays stay indexed by the global env / row)
    const int64_t local = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / G;
    const int64_t env = env0 + local;
    const int j = threadIdx.x % G;
    const int A = e.n_agents, L = e.n_landmarks;
    const bool live = local < n_sub;
    const bool is_agent = live && j < A;
    const int64_t row = env * A + j;

    float2 p = make_float2(0.f, 0.f), v = make_float2(0.f, 0.f), lm = make_float2(0.f, 0.f);
    int a = 0;
    if (is_agent) {
        p = reinterpret_cast<const float2*>(e.pos)[row];
        v = reinterpret_cast<const float2*>(e.vel)[row];
    }
    if (live && j < L) lm = reinterpret_cast<const float2*>(e.lm)[env * L + j];
    // Rollout chain (collect.cu): launched with programmatic stream serialization behind the policy forward that writes `act`.
    // The world state loaded above was written by the previous world step, which completed before that forward passed its own
    // wait (and only then let this kernel be scheduled).  No-ops for a normal launch.
    pdl_wait();
    pdl_launch_dependents();
    // continuous_actions=True (simple_env._set_action): a row's action is a Box(0, 1, (5,)) vector, u = (a[2] - a[1], a[4] - a[3])
    // in float32, after ClipOutOfBoundsWrapper (mpe make_env wraps continuous envs with it) clipped it to the box
    float ux = 0.f, uy = 0.f;
    if (is_agent) {
        if (act_f) {
            const float* af = act_f + row * 5;
            const float a1 = fminf(fmaxf(af[1], 0.f), 1.f), a2 = fminf(fmaxf(af[2], 0.f), 1.f);
            const float a3 = fminf(fmaxf(af[3], 0.f), 1.f), a4 = fminf(fmaxf(af[4], 0.f), 1.f);
            ux = __fsub_rn(a2, a1);
            uy = __fsub_rn(a4, a3);
        } else {
            a = act[row];
            ux = a == 1 ? -1.f : (a == 2 ? 1.f : 0.f);
            uy = a == 3 ? -1.f : (a == 4 ? 1.f : 0.f);
        }
    }
    const float my_size = agent_size(e, j);

    // simple_env._set_action (discrete) + World.apply_action_force
    const float sens = agent_accel(e, j);
    float fx = __fmul_rn(ux, sens);
    float fy = __fmul_rn(uy, sens);
    // World.apply_environment_force: partners in ascending entity order (agents, then landmarks); simple_adversary has no
    // collidable entity at all (collide = False everywhere)
    for (int k = 0; k < (e.scenario == TSM_MPE_ADVERSARY ? 0 : A); ++k) {
        const float2 pk = shfl2<G>(p, k);
        if (k != j) {
            const float2 f = collision_force(p, pk, my_size + agent_size(e, k));
            fx = __fadd_rn(f.x, fx);
            fy = __fadd_rn(f.y, fy);
        }
    }
    if (e.scenario == TSM_MPE_TAG) {  // obstacles collide (simple_tag.make_world); spread landmarks do not
        for (int l = 0; l < L; ++l) {
            const float2 pl = shfl2<G>(lm, l);
            const float2 f = collision_force(p, pl, my_size + landmark_size(e));
            fx = __fadd_rn(f.x, fx);
            fy = __fadd_rn(f.y, fy);
        }
    }
    // World.integrate_state (mass 1)
    if (e.pos_first) {
        p.x = __fadd_rn(p.x, __fmul_rn(v.x, DT));
        p.y = __fadd_rn(p.y, __fmul_rn(v.y, DT));
    }
    v.x = __fadd_rn(__fmul_rn(v.x, 1.f - DAMPING), __fmul_rn(fx, DT));
    v.y = __fadd_rn(__fmul_rn(v.y, 1.f - DAMPING), __fmul_rn(fy, DT));
    const float ms = agent_max_speed(e, j);
    const float speed = sqrtf(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)));
    if (speed > ms) {
        v.x = v.x / speed * ms;
        v.y = v.y / speed * ms;
    }
    if (!e.pos_first) {
        p.x = __fadd_rn(p.x, __fmul_rn(v.x, DT));
        p.y = __fadd_rn(p.y, __fmul_rn(v.y, DT));
    }

    // rewards on the post-step state
    float r = 0.f;
    int goal = 0;  // simple_adversary: goal landmark of the running episode
    if (e.scenario == TSM_MPE_SPREAD) {
        float local = 0.f;
        for (int k = 0; k < A; ++k) {
            const float2 pk = shfl2<G>(p, k);
            if (k != j && dist2d(pk, p) < agent_size(e, k) + my_size) local -= 1.f;
        }
        // global_reward: lane l holds landmark l; min over agents, then ordered sum over landmarks
        float mind = INFINITY;
        for (int k = 0; k < A; ++k) {
            const float2 pk = shfl2<G>(p, k);
            mind = fminf(mind, dist2d(pk, lm));
        }
        float g = 0.f;
        for (int l = 0; l < L; ++l) g -= shfl1<G>(mind, l);
        r = __fadd_rn(__fmul_rn(g, 1.f - e.local_ratio), __fmul_rn(local, e.local_ratio));
    } else if (e.scenario == TSM_MPE_ADVERSARY) {
        // simple_adversary.reward (shaped): adversary -|p - goal|; good agents sum_adv |p_adv - goal| - min_good |p_good - goal|
        goal = live ? goal_index(e, env, e.episode[env] - 1) : 0;
        const float2 pg = shfl2<G>(lm, goal);
        const float dg = dist2d(p, pg);
        float adv_rew = 0.f, mind = INFINITY;
        for (int k = 0; k < A; ++k) {
            const float dk = shfl1<G>(dg, k);
            if (k < e.n_adversaries) adv_rew += dk;
            else mind = fminf(mind, dk);
        }
        r = j < e.n_adversaries ? -dg : __fadd_rn(-mind, adv_rew);
    } else {
        const bool adv = j < e.n_adversaries;
        // number of adversaries in collision with me (meaningful on good-agent lanes)
        float caught = 0.f;
        for (int k = 0; k < e.n_adversaries; ++k) {
            const float2 pk = shfl2<G>(p, k);
            if (dist2d(pk, p) < agent_size(e, k) + my_size) caught += 1.f;
        }
        float total = 0.f;  // sum over good agents, in agent order
        for (int k = e.n_adversaries; k < A; ++k) total += shfl1<G>(caught, k);
        if (adv) {
            r = 10.f * total;
        } else {
            r = -10.f * caught;
            const float ax[2] = {fabsf(p.x), fabsf(p.y)};
#pragma unroll
            for (int d = 0; d < 2; ++d) {
                const float x = ax[d];
                float b = 0.f;
                if (x >= 0.9f) b = x < 1.0f ? (x - 0.9f) * 10.f : fminf(expf(2.f * x - 2.f), 10.f);
                r -= b;
            }
        }
    }

    // bookkeeping: steps, truncation (simple_env.step), no termination in these scenarios
    int st = 0;
    if (live) st = e.steps[env] + 1;
    const bool tr = live && st >= e.max_cycles;
    if (is_agent) {
        rew[row] = r;
        term[row] = 0;
        trunc[row] = tr ? 1 : 0;
    }
    const bool do_reset = autoreset && tr;
    // observation of the post-step state goes to obs_next (terminal obs for done envs) and, when
    // the env keeps running, also straight to the carry slot the next policy forward reads
    extern __shared__ float2 s_obs[];                 // STAGE: [worlds of the block][A * obs_dim / 2]
    __shared__ unsigned char s_reset[THREADS / 4];    // STAGE: the world's carry observation comes from the reset
    constexpr int W = THREADS / G;
    const int wl = (int)(threadIdx.x / G), row_f2 = e.obs_dim / 2, world_f2 = A * row_f2;
    const int64_t first_local = (int64_t)blockIdx.x * W;
    const int total_f2 = (int)min((int64_t)W, n_sub - first_local) * world_f2;  // (<= 0 for a block past the range)
    float2* blk_next = obs_next ? reinterpret_cast<float2*>(obs_next) + (env0 + first_local) * world_f2 : nullptr;
    float2* blk_carry = obs_carry ? reinterpret_cast<float2*>(obs_carry) + (env0 + first_local) * world_f2 : nullptr;
    if constexpr (STAGE) {
        write_obs<G>(e, j, is_agent, p, v, lm, reinterpret_cast<float*>(s_obs + wl * world_f2 + j * row_f2), nullptr, goal);
        if (j == 0) s_reset[wl] = do_reset ? 1 : 0;
        __syncthreads();
        for (int i = threadIdx.x; i < total_f2; i += THREADS) {
            const float2 val = s_obs[i];
            if (blk_next) blk_next[i] = val;
            if (blk_carry && !s_reset[i / world_f2]) blk_carry[i] = val;
        }
    } else {
        write_obs<G>(e, j, is_agent, p, v, lm, obs_next ? obs_next + row * e.obs_dim : nullptr,
                     (obs_carry && !do_reset) ? obs_carry + row * e.obs_dim : nullptr, goal);
    }
    if (do_reset) {
        const int64_t ep = e.episode[env];
        v = make_float2(0.f, 0.f);
        if (j < A) p = draw_entity(e, env, ep, j);
        if (j < L) {
            lm = draw_entity(e, env, ep, A + j);
            reinterpret_cast<float2*>(e.lm)[env * L + j] = lm;
        }
    }
    // the whole group takes the same branch (tr is per env); shuffles inside use the group mask
    if constexpr (STAGE) {
        if (__syncthreads_or(do_reset ? 1 : 0)) {  // (also orders the copy loop above before the staging area is reused)
            if (do_reset)
                write_obs<G>(e, j, is_agent, p, v, lm, reinterpret_cast<float*>(s_obs + wl * world_f2 + j * row_f2), nullptr,
                             e.scenario == TSM_MPE_ADVERSARY ? goal_index(e, env, e.episode[env]) : 0);
            __syncthreads();
            if (blk_carry)
                for (int i = threadIdx.x; i < total_f2; i += THREADS)
                    if (s_reset[i / world_f2]) blk_carry[i] = s_obs[i];
        }
    } else if (do_reset) {
        write_obs<G>(e, j, is_agent, p, v, lm, obs_carry ? obs_carry + row * e.obs_dim : nullptr, nullptr,
                     e.scenario == TSM_MPE_ADVERSARY ? goal_index(e, env, e.episode[env]) : 0);
    }
    if (is_agent) {
        reinterpret_cast<float2*>(e.pos)[row] = p;
        reinterpret_cast<float2*>(e.vel)[row] = v;
    }
    if (live && j == 0) {
        e.steps[env] = do_reset ? 0 : st;
        if (do_reset) e.episode[env] += 1;
    }
    if constexpr (TAIL) {
        // ---- ring counters of sub-buffer `row` (one per row, buffer_ids = arange) ----
        bool dn = false;
        double ret_done = 0.0;
        int64_t len_done = 0;
        if (is_agent) {
            const tsm_buf& b = t.b;
            const int64_t off = row * b.size, cur = b.ins[row];
            const int64_t nsz = min(b.sz[row] + 1, b.size), nins = (cur + 1) % b.size;
            dn = tr;  // no termination in these scenarios
            const double eret = b.ep_ret[row] + (double)r;
            const int64_t elen = b.ep_len[row] + 1, estart = b.ep_start[row];
            const int64_t slot = cur * b.n_bufs + row;  // time-major payload
            b.act[slot] = a;
            b.rew[slot] = r;
            b.terminated[slot] = 0;
            b.truncated[slot] = tr ? 1 : 0;
            b.done[slot] = dn ? 1 : 0;
            ret_done = dn ? eret : 0.0;
            len_done = dn ? elen : 0;
            if (t.ep_ret_out) t.ep_ret_out[row] = ret_done;
            if (t.ep_len_out) t.ep_len_out[row] = len_done;
            b.ins[row] = nins;
            b.sz[row] = nsz;
            b.last_index[row] = cur + off;
            b.ep_ret[row] = dn ? 0.0 : eret;
            b.ep_len[row] = dn ? 0 : elen;
            b.ep_start[row] = dn ? nins : estart;  // buffer_base.py:408
        }
        // ---- episode statistics: only on steps where some episode of this block ended ----
        if (__syncthreads_or(dn ? 1 : 0)) {
            __shared__ double s_part[THREADS / 32][9];
            const double dl = (double)len_done;
            double v[9] = {dn ? 1.0 : 0.0, ret_done, ret_done * ret_done, dn ? ret_done : INFINITY, dn ? ret_done : -INFINITY,
                           dl, dl * dl, dn ? dl : INFINITY, dn ? dl : -INFINITY};
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const double o = __shfl_down_sync(0xffffffffu, v[k], d);
                    v[k] = (k == 3 || k == 7) ? fmin(v[k], o) : ((k == 4 || k == 8) ? fmax(v[k], o) : v[k] + o);
                }
            }
            const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
            if (lane == 0)
#pragma unroll
                for (int k = 0; k < 9; ++k) s_part[w][k] = v[k];
            if (t.list_count) {  // one counter bump per warp, then every finished row writes its own entry
                const unsigned m = __ballot_sync(0xffffffffu, dn);
                if (m) {
                    int64_t base = 0;
                    if (lane == __ffs(m) - 1) base = (int64_t)atomicAdd(reinterpret_cast<unsigned long long*>(t.list_count), (unsigned long long)__popc(m));
                    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                    const int64_t at = base + __popc(m & ((1u << lane) - 1u));
                    if (dn && at < t.list_cap) { t.list_ret[at] = ret_done; t.list_len[at] = len_done; }
                }
            }
            __syncthreads();
            if (threadIdx.x < 9) {
                const int k = threadIdx.x;
                double x = s_part[0][k];
                for (int ww = 1; ww < THREADS / 32; ++ww) {
                    const double o = s_part[ww][k];
                    x = (k == 3 || k == 7) ? fmin(x, o) : ((k == 4 || k == 8) ? fmax(x, o) : x + o);
                }
                if (k == 3 || k == 7) atomic_min_f64(t.acc + k, x);
                else if (k == 4 || k == 8) atomic_max_f64(t.acc + k, x);
                else atomicAdd(t.acc + k, x);
            }
        }
    }
}

}  // namespace mpe
}  // namespace tsm

using namespace tsm;
using namespace tsm::mpe;

static int check_env(const tsm_mpe* e, const char* fn) {
    if (!e) return fail(TSM_ERR_INVALID, "%s: env is NULL", fn);
    if (e->scenario != TSM_MPE_SPREAD && e->scenario != TSM_MPE_TAG && e->scenario != TSM_MPE_ADVERSARY)
        return fail(TSM_ERR_INVALID, "%s: unknown scenario", fn);
    if (e->n_envs <= 0 || e->n_agents <= 0 || e->n_agents > 32 || e->n_landmarks < 0 || e->n_landmarks > 32)
        return fail(TSM_ERR_INVALID, "%s: need n_envs > 0, 1 <= n_agents <= 32, 0 <= n_landmarks <= 32", fn);
    if (e->scenario != TSM_MPE_SPREAD && (e->n_adversaries < 0 || e->n_adversaries > e->n_agents))
        return fail(TSM_ERR_INVALID, "%s: bad n_adversaries", fn);
    if (e->scenario == TSM_MPE_ADVERSARY && e->n_landmarks < 1) return fail(TSM_ERR_INVALID, "%s: simple_adversary needs a landmark", fn);
    const int want = e->scenario == TSM_MPE_SPREAD      ? 6 * e->n_agents
                     : e->scenario == TSM_MPE_ADVERSARY ? 2 + 2 * e->n_landmarks + 2 * (e->n_agents - 1)
                                                        : 4 + 2 * e->n_landmarks + 2 * (e->n_agents - 1) + 2 * (e->n_agents - e->n_adversaries);
    if (e->obs_dim != want) return fail(TSM_ERR_INVALID, "%s: obs_dim does not match the scenario (want %lld)", fn, want);
    if (!e->pos || !e->vel || !e->steps || !e->episode || (e->n_landmarks > 0 && !e->lm))
        return fail(TSM_ERR_INVALID, "%s: state arrays must not be NULL", fn);
    return TSM_OK;
}

static int group_width(const tsm_mpe* e) {
    int m = e->n_agents > e->n_landmarks ? e->n_agents : e->n_landmarks;
    int g = 4;
    while (g < m) g <<= 1;
    return g;
}

extern "C" int tsm_mpe_reset(const tsm_mpe* e, const int64_t* env_ids, int64_t n, const float* init_pos, float* obs_out,
                             tsm_stream_t stream) {
    int rc = check_env(e, __func__);
    if (rc) return rc;
    TSM_REQUIRE(n >= 0 && n <= e->n_envs, "0 <= n <= n_envs");
    TSM_REQUIRE(env_ids != nullptr || n == e->n_envs, "env_ids == NULL needs n == n_envs");
    if (n == 0) return TSM_OK;
    const int G = group_width(e);
    const unsigned grid = (unsigned)ceil_div(n * G, THREADS);
    cudaStream_t s = as_stream(stream);
    switch (G) {
        case 4: reset_kernel<4><<<grid, THREADS, 0, s>>>(*e, env_ids, n, init_pos, obs_out); break;
        case 8: reset_kernel<8><<<grid, THREADS, 0, s>>>(*e, env_ids, n, init_pos, obs_out); break;
        case 16: reset_kernel<16><<<grid, THREADS, 0, s>>>(*e, env_ids, n, init_pos, obs_out); break;
        default: reset_kernel<32><<<grid, THREADS, 0, s>>>(*e, env_ids, n, init_pos, obs_out); break;
    }
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

template <bool TAIL>
static int launch_step(const tsm_mpe* e, const int32_t* act, float* obs_next, float* rew, uint8_t* terminated, uint8_t* truncated,
                       int autoreset, float* obs_carry, const RolloutTail& t, int64_t env0, int64_t n_sub, cudaStream_t s,
                       const float* act_f = nullptr) {
    const int G = group_width(e);
    const unsigned grid = (unsigned)ceil_div(n_sub * G, THREADS);
    // staging area: the observations of the block's worlds; direct per-lane writes when it would not fit the default 48 KB
    const size_t stage = (size_t)(THREADS / G) * e->n_agents * e->obs_dim * sizeof(float);
    static const bool no_stage = getenv("TSM_MPE_NO_STAGE") != nullptr;
    const bool st = stage <= 48 * 1024 - 2048 && (e->obs_dim % 2) == 0 && !no_stage;
#define TSM_STEP(GG)                                                                                                                          \
    if (st) TSM_CUDA(launch_pdl(g_pdl_hint, step_kernel<GG, TAIL, true>, grid, THREADS, stage, s, *e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry, t, env0, n_sub, act_f)); \
    else TSM_CUDA(launch_pdl(g_pdl_hint, step_kernel<GG, TAIL, false>, grid, THREADS, 0, s, *e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry, t, env0, n_sub, act_f));
    switch (G) {
        case 4: TSM_STEP(4) break;
        case 8: TSM_STEP(8) break;
        case 16: TSM_STEP(16) break;
        default: TSM_STEP(32) break;
    }
#undef TSM_STEP
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_mpe_step(const tsm_mpe* e, const int32_t* act, float* obs_next, float* rew, uint8_t* terminated,
                            uint8_t* truncated, int autoreset, float* obs_carry, tsm_stream_t stream) {
    int rc = check_env(e, __func__);
    if (rc) return rc;
    TSM_REQUIRE(act && rew && terminated && truncated, "act/rew/terminated/truncated != NULL");
    return launch_step<false>(e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry, RolloutTail{}, 0, e->n_envs, as_stream(stream));
}

extern "C" int tsm_mpe_step_continuous(const tsm_mpe* e, const float* act, float* obs_next, float* rew, uint8_t* terminated,
                                       uint8_t* truncated, int autoreset, float* obs_carry, tsm_stream_t stream) {
    int rc = check_env(e, __func__);
    if (rc) return rc;
    TSM_REQUIRE(act && rew && terminated && truncated, "act/rew/terminated/truncated != NULL");
    return launch_step<false>(e, nullptr, obs_next, rew, terminated, truncated, autoreset, obs_carry, RolloutTail{}, 0, e->n_envs, as_stream(stream), act);
}

extern "C" int tsm_rollout_step(const tsm_mpe* e, const tsm_buf* buf, const int32_t* act, float* obs_next, float* rew,
                                uint8_t* terminated, uint8_t* truncated, float* obs_carry, double* ep_ret_out, int64_t* ep_len_out,
                                double* acc, double* list_ret, int64_t* list_len, int64_t* list_count, int64_t list_capacity,
                                int64_t env_begin, int64_t env_count, tsm_stream_t stream) {
    int rc = check_env(e, __func__);
    if (rc) return rc;
    TSM_REQUIRE(buf && act && rew && terminated && truncated && acc, "buf / act / rew / terminated / truncated / acc != NULL");
    TSM_REQUIRE(buf->n_bufs == e->n_envs * e->n_agents && buf->size > 0, "one sub-buffer per row (n_bufs == n_envs * n_agents)");
    TSM_REQUIRE(buf->ins && buf->sz && buf->last_index && buf->ep_len && buf->ep_start && buf->ep_ret, "counter arrays != NULL");
    TSM_REQUIRE(buf->act && buf->rew && buf->terminated && buf->truncated && buf->done, "payload arrays != NULL");
    TSM_REQUIRE(!list_count || (list_ret && list_len && list_capacity >= 0), "episode list arrays");
    RolloutTail t{*buf, ep_ret_out, ep_len_out, acc, list_ret, list_len, list_count, list_capacity};
    if (env_count <= 0) { env_begin = 0; env_count = e->n_envs; }  // the whole vector env
    TSM_REQUIRE(env_begin >= 0 && env_begin + env_count <= e->n_envs, "0 <= env_begin, env_begin + env_count <= n_envs");
    return launch_step<true>(e, act, obs_next, rew, terminated, truncated, 1, obs_carry, t, env_begin, env_count, as_stream(stream));
}
