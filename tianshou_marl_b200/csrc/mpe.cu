// mpe.cu -- K1: batched MPE simple_spread / simple_tag world step (sm_100a).
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
#include "common.cuh"
#include "philox.cuh"

namespace tsm {
namespace mpe {

constexpr float DT = 0.1f, DAMPING = 0.25f, CONTACT_FORCE = 1e2f, CONTACT_MARGIN = 1e-3f;
constexpr int THREADS = 256;

__device__ __forceinline__ float agent_size(const tsm_mpe& e, int j) {
    if (e.scenario == TSM_MPE_SPREAD) return 0.15f;
    return j < e.n_adversaries ? 0.075f : 0.05f;
}
__device__ __forceinline__ float agent_accel(const tsm_mpe& e, int j) {
    if (e.scenario == TSM_MPE_SPREAD) return 5.0f;  // simple_env._set_action default sensitivity
    return j < e.n_adversaries ? 3.0f : 4.0f;
}
__device__ __forceinline__ float agent_max_speed(const tsm_mpe& e, int j) {
    if (e.scenario == TSM_MPE_SPREAD) return INFINITY;
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

// scenario.observation for agent j (lane j); all lanes of the group must call (shuffles inside)
template <int G>
__device__ __forceinline__ void write_obs(const tsm_mpe& e, int j, bool valid, float2 p, float2 v, float2 lm,
                                          float* __restrict__ dst0, float* __restrict__ dst1) {
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
    write_obs<G>(e, j, live && j < A && obs_out != nullptr, p, v, lm,
                 obs_out ? obs_out + (gid * A + j) * e.obs_dim : nullptr, nullptr);
    if (live && j == 0) {
        e.steps[env] = 0;
        e.episode[env] += 1;
    }
}

template <int G>
__global__ void __launch_bounds__(THREADS) step_kernel(tsm_mpe e, const int32_t* __restrict__ act,
                                                      float* __restrict__ obs_next, float* __restrict__ rew,
                                                      uint8_t* __restrict__ term, uint8_t* __restrict__ trunc, int autoreset,
                                                      float* __restrict__ obs_carry) {
    const int64_t env = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / G;
    const int j = threadIdx.x % G;
    const int A = e.n_agents, L = e.n_landmarks;
    const bool live = env < e.n_envs;
    const bool is_agent = live && j < A;
    const int64_t row = env * A + j;

    float2 p = make_float2(0.f, 0.f), v = make_float2(0.f, 0.f), lm = make_float2(0.f, 0.f);
    int a = 0;
    if (is_agent) {
        p = reinterpret_cast<const float2*>(e.pos)[row];
        v = reinterpret_cast<const float2*>(e.vel)[row];
        a = act[row];
    }
    if (live && j < L) lm = reinterpret_cast<const float2*>(e.lm)[env * L + j];
    const float my_size = agent_size(e, j);

    // simple_env._set_action (discrete) + World.apply_action_force
    const float sens = agent_accel(e, j);
    float fx = (a == 1 ? -1.f : (a == 2 ? 1.f : 0.f)) * sens;
    float fy = (a == 3 ? -1.f : (a == 4 ? 1.f : 0.f)) * sens;
    // World.apply_environment_force: partners in ascending entity order (agents, then landmarks)
    for (int k = 0; k < A; ++k) {
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
    write_obs<G>(e, j, is_agent, p, v, lm, obs_next ? obs_next + row * e.obs_dim : nullptr,
                 (obs_carry && !do_reset) ? obs_carry + row * e.obs_dim : nullptr);
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
    if (do_reset)
        write_obs<G>(e, j, is_agent, p, v, lm, obs_carry ? obs_carry + row * e.obs_dim : nullptr, nullptr);
    if (is_agent) {
        reinterpret_cast<float2*>(e.pos)[row] = p;
        reinterpret_cast<float2*>(e.vel)[row] = v;
    }
    if (live && j == 0) {
        e.steps[env] = do_reset ? 0 : st;
        if (do_reset) e.episode[env] += 1;
    }
}

}  // namespace mpe
}  // namespace tsm

using namespace tsm;
using namespace tsm::mpe;

static int check_env(const tsm_mpe* e, const char* fn) {
    if (!e) return fail(TSM_ERR_INVALID, "%s: env is NULL", fn);
    if (e->scenario != TSM_MPE_SPREAD && e->scenario != TSM_MPE_TAG) return fail(TSM_ERR_INVALID, "%s: unknown scenario", fn);
    if (e->n_envs <= 0 || e->n_agents <= 0 || e->n_agents > 32 || e->n_landmarks < 0 || e->n_landmarks > 32)
        return fail(TSM_ERR_INVALID, "%s: need n_envs > 0, 1 <= n_agents <= 32, 0 <= n_landmarks <= 32", fn);
    if (e->scenario == TSM_MPE_TAG && (e->n_adversaries < 0 || e->n_adversaries > e->n_agents))
        return fail(TSM_ERR_INVALID, "%s: bad n_adversaries", fn);
    const int want = e->scenario == TSM_MPE_SPREAD ? 6 * e->n_agents
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

extern "C" int tsm_mpe_step(const tsm_mpe* e, const int32_t* act, float* obs_next, float* rew, uint8_t* terminated,
                            uint8_t* truncated, int autoreset, float* obs_carry, tsm_stream_t stream) {
    int rc = check_env(e, __func__);
    if (rc) return rc;
    TSM_REQUIRE(act && rew && terminated && truncated, "act/rew/terminated/truncated != NULL");
    const int G = group_width(e);
    const unsigned grid = (unsigned)ceil_div(e->n_envs * G, THREADS);
    cudaStream_t s = as_stream(stream);
    switch (G) {
        case 4: step_kernel<4><<<grid, THREADS, 0, s>>>(*e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry); break;
        case 8: step_kernel<8><<<grid, THREADS, 0, s>>>(*e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry); break;
        case 16: step_kernel<16><<<grid, THREADS, 0, s>>>(*e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry); break;
        default: step_kernel<32><<<grid, THREADS, 0, s>>>(*e, act, obs_next, rew, terminated, truncated, autoreset, obs_carry); break;
    }
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
