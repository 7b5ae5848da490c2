// ctde.cu -- centralized-training helpers: global state construction and the one-step TD target.
//   GlobalStateConstructor.build   tianshou/algorithm/multiagent/ctde.py:282-343 (concatenate, mean)
//   CTDEPolicy.learn               ctde.py:149-178 (value = mean over critic heads, TD target, advantage)
// Both are streaming, HBM-bound passes: one thread per output float4 / per row.
#include <algorithm>
#include "common.cuh"

namespace tsm {
namespace ctde {

// out[(e * A + a)][:] = concat_j obs[e][j][:]  -- the A*O floats of world e are already contiguous in obs, so every agent
// row of the world receives a straight copy of that block (the concatenation itself is a reshape)
__global__ void concat_kernel(const float* __restrict__ obs, int64_t n_envs, int A, int AO, float* __restrict__ out) {
    const int64_t total = n_envs * A * (int64_t)AO;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / AO;
        const int k = (int)(i - row * AO);
        out[i] = __ldg(obs + (row / A) * AO + k);
    }
}

__global__ void mean_kernel(const float* __restrict__ obs, int64_t n_envs, int A, int O, float* __restrict__ out) {
    const int64_t total = n_envs * A * (int64_t)O;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / O;
        const int k = (int)(i - row * O);
        const float* p = obs + (row / A) * (int64_t)A * O + k;
        float s = 0.f;
        for (int j = 0; j < A; ++j) s += __ldg(p + (int64_t)j * O);  // torch.stack(...).mean(0): sequential sum, then / A
        out[i] = s / (float)A;
    }
}

__global__ void td_kernel(const float* __restrict__ v_cur, const float* __restrict__ v_next, int heads, const float* __restrict__ rew,
                          const uint8_t* __restrict__ term, float gamma, int64_t n, float* __restrict__ td, float* __restrict__ adv,
                          float* __restrict__ value) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float a = 0.f, b = 0.f;
    for (int j = 0; j < heads; ++j) {
        a += v_cur[i * heads + j];
        b += v_next[i * heads + j];
    }
    a /= (float)heads;
    b /= (float)heads;
    const float t = rew[i] + gamma * b * (term[i] ? 0.f : 1.f);
    td[i] = t;
    if (adv) adv[i] = t - a;
    if (value) value[i] = a;
}

}  // namespace ctde
}  // namespace tsm

using namespace tsm;

extern "C" int tsm_global_state(const float* obs, int64_t n_envs, int n_agents, int obs_dim, int mode, float* out, tsm_stream_t stream) {
    TSM_REQUIRE(obs && out, "obs / out != NULL");
    TSM_REQUIRE(n_envs >= 0 && n_agents >= 1 && obs_dim >= 1, "n_envs >= 0, n_agents, obs_dim >= 1");
    TSM_REQUIRE(mode == 0 || mode == 1, "mode: 0 concatenate, 1 mean");
    if (n_envs == 0) return TSM_OK;
    const int64_t total = n_envs * n_agents * (int64_t)(mode == 0 ? n_agents * obs_dim : obs_dim);
    const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)sm_count() * 16);
    if (mode == 0) ctde::concat_kernel<<<grid, 256, 0, as_stream(stream)>>>(obs, n_envs, n_agents, n_agents * obs_dim, out);
    else ctde::mean_kernel<<<grid, 256, 0, as_stream(stream)>>>(obs, n_envs, n_agents, obs_dim, out);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

extern "C" int tsm_td_target(const float* v_cur, const float* v_next, int heads, const float* rew, const uint8_t* terminated,
                             float gamma, int64_t n, float* td_out, float* adv_out, float* value_out, tsm_stream_t stream) {
    TSM_REQUIRE(v_cur && v_next && rew && terminated && td_out, "v_cur / v_next / rew / terminated / td_out != NULL");
    TSM_REQUIRE(heads >= 1 && n >= 0, "heads >= 1, n >= 0");
    if (n == 0) return TSM_OK;
    ctde::td_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(v_cur, v_next, heads, rew, terminated, gamma, n, td_out,
                                                                                adv_out, value_out);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
