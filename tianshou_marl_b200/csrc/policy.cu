// policy.cu -- action sampling entry points of the rollout (sm_100a).
//
// tsm_random_actions replaces the `random=True` branch of the collector
//   tianshou/data/collector.py:711-726  (act = [action_space[i].sample() for i in ready_env_ids])
// with one Philox draw per row (the reference draws from numpy's per-space generator, which a
// GPU cannot reproduce; oracle/philox.py restates this stream bit-for-bit).
#include "common.cuh"
#include "philox.cuh"

namespace tsm {
namespace policy {

constexpr int THREADS = 256;

// counter (row_lo, row_hi, step_lo, step_hi), key (seed_lo, seed_hi ^ STREAM_ACT); action = mulhi(x, n_act)
__global__ void __launch_bounds__(THREADS) random_actions_kernel(int32_t* __restrict__ act, int64_t n, uint32_t n_act,
                                                                uint64_t seed, uint64_t step) {
    const int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x;
    if (i >= n) return;
    const Philox4 r = philox4x32((uint32_t)i, (uint32_t)((uint64_t)i >> 32), (uint32_t)step, (uint32_t)(step >> 32),
                                 (uint32_t)seed, (uint32_t)(seed >> 32) ^ STREAM_ACT);
    act[i] = (int32_t)__umulhi(r.x, n_act);
}

}  // namespace policy
}  // namespace tsm

using namespace tsm;

extern "C" int tsm_random_actions(int32_t* act, int64_t n, int n_act, uint64_t seed, uint64_t step, tsm_stream_t stream) {
    TSM_REQUIRE(n >= 0 && n_act > 0, "n >= 0, n_act > 0");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(act != nullptr, "act != NULL");
    policy::random_actions_kernel<<<(unsigned)ceil_div(n, policy::THREADS), policy::THREADS, 0, as_stream(stream)>>>(
        act, n, (uint32_t)n_act, seed, step);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}

// Policy.map_action for Box action spaces (reference tianshou/algorithm/algorithm_base.py:255-288): bounding (clip to [-1, 1] or
// tanh) then scaling low + (high - low) * (act + 1) / 2, elementwise over [n][dim] raw actions; the mapped action only goes to the
// environment, the buffer keeps the raw one.  *out_of_range is set when scaling meets a raw action outside [-1, 1] (the
// reference asserts there).
namespace tsm {
namespace policy {
__global__ void __launch_bounds__(THREADS) map_action_kernel(const float* __restrict__ act, int64_t n, int dim, int bound, int scaling,
                                                            const float* __restrict__ low, const float* __restrict__ high,
                                                            float* __restrict__ out, int* __restrict__ out_of_range) {
    const int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x;
    if (i >= n * dim) return;
    float a = act[i];
    if (bound == 1) a = fminf(fmaxf(a, -1.0f), 1.0f);
    else if (bound == 2) a = tanhf(a);
    if (scaling) {
        if (!(a >= -1.0f && a <= 1.0f) && out_of_range) *out_of_range = 1;
        const int d = (int)(i % dim);
        const float lo = low[d], hi = high[d];
        a = lo + (hi - lo) * (a + 1.0f) / 2.0f;
    }
    out[i] = a;
}
}  // namespace policy
}  // namespace tsm

extern "C" int tsm_map_action(const float* act, int64_t n, int dim, int bound_method, int scaling, const float* low, const float* high,
                              float* out, int* out_of_range, tsm_stream_t stream) {
    TSM_REQUIRE(n >= 0 && dim >= 1 && bound_method >= 0 && bound_method <= 2, "n >= 0, dim >= 1, bound_method in {0 none, 1 clip, 2 tanh}");
    if (n == 0) return TSM_OK;
    TSM_REQUIRE(act && out && (!scaling || (low && high)), "act / out != NULL; scaling needs low and high");
    policy::map_action_kernel<<<(unsigned)ceil_div(n * dim, policy::THREADS), policy::THREADS, 0, as_stream(stream)>>>(
        act, n, dim, bound_method, scaling, low, high, out, out_of_range);
    TSM_LAUNCH_CHECK();
    return TSM_OK;
}
