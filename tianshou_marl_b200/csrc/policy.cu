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
