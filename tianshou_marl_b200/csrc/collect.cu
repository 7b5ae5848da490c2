// collect.cu -- the lock-step rollout loop as ONE C call (launches only; no kernel of its own).
//
// Replaces the per-step Python loop of Collector._collect (reference tianshou/data/collector.py:770-1098)
// for the case every environment advances every step: per vector step
//   policy forward + sample -> act / logp / value written straight into buffer slot t   (collector.py:704-767)
//   batched MPE step        -> obs_next / rew / flags into slot t, next obs into slot t+1 (venvs.py:237-322)
//   + ring counters, episode return / length (manager.py:131-193) and episode statistics (collector.py:983-1024) in
//     the same launch (tsm_rollout_step)
// The host loop costs ~3 us per launch instead of ~20 us of interpreter time, which is what bounds a rollout of
// 2 small kernels per step.
#include <algorithm>
#include <cstdlib>
#include "common.cuh"

using namespace tsm;

extern "C" int tsm_collect_lockstep(const tsm_net* net, const tsm_mpe* env, const tsm_buf* buf, const tsm_rollout* ro, int64_t n_steps,
                                    int mode, uint64_t seed, uint64_t counter0, int64_t lockstep_len, int n_random_act,
                                    int64_t* slot_out, int64_t* lockstep_len_out, int* obs_in_carry_out, tsm_stream_t stream) {
    TSM_REQUIRE(env && buf && ro, "env / buf / rollout != NULL");
    TSM_REQUIRE(n_steps >= 0 && ro->size > 0 && ro->slot >= 0 && ro->slot < ro->size, "n_steps >= 0, 0 <= slot < size");
    TSM_REQUIRE(ro->obs && ro->act && ro->rew && ro->terminated && ro->truncated && ro->obs_next && ro->carry, "payload pointers != NULL");
    TSM_REQUIRE(ro->ep_ret && ro->ep_len && ro->acc, "episode scratch != NULL");
    TSM_REQUIRE(net || n_random_act > 0, "a net or n_random_act > 0");
    TSM_REQUIRE(!net || (mode == 1 || mode == 2), "mode: 1 sample, 2 argmax");
    const int64_t R = ro->n_rows, O = ro->obs_dim, size = ro->size;
    TSM_REQUIRE(R > 0 && O > 0, "n_rows, obs_dim > 0");
    int64_t slot = ro->slot, len = lockstep_len;
    bool in_carry = false;  // the observation to act on sits in ro->carry instead of obs[slot]
    cudaStream_t s = as_stream(stream);
    // Environments never interact, so the rollout of one half of them is independent of the other half's for ALL steps:
    // with a policy net the environments are split into `parts` ranges, each range runs its own forward -> step chain on
    // its own stream (forked from / joined into the caller's stream), and the GPU overlaps one range's latency-bound
    // tensor-core forward with the other's CUDA-core world step.  Sampling keys use the global row index
    // (tsm_net_forward_range), so the actions are those of the unsplit loop.
    static const bool pdl_on = !getenv("TSM_COLLECT_NO_PDL");
    const bool pdl = pdl_on && net && net->engine == 1;
    static const int want_parts = getenv("TSM_COLLECT_STREAMS") ? atoi(getenv("TSM_COLLECT_STREAMS")) : 2;
    constexpr int MAX_PARTS = 4;
    const int64_t E = env->n_envs, A = env->n_agents;
    int parts = (net && E >= 8192) ? std::min(std::max(want_parts, 1), MAX_PARTS) : 1;
    static cudaStream_t part_stream[MAX_PARTS] = {};
    static cudaEvent_t fork_ev = nullptr, join_ev[MAX_PARTS] = {};
    static int ev_device = -1;
    if (parts > 1) {
        int dev = 0;
        TSM_CUDA(cudaGetDevice(&dev));
        if (ev_device != dev) {  // (one device per process)
            for (int p = 0; p < MAX_PARTS; ++p) {
                TSM_CUDA(cudaStreamCreateWithFlags(&part_stream[p], cudaStreamNonBlocking));
                TSM_CUDA(cudaEventCreateWithFlags(&join_ev[p], cudaEventDisableTiming));
            }
            TSM_CUDA(cudaEventCreateWithFlags(&fork_ev, cudaEventDisableTiming));
            ev_device = dev;
        }
        TSM_CUDA(cudaEventRecord(fork_ev, s));
        for (int p = 0; p < parts; ++p) TSM_CUDA(cudaStreamWaitEvent(part_stream[p], fork_ev, 0));
    }
    for (int64_t k = 0; k < n_steps; ++k) {
        float* obs_slot = ro->obs + slot * R * O;
        int32_t* act_slot = ro->act + slot * R;
        // the next observation goes straight into the next slot unless this is the last step or the ring would wrap onto live data
        const int64_t nxt = (slot + 1) % size;
        const bool to_carry = (k == n_steps - 1) || (len + 1 >= size);
        float* carry = to_carry ? ro->carry : ro->obs + nxt * R * O;
        for (int p = 0; p < parts; ++p) {
            // engine 2 packs each range's rows into 128-row tiles of the net's scratch image: ranges start on a tile boundary
            const int64_t al = (net && net->engine == 2) ? 128 : 1;
            const int64_t e0 = p == 0 ? 0 : E * p / parts / al * al, e1 = p == parts - 1 ? E : E * (p + 1) / parts / al * al;
            const int64_t r0 = e0 * A, nr = (e1 - e0) * A;
            if (nr <= 0) continue;
            cudaStream_t ps = parts > 1 ? part_stream[p] : s;
            tsm_stream_t pst = parts > 1 ? reinterpret_cast<tsm_stream_t>(ps) : stream;
            if (in_carry)
                TSM_CUDA(cudaMemcpyAsync(obs_slot + r0 * O, ro->carry + r0 * O, sizeof(float) * nr * O, cudaMemcpyDeviceToDevice, ps));
            int rc;
            if (net) {
                // engine 1's forward starts under the tail of the previous world step of this range (k > 0: the predecessor in
                // this stream is that launch; before it comes whatever the caller enqueued, e.g. the optimizer's image write)
                g_pdl_hint = pdl && k > 0 && !in_carry;
                rc = tsm_net_forward_range(net, obs_slot, r0, nr, mode, seed, counter0 + (uint64_t)k, act_slot,
                                           ro->logp ? ro->logp + slot * R : nullptr,
                                           (ro->value && net->value_index >= 0) ? ro->value + slot * R : nullptr, pst);
            } else {
                rc = tsm_random_actions(act_slot, R, n_random_act, seed, counter0 + (uint64_t)k, stream);
                if (!rc && ro->logp) TSM_CUDA(cudaMemsetAsync(ro->logp + slot * R, 0, sizeof(float) * R, s));
                if (!rc && ro->value) TSM_CUDA(cudaMemsetAsync(ro->value + slot * R, 0, sizeof(float) * R, s));
            }
            g_pdl_hint = false;
            if (rc) return rc;
            // world step + ring counters + episode statistics: one launch (its loads of the world state run under the forward's tail)
            g_pdl_hint = pdl && net != nullptr;
            rc = tsm_rollout_step(env, buf, act_slot, ro->obs_next + slot * R * O, ro->rew + slot * R, ro->terminated + slot * R,
                                  ro->truncated + slot * R, carry, ro->ep_ret, ro->ep_len, ro->acc, ro->list_ret, ro->list_len,
                                  ro->list_count, ro->list_cap, e0, e1 - e0, pst);
            g_pdl_hint = false;
            if (rc) return rc;
        }
        len = len + 1 < size ? len + 1 : size;
        slot = nxt;
        in_carry = to_carry;
    }
    if (parts > 1) {
        for (int p = 0; p < parts; ++p) {
            TSM_CUDA(cudaEventRecord(join_ev[p], part_stream[p]));
            TSM_CUDA(cudaStreamWaitEvent(s, join_ev[p], 0));
        }
    }
    if (slot_out) *slot_out = slot;
    if (lockstep_len_out) *lockstep_len_out = len;
    if (obs_in_carry_out) *obs_in_carry_out = in_carry ? 1 : 0;
    return TSM_OK;
}

// The same loop for GROUPED policies (FlexibleMultiAgentPolicyManager.forward inside the rollout, reference
// multiagent/flexible_policy.py:189-231 -> marl.py:108-185): with rows flattened as row = world * A + agent the rows of a policy
// group are a fixed strided set, so every vector step is one fused forward launch PER POLICY GROUP (row gather + scatter of
// act / logp / value straight into buffer slot t; sampling key = the global row, seed and counter of the group's policy) and one
// tsm_rollout_step over all worlds.  No Python per step.
extern "C" int tsm_collect_lockstep_groups(const tsm_net* const* nets, const int64_t* const* rows, const int64_t* n_rows, const uint64_t* seeds,
                                           const uint64_t* counters0, int n_groups, const tsm_mpe* env, const tsm_buf* buf, const tsm_rollout* ro,
                                           int64_t n_steps, int mode, int64_t lockstep_len, int64_t* slot_out, int64_t* lockstep_len_out,
                                           int* obs_in_carry_out, tsm_stream_t stream) {
    TSM_REQUIRE(nets && rows && n_rows && seeds && counters0 && n_groups >= 1 && n_groups <= 64, "group arrays != NULL, 1 <= n_groups <= 64");
    TSM_REQUIRE(env && buf && ro, "env / buf / rollout != NULL");
    TSM_REQUIRE(n_steps >= 0 && ro->size > 0 && ro->slot >= 0 && ro->slot < ro->size, "n_steps >= 0, 0 <= slot < size");
    TSM_REQUIRE(ro->obs && ro->act && ro->rew && ro->terminated && ro->truncated && ro->obs_next && ro->carry, "payload pointers != NULL");
    TSM_REQUIRE(ro->ep_ret && ro->ep_len && ro->acc, "episode scratch != NULL");
    TSM_REQUIRE(mode == 1 || mode == 2, "mode: 1 sample, 2 argmax");
    const int64_t R = ro->n_rows, O = ro->obs_dim, size = ro->size;
    TSM_REQUIRE(R > 0 && O > 0, "n_rows, obs_dim > 0");
    for (int g = 0; g < n_groups; ++g) TSM_REQUIRE(nets[g] && rows[g] && n_rows[g] > 0, "every group needs a net and a row list");
    bool pdl = !getenv("TSM_COLLECT_NO_PDL");
    for (int g = 0; g < n_groups; ++g) pdl = pdl && nets[g]->engine == 1;  // (engine 1's forward is the kernel that knows the protocol)
    int64_t slot = ro->slot, len = lockstep_len;
    bool in_carry = false;
    cudaStream_t s = as_stream(stream);
    for (int64_t k = 0; k < n_steps; ++k) {
        float* obs_slot = ro->obs + slot * R * O;
        int32_t* act_slot = ro->act + slot * R;
        const int64_t nxt = (slot + 1) % size;
        const bool to_carry = (k == n_steps - 1) || (len + 1 >= size);
        float* carry = to_carry ? ro->carry : ro->obs + nxt * R * O;
        if (in_carry) TSM_CUDA(cudaMemcpyAsync(obs_slot, ro->carry, sizeof(float) * R * O, cudaMemcpyDeviceToDevice, s));
        for (int g = 0; g < n_groups; ++g) {
            // the chain forward(group 0) -> ... -> forward(group n-1) -> world step runs under programmatic dependent launch like
            // the single-policy loop above (every kernel waits for its predecessor before it touches what that one wrote)
            g_pdl_hint = pdl && (g > 0 || (k > 0 && !in_carry));
            int rc = tsm_net_forward(nets[g], obs_slot, rows[g], n_rows[g], mode, 1, seeds[g], counters0[g] + (uint64_t)k, nullptr, act_slot,
                                     ro->logp ? ro->logp + slot * R : nullptr,
                                     (ro->value && nets[g]->value_index >= 0) ? ro->value + slot * R : nullptr, nullptr, stream);
            g_pdl_hint = false;
            if (rc) return rc;
        }
        g_pdl_hint = pdl;
        int rc = tsm_rollout_step(env, buf, act_slot, ro->obs_next + slot * R * O, ro->rew + slot * R, ro->terminated + slot * R,
                                  ro->truncated + slot * R, carry, ro->ep_ret, ro->ep_len, ro->acc, ro->list_ret, ro->list_len, ro->list_count,
                                  ro->list_cap, 0, env->n_envs, stream);
        g_pdl_hint = false;
        if (rc) return rc;
        len = len + 1 < size ? len + 1 : size;
        slot = nxt;
        in_carry = to_carry;
    }
    if (slot_out) *slot_out = slot;
    if (lockstep_len_out) *lockstep_len_out = len;
    if (obs_in_carry_out) *obs_in_carry_out = in_carry ? 1 : 0;
    return TSM_OK;
}
