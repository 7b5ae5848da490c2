// ppo_math.cuh -- pieces shared by the fp32 (mlp.cu) and tcgen05 (mlp_tc.cu) actor-critic kernels:
// flat parameter layout, batch-position -> buffer-slot mapping, softmax pieces and the per-row PPO
// loss with its gradient w.r.t. the head outputs (tianshou/algorithm/modelfree/ppo.py:181-211).
#pragma once
#include <stdint.h>
#include <math_constants.h>
#include "feistel.cuh"

namespace tsm {
namespace ppo {

constexpr int OUTP = 8;  // head width padded to 8 (out_dim <= 8)

struct Offsets {
    int64_t w1, b1, w2, b2, w3, b3, total;
};
__host__ __device__ inline Offsets offsets(int in_dim, int H, int out_dim) {
    Offsets o;
    o.w1 = 0;
    o.b1 = (int64_t)H * in_dim;
    o.w2 = o.b1 + H;
    o.b2 = o.w2 + (int64_t)H * H;
    o.w3 = o.b2 + H;
    o.b3 = o.w3 + (int64_t)out_dim * H;
    o.total = o.b3 + out_dim;
    return o;
}

// where batch position p lives in the time-major device buffer
struct BatchView {
    const int64_t* indices;  // explicit env-major buffer indices per position (may be NULL)
    int64_t size, n_rows;    // ring: slots per row, rows
    int64_t start, len;      // lock-step view: row r holds `len` valid slots from ring slot `start`
    uint64_t perm_key;       // Feistel key of the minibatch permutation
    int64_t perm_n;          // domain of the permutation (0 = identity)
    FeistelPre pre;          // round keys / masks of (perm_n, perm_key), prepared by make_batch_view()
};
inline BatchView make_batch_view(const int64_t* indices, int64_t size, int64_t n_rows, int64_t start, int64_t len, uint64_t perm_key,
                                 int64_t perm_n) {
    BatchView v{};
    v.indices = indices; v.size = size; v.n_rows = n_rows; v.start = start; v.len = len; v.perm_key = perm_key; v.perm_n = perm_n;
    if (perm_n > 0) v.pre = feistel_prepare((uint64_t)perm_n, perm_key);
    return v;
}
// the batch's slot table if it can be used (32-bit slots), else NULL
inline int32_t* slots_of(const tsm_ppo_batch* b) {
    return (b->slots && b->size * b->n_rows < (int64_t)1 << 31) ? b->slots : nullptr;
}
__device__ __forceinline__ int64_t slot_of(const BatchView& b, int64_t p) {
    if (b.indices) {
        const int64_t i = b.indices[p];
        const int64_t row = i / b.size, t = i - row * b.size;
        return t * b.n_rows + row;
    }
    const int64_t q = b.perm_n > 0 ? (int64_t)feistel_permute((uint64_t)p, (uint64_t)b.perm_n, b.pre) : p;
    int64_t row, j;
    if (((uint64_t)q | (uint64_t)b.len) >> 32) {
        row = q / b.len; j = q - row * b.len;
    } else {  // 32-bit division: the 64-bit one is emulated in about 70 instructions
        const uint32_t r32 = (uint32_t)q / (uint32_t)b.len;
        row = r32; j = (uint32_t)q - r32 * (uint32_t)b.len;
    }
    int64_t t = b.start + j;
    if (t >= b.size) t -= b.size;
    return t * b.n_rows + row;
}

// softmax pieces of one row
struct Soft {
    float logp[OUTP];
    float p[OUTP];
    float ent;
    unsigned clamped;  // probs parameterisation only: bit j set where log p_j hit a clamp (no gradient flows through it)
};
// FAST: hardware exp2 / log2 / reciprocal approximations (about 2 ulp), used by the tcgen05 engine whose
// GEMM operands are already rounded to fp16; the fp32 engine keeps the IEEE library versions.
template <bool FAST> __device__ __forceinline__ float exp_t(float x) { return FAST ? __expf(x) : expf(x); }
template <bool FAST> __device__ __forceinline__ float log_t(float x) { return FAST ? __logf(x) : logf(x); }
template <bool FAST> __device__ __forceinline__ float div_t(float a, float b) { return FAST ? __fdividef(a, b) : a / b; }

// probs: the reference's DiscreteActor(softmax_output=True) + Categorical(probs=softmax(z)) (utils/net/discrete.py:85-86,
// test/discrete/test_ppo_discrete.py:97-113).  torch turns the probabilities back into logits with
// log(clamp(p, eps, 1 - eps)), eps = 2^-23 (torch.distributions.utils.probs_to_logits); log is monotonic, so that is
// clamp(log p, log eps, log(1 - eps)).  Entropy and log-probabilities use the clamped logs; a clamped entry has no gradient.
constexpr float PROBS_LOG_LO = -15.942385f;      // log(2^-23)
constexpr float PROBS_LOG_HI = -1.1920930e-7f;   // log(1 - 2^-23)
// NA > 0: the number of actions as a compile-time constant (the caller has checked n_act == NA): the same operations in the
// same order on entries 0..NA-1, none of the predication on the others
template <bool FAST = false, int NA = 0>
__device__ __forceinline__ Soft softmax_row(const float* z, int n_act_in, bool probs = false) {
    const int n_act = NA ? NA : n_act_in;
    Soft s;
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < OUTP; ++j)
        if (j < n_act) m = fmaxf(m, z[j]);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < OUTP; ++j) {
        s.p[j] = j < n_act ? exp_t<FAST>(z[j] - m) : 0.f;
        if (!NA || j < NA) sum += s.p[j];
    }
    const float lse = log_t<FAST>(sum);
    const float inv_sum = div_t<FAST>(1.f, sum);
    s.ent = 0.f;
    s.clamped = 0u;
#pragma unroll
    for (int j = 0; j < OUTP; ++j) {
        float lp = (z[j] - m) - lse;
        if (probs && j < n_act) {
            const float c = fminf(fmaxf(lp, PROBS_LOG_LO), PROBS_LOG_HI);
            if (c != lp) s.clamped |= 1u << j;
            lp = c;
        }
        s.logp[j] = j < n_act ? lp : 0.f;
        s.p[j] = j < n_act ? (FAST ? s.p[j] * inv_sum : s.p[j] / sum) : 0.f;
        if (j < n_act) s.ent -= s.p[j] * s.logp[j];
    }
    return s;
}

struct PpoCfg {
    float eps_clip, dual_clip, vf_coef, ent_coef;
    int value_clip, adv_norm;
    int plain_pg;     // 1: policy term = -logp(act) * adv (one-step actor-critic) instead of the clipped surrogate;
                      // 2: the reference's literal CTDEPolicy.learn, where logp [n] * adv [n,1] broadcasts to [n,n]
                      //    (ctde.py:181-183): every row is weighted by the minibatch MEAN advantage (mb_stats[0])
    int value_heads;  // > 1: the value is the mean of that many head outputs from value_index on (ctde.py:153-156)
    int probs_param;  // the net's softmax_output flag (tsm_net.probs_param): Categorical(probs = softmax(logits))
};


// loss terms of one row and d(mean loss)/d(head outputs) (dz), ppo.py:181-211
struct RowLoss {
    float clip, vf, ent;
};
// SIMPLE: the caller has checked (once, on uniform values) that cfg.probs_param == 0, cfg.value_heads <= 1 and
// cfg.plain_pg == 0 -- plain PPO on a logits head with one value output.  The same arithmetic with those three folded in
// as constants: the loss stage of the tcgen05 engine is a dependent chain of one warp per scheduler, so its length is
// its instruction count (570 -> 300 SASS instructions per row).
// NA > 0 (with SIMPLE): additionally n_act == NA and value_index == NA, checked by the caller.
template <bool FAST = false, bool SIMPLE = false, int NA = 0>
__device__ __forceinline__ RowLoss row_loss(const float (&z)[OUTP], int n_act_in, int value_index_in, int act, float adv, float ret,
                                            float logp_old, float v_old, float mean, float rstd, const PpoCfg& cfg_in, float inv_n,
                                            float (&dz)[OUTP]) {
    static_assert(NA == 0 || (SIMPLE && NA < OUTP), "NA needs SIMPLE and room for the value output");
    const int n_act = NA ? NA : n_act_in, value_index = NA ? NA : value_index_in;
    PpoCfg cfg = cfg_in;
    if (SIMPLE) { cfg.probs_param = 0; cfg.value_heads = 1; cfg.plain_pg = 0; }
    RowLoss out{0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < OUTP; ++j) dz[j] = 0.f;
    if (n_act > 0) {
        const Soft sm = softmax_row<FAST, NA>(z, n_act, cfg.probs_param != 0);
        float lp = 0.f;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j == act) lp = sm.logp[j];
        float advn = adv;
        if (cfg.adv_norm) advn = (advn - mean) * rstd;
        if (cfg.plain_pg == 2) advn = mean;
        float li, dlp;  // loss term of this row and d loss / d logp(act)
        if (cfg.plain_pg) {
            li = -lp * advn;
            dlp = -advn * inv_n;
        } else {
            const float ratio = exp_t<FAST>(lp - logp_old);
            const float lo = 1.f - cfg.eps_clip, hi = 1.f + cfg.eps_clip;
            const float s1 = ratio * advn, s2 = fminf(fmaxf(ratio, lo), hi) * advn;
            const bool inside = ratio >= lo && ratio <= hi;
            const float clip1 = fminf(s1, s2);
            float g_ratio = (inside || s1 < s2) ? advn : 0.f;  // d min(s1,s2) / d ratio
            li = -clip1;
            if (cfg.dual_clip > 0.f && advn < 0.f) {
                const float c2 = cfg.dual_clip * advn;
                if (c2 > clip1) { li = -c2; g_ratio = 0.f; }
            }
            dlp = -g_ratio * ratio * inv_n;
        }
        out.clip = li;
        out.ent = sm.ent;
        const float ec = cfg.ent_coef * inv_n;
        if (sm.clamped == 0u) {
#pragma unroll
            for (int j = 0; j < OUTP; ++j)
                if (j < n_act) dz[j] = dlp * ((j == act ? 1.f : 0.f) - sm.p[j]) + ec * sm.p[j] * (sm.logp[j] + sm.ent);
        } else {
            // some log p_j sit on a clamp: d logp(act) = 0 if act is clamped; d(-ent)/dz_k = p_k (L_k + ent + u_k - U) with
            // u_k = [k unclamped], U = sum of the unclamped p_j
            float U = 0.f;
#pragma unroll
            for (int j = 0; j < OUTP; ++j)
                if (j < n_act && !((sm.clamped >> j) & 1u)) U += sm.p[j];
            const float dl = ((sm.clamped >> act) & 1u) ? 0.f : dlp;
#pragma unroll
            for (int j = 0; j < OUTP; ++j)
                if (j < n_act) {
                    const float u = ((sm.clamped >> j) & 1u) ? 0.f : 1.f;
                    dz[j] = dl * ((j == act ? 1.f : 0.f) - sm.p[j]) + ec * sm.p[j] * (sm.logp[j] + sm.ent + u - U);
                }
        }
    }
    if (value_index >= 0) {
        const int nh = cfg.value_heads > 1 ? cfg.value_heads : 1;
        float v = 0.f;
        if (SIMPLE) {
#pragma unroll
            for (int j = 0; j < OUTP; ++j)
                if (j == value_index) v = z[j];
        } else {
#pragma unroll
            for (int j = 0; j < OUTP; ++j)
                if (j >= value_index && j < value_index + nh) v += z[j];
            v /= (float)nh;
        }
        float vf = (ret - v) * (ret - v), dv = 2.f * (v - ret);
        if (cfg.value_clip) {
            const float d = v - v_old;
            const float vc = v_old + fminf(fmaxf(d, -cfg.eps_clip), cfg.eps_clip);
            const float vf2 = (ret - vc) * (ret - vc);
            if (vf2 > vf) {
                vf = vf2;
                dv = (d >= -cfg.eps_clip && d <= cfg.eps_clip) ? 2.f * (vc - ret) : 0.f;
            }
        }
        out.vf = vf;
        const float dvs = SIMPLE ? cfg.vf_coef * inv_n * dv : cfg.vf_coef * inv_n * dv / (float)nh;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (SIMPLE ? j == value_index : (j >= value_index && j < value_index + nh)) dz[j] = dvs;
    }
    return out;
}

// The same row for a diagonal Gaussian policy with state-independent scale: the reference's ContinuousActorProbabilistic(unbounded
// = True, conditioned_sigma = False) feeding Independent(Normal(mu, sigma), 1) (utils/net/continuous.py:229-247,
// test/continuous/test_ppo.py:104-110).  z[0..n_act) = mu, log_sigma[0..n_act) = the actor's sigma_param, a[0..n_act) = the
// stored action (before map_action).  torch.distributions.Normal: log_prob = -(a - mu)^2 / (2 sigma^2) - log sigma - log sqrt(2 pi),
// entropy = 0.5 + 0.5 log(2 pi) + log sigma, both summed over the action dimensions.  dz as in row_loss; dls[j] += d(mean loss) /
// d log_sigma[j] of this row (the caller reduces it over the rows of the minibatch).
__device__ __forceinline__ RowLoss row_loss_gauss(const float (&z)[OUTP], int n_act, int value_index, const float* __restrict__ a,
                                                  const float* __restrict__ log_sigma, float adv, float ret, float logp_old, float v_old,
                                                  float mean, float rstd, const PpoCfg& cfg, float inv_n, float (&dz)[OUTP], float (&dls)[OUTP],
                                                  float mu_bound = 0.f, bool cond_sigma = false) {
    RowLoss out{0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < OUTP; ++j) dz[j] = 0.f;
    float lp = 0.f, ent = 0.f;
    float dmu[OUTP], dsg[OUTP];  // d logp / d mu_j, d logp / d log_sigma_j
#pragma unroll
    for (int j = 0; j < OUTP; ++j) {
        dmu[j] = dsg[j] = 0.f;
        if (j < n_act) {
            // unbounded=False (continuous.py:239-241): mu = max_action * tanh(z), d mu / d z = max_action * (1 - tanh^2 z)
            const float th = mu_bound > 0.f ? tanhf(z[j]) : 0.f;
            const float mu = mu_bound > 0.f ? mu_bound * th : z[j], dmu_dz = mu_bound > 0.f ? mu_bound * (1.f - th * th) : 1.f;
            // conditioned_sigma=True (continuous.py:242): log sigma = clamp(head output n_act + j, -20, 2), per row
            float ls_raw = 0.f;
            if (cond_sigma) {
#pragma unroll
                for (int k = 0; k < OUTP; ++k)
                    if (k == n_act + j) ls_raw = z[k];
            }
            const float ls = cond_sigma ? fminf(fmaxf(ls_raw, -20.f), 2.f) : log_sigma[j];
            const float sg = expf(ls), var = sg * sg, d = a[j] - mu;
            lp += -(d * d) / (2.f * var) - ls - 0.9189385332046727f;  // log sqrt(2 pi)
            ent += 0.5f + 0.9189385332046727f + ls;                   // 0.5 + 0.5 log(2 pi) + log sigma
            dmu[j] = d / var * dmu_dz;
            dsg[j] = d * d / var - 1.f;
            if (cond_sigma && (ls_raw < -20.f || ls_raw > 2.f)) dsg[j] = CUDART_NAN_F;  // marks "clamped": no gradient (below)
        }
    }
    float advn = adv;
    if (cfg.adv_norm) advn = (advn - mean) * rstd;
    float li, dlp;
    if (cfg.plain_pg) {
        li = -lp * advn;
        dlp = -advn * inv_n;
    } else {
        const float ratio = expf(lp - logp_old);
        const float lo = 1.f - cfg.eps_clip, hi = 1.f + cfg.eps_clip;
        const float s1 = ratio * advn, s2 = fminf(fmaxf(ratio, lo), hi) * advn;
        const bool inside = ratio >= lo && ratio <= hi;
        const float clip1 = fminf(s1, s2);
        float g_ratio = (inside || s1 < s2) ? advn : 0.f;
        li = -clip1;
        if (cfg.dual_clip > 0.f && advn < 0.f) {
            const float c2 = cfg.dual_clip * advn;
            if (c2 > clip1) { li = -c2; g_ratio = 0.f; }
        }
        dlp = -g_ratio * ratio * inv_n;
    }
    out.clip = li;
    out.ent = ent;
    const float ec = cfg.ent_coef * inv_n;
#pragma unroll
    for (int j = 0; j < OUTP; ++j)
        if (j < n_act) {
            dz[j] = dlp * dmu[j];
            if (!cond_sigma) {
                dls[j] += dlp * dsg[j] - ec;
            } else if (dsg[j] == dsg[j]) {  // inside the clamp: d loss / d (head output n_act + j); the entropy term is per row here
#pragma unroll
                for (int k = 0; k < OUTP; ++k)
                    if (k == n_act + j) dz[k] = dlp * dsg[j] - ec;
            }
        }
    if (value_index >= 0) {
        float v = 0.f;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j == value_index) v = z[j];
        float vf = (ret - v) * (ret - v), dv = 2.f * (v - ret);
        if (cfg.value_clip) {
            const float d = v - v_old;
            const float vc = v_old + fminf(fmaxf(d, -cfg.eps_clip), cfg.eps_clip);
            const float vf2 = (ret - vc) * (ret - vc);
            if (vf2 > vf) {
                vf = vf2;
                dv = (d >= -cfg.eps_clip && d <= cfg.eps_clip) ? 2.f * (vc - ret) : 0.f;
            }
        }
        out.vf = vf;
        const float dvs = cfg.vf_coef * inv_n * dv;
#pragma unroll
        for (int j = 0; j < OUTP; ++j)
            if (j == value_index) dz[j] = dvs;
    }
    return out;
}

}  // namespace ppo
}  // namespace tsm
