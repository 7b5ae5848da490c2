/* Oracle (C part): CPU restatement of the serial/hot loops of the reference path.
 * TEST INFRASTRUCTURE ONLY -- used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs.  Never linked into the product library.
 *
 * orc_gae           tianshou/algorithm/algorithm_base.py:1126-1134  (_gae, numba, float64)
 * orc_buf_add       tianshou/data/buffer/buffer_base.py:373-410 + manager.py:162-177
 * orc_mpe_spread_step   pettingzoo 1.24.2 mpe core.World.step + simple_spread scenario
 *                       (third-party; PARITY UNPINNED, see oracle/mpe.py)
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

void orc_gae(const double* v_s, const double* v_s_, const double* rew, const uint8_t* end_flag,
             int64_t n, double gamma, double lam, double* out) {
    double gae = 0.0;
    const double gl = gamma * lam;
    for (int64_t i = n - 1; i >= 0; --i) {
        double delta = rew[i] + v_s_[i] * gamma - v_s[i];
        double discount = (1.0 - (double)end_flag[i]) * gl;
        gae = delta + discount * gae;
        out[i] = gae;
    }
}

/* Same recurrence with float32 inputs as the buffer holds them (v_s, v_s_ float32; rew
 * promoted to double first, exactly what numpy/numba promotion does). */
void orc_gae_f32in(const float* v_s, const float* v_s_, const float* rew, const uint8_t* end_flag,
                   int64_t n, double gamma, double lam, double* out) {
    double gae = 0.0;
    const double gl = gamma * lam;
    for (int64_t i = n - 1; i >= 0; --i) {
        double delta = (double)rew[i] + (double)v_s_[i] * gamma - (double)v_s[i];
        double discount = (1.0 - (double)end_flag[i]) * gl;
        gae = delta + discount * gae;
        out[i] = gae;
    }
}

/* Per-sub-buffer add state machine, one call = one vector step over n rows. */
void orc_buf_add(int64_t* ins, int64_t* sz, double* ep_ret, int64_t* ep_len, int64_t* ep_start,
                 int64_t* last_index, int64_t* lengths, int64_t size,
                 const int64_t* buffer_ids, const double* rew, const uint8_t* done, int64_t n,
                 int64_t* ptr_out, double* ep_ret_out, int64_t* ep_len_out, int64_t* ep_start_out) {
    for (int64_t k = 0; k < n; ++k) {
        int64_t b = buffer_ids ? buffer_ids[k] : k;
        int64_t off = b * size;
        int64_t cur = ins[b];
        sz[b] = sz[b] + 1 < size ? sz[b] + 1 : size;
        ins[b] = (ins[b] + 1) % size;
        ep_ret[b] += rew[k];
        ep_len[b] += 1;
        if (done[k]) { ep_ret_out[k] = ep_ret[b]; ep_len_out[k] = ep_len[b]; }
        else { ep_ret_out[k] = 0.0; ep_len_out[k] = 0; }
        ep_start_out[k] = ep_start[b] + off;
        if (done[k]) { ep_ret[b] = 0.0; ep_len[b] = 0; ep_start[b] = ins[b]; }
        ptr_out[k] = cur + off;
        last_index[b] = cur + off;
        lengths[b] = sz[b];
    }
}

/* simple_spread world step for envs [0, E): state float64, obs float32.
 * pos: (E, 2N, 2) agents then landmarks; vel: (E, N, 2); act: (E, N) int32.
 * Outputs obs (E, N, 6N) f32, rew (E, N) f64. steps/truncation handled by caller. */
void orc_mpe_spread_step(double* pos, double* vel, const int32_t* act, int64_t E, int N,
                         double local_ratio, int pos_first, float* obs, double* rew) {
    const double dt = 0.1, damping = 0.25, cf = 1e2, k = 1e-3, size = 0.15, accel = 5.0;
    const int O = 6 * N;
    #pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < E; ++e) {
        double* p = pos + e * 4 * N;
        double* v = vel + e * 2 * N;
        double f[64][2];
        for (int a = 0; a < N; ++a) {
            int c = act[e * N + a];
            f[a][0] = (c == 1 ? -1.0 : (c == 2 ? 1.0 : 0.0)) * accel;
            f[a][1] = (c == 3 ? -1.0 : (c == 4 ? 1.0 : 0.0)) * accel;
        }
        for (int a = 0; a < N; ++a)
            for (int b = a + 1; b < N; ++b) {
                double dx = p[2 * a] - p[2 * b], dy = p[2 * a + 1] - p[2 * b + 1];
                double dist = sqrt(dx * dx + dy * dy);
                double x = -(dist - 2 * size) / k;
                double pen = (x > 0 ? x + log1p(exp(-x)) : log1p(exp(x))) * k;
                double fx = cf * dx / dist * pen, fy = cf * dy / dist * pen;
                f[a][0] += fx; f[a][1] += fy; f[b][0] -= fx; f[b][1] -= fy;
            }
        for (int a = 0; a < N; ++a)
            for (int d = 0; d < 2; ++d) {
                if (pos_first) p[2 * a + d] += v[2 * a + d] * dt;
                v[2 * a + d] = v[2 * a + d] * (1 - damping) + f[a][d] * dt;
                if (!pos_first) p[2 * a + d] += v[2 * a + d] * dt;
            }
        double g = 0.0;
        for (int l = 0; l < N; ++l) {
            double m = 1e300;
            for (int a = 0; a < N; ++a) {
                double dx = p[2 * a] - p[2 * (N + l)], dy = p[2 * a + 1] - p[2 * (N + l) + 1];
                double d = sqrt(dx * dx + dy * dy);
                if (d < m) m = d;
            }
            g -= m;
        }
        for (int a = 0; a < N; ++a) {
            double r = 0.0;
            for (int b = 0; b < N; ++b) {
                if (b == a) continue;
                double dx = p[2 * a] - p[2 * b], dy = p[2 * a + 1] - p[2 * b + 1];
                if (sqrt(dx * dx + dy * dy) < 2 * size) r -= 1.0;
            }
            rew[e * N + a] = g * (1 - local_ratio) + r * local_ratio;
            float* o = obs + (e * N + a) * O;
            int q = 0;
            o[q++] = (float)v[2 * a]; o[q++] = (float)v[2 * a + 1];
            o[q++] = (float)p[2 * a]; o[q++] = (float)p[2 * a + 1];
            for (int l = 0; l < N; ++l) { o[q++] = (float)(p[2 * (N + l)] - p[2 * a]); o[q++] = (float)(p[2 * (N + l) + 1] - p[2 * a + 1]); }
            for (int b = 0; b < N; ++b) { if (b == a) continue; o[q++] = (float)(p[2 * b] - p[2 * a]); o[q++] = (float)(p[2 * b + 1] - p[2 * a + 1]); }
            for (; q < O; ++q) o[q] = 0.0f;
        }
    }
}
