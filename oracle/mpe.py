"""Oracle: MPE simple_spread / simple_tag physics in numpy float64.  Test infrastructure only.

PARITY UNPINNED: the simulator is pettingzoo 1.24.2 (poetry.lock:4126-4127 of the
reference), a third-party dependency that is not in the reference tree and not installable
offline.  This file restates its published algorithm:
  pettingzoo/mpe/_mpe_utils/core.py          World.step / apply_action_force /
                                             apply_environment_force / get_collision_force /
                                             integrate_state
  pettingzoo/mpe/_mpe_utils/simple_env.py    _set_action / _execute_world_step / step
  pettingzoo/mpe/scenarios/simple_spread.py  make_world / reset_world / reward /
                                             global_reward / observation
  pettingzoo/mpe/scenarios/simple_tag.py     make_world / reset_world / agent_reward /
                                             adversary_reward / observation
The reference reaches it through tianshou/env/pettingzoo_env.py:38-120 and
tianshou/env/enhanced_pettingzoo_env.py:90-222.  No reference test pins its numbers.

Vectorised over environments: state arrays are (E, n_entities, 2) float64; entity loops run
in PettingZoo's order.  Reset positions come from oracle.philox (our RNG, f32-exact) or from
the caller.
"""
from __future__ import annotations

import numpy as np

from . import philox

DT = 0.1
DAMPING = 0.25
CONTACT_FORCE = 1e2
CONTACT_MARGIN = 1e-3

SPREAD, TAG, ADVERSARY = 0, 1, 2


class OracleMPE:
    def __init__(self, scenario: int, n_envs: int, n_agents: int = 3, max_cycles: int = 25,
                 local_ratio: float = 0.5, pos_first: bool = True, seed: int = 0,
                 num_good: int = 1, num_adversaries: int = 3, num_obstacles: int = 2) -> None:
        self.scenario, self.E, self.max_cycles, self.pos_first, self.seed = scenario, n_envs, max_cycles, pos_first, seed
        if scenario == SPREAD:
            A = L = n_agents
            self.size = np.array([0.15] * A + [0.05] * L)
            self.collide = np.array([True] * A + [False] * L)
            self.accel = np.full(A, 5.0)  # sensitivity default, simple_env._set_action
            self.max_speed = np.full(A, np.inf)
            self.adversary = np.zeros(A, bool)
            self.local_ratio = local_ratio
            self.lm_lo, self.lm_hi = -1.0, 1.0
            self.obs_dim = 6 * A
        elif scenario == TAG:
            A, L = num_adversaries + num_good, num_obstacles
            self.adversary = np.arange(A) < num_adversaries
            self.size = np.array([0.075 if a else 0.05 for a in self.adversary] + [0.2] * L)
            self.collide = np.array([True] * (A + L))
            self.accel = np.where(self.adversary, 3.0, 4.0)
            self.max_speed = np.where(self.adversary, 1.0, 1.3)
            self.local_ratio = None
            self.lm_lo, self.lm_hi = -0.9, 0.9
            n_good = num_good
            # adversary obs: 4 + 2L + 2(A-1) + 2*n_good ; good agent: 4 + 2L + 2(A-1) + 2*(n_good-1)
            self.obs_dim = 4 + 2 * L + 2 * (A - 1) + 2 * n_good
        elif scenario == ADVERSARY:
            # scenarios/simple_adversary.py: 1 adversary first, then N good agents; N landmarks, one of them the goal; nothing collides
            A, L = num_adversaries + num_good, num_good
            self.adversary = np.arange(A) < num_adversaries
            self.size = np.array([0.15] * A + [0.08] * L)
            self.collide = np.array([False] * (A + L))
            self.accel = np.full(A, 5.0)
            self.max_speed = np.full(A, np.inf)
            self.local_ratio = None
            self.lm_lo, self.lm_hi = -1.0, 1.0
            self.obs_dim = 2 + 2 * L + 2 * (A - 1)
            self.goal = np.zeros(n_envs, np.int64)
        else:
            raise ValueError(scenario)
        self.A, self.L = A, L
        self.pos = np.zeros((n_envs, A + L, 2))
        self.vel = np.zeros((n_envs, A, 2))
        self.steps = np.zeros(n_envs, np.int64)
        self.episode = np.zeros(n_envs, np.int64)  # number of resets so far, RNG counter

    # -- reset ------------------------------------------------------------------------------
    def draw_positions(self, env_ids, episode) -> np.ndarray:
        """(n, A+L, 2) float32 positions from Philox: counter (env_lo, env_hi, episode, block),
        key (seed_lo, seed_hi ^ STREAM_RESET); block b yields entities 2b (x,y) and 2b+1 (x,y)."""
        env_ids = np.asarray(env_ids, dtype=np.uint64)
        n, ne = len(env_ids), self.A + self.L
        out = np.zeros((n, ne, 2), np.float32)
        nblk = (ne + 1) // 2
        for b in range(nblk):
            r = philox.philox4x32(env_ids & 0xFFFFFFFF, env_ids >> np.uint64(32), np.asarray(episode, np.uint64) & 0xFFFFFFFF,
                                  np.uint32(b), np.uint32(self.seed & 0xFFFFFFFF),
                                  np.uint32(((self.seed >> 32) & 0xFFFFFFFF) ^ philox.STREAM_RESET))
            u = [philox.u01(x) for x in r]
            for j in range(2):
                e = 2 * b + j
                if e >= ne:
                    break
                lo, hi = (-1.0, 1.0) if e < self.A else (self.lm_lo, self.lm_hi)
                lo32, w32 = np.float32(lo), np.float32(hi - lo)
                out[:, e, 0] = lo32 + w32 * u[2 * j]  # two rounded f32 ops, like __fmul_rn/__fadd_rn
                out[:, e, 1] = lo32 + w32 * u[2 * j + 1]
        return out

    def draw_goal(self, env_ids, episode) -> np.ndarray:
        """simple_adversary reset_world: goal = np_random.choice(landmarks); here the Philox draw of the entity slot behind the
        last landmark (same counter layout as draw_positions), floor(u * L)."""
        env_ids = np.asarray(env_ids, dtype=np.uint64)
        ent = self.A + self.L
        r = philox.philox4x32(env_ids & 0xFFFFFFFF, env_ids >> np.uint64(32), np.asarray(episode, np.uint64) & 0xFFFFFFFF,
                              np.uint32(ent >> 1), np.uint32(self.seed & 0xFFFFFFFF),
                              np.uint32(((self.seed >> 32) & 0xFFFFFFFF) ^ philox.STREAM_RESET))
        u = philox.u01(r[2] if ent & 1 else r[0])
        return np.minimum((u * np.float32(self.L)).astype(np.int64), self.L - 1)

    def reset(self, env_ids=None, init_pos=None) -> np.ndarray:
        """reset_world (simple_spread.py / simple_tag.py): agents then landmarks, zero velocity.
        Returns obs (n, A, obs_dim) float32."""
        env_ids = np.arange(self.E) if env_ids is None else np.asarray(env_ids, dtype=np.int64)
        if init_pos is None:
            init_pos = self.draw_positions(env_ids, self.episode[env_ids])
        if self.scenario == ADVERSARY:
            self.goal[env_ids] = self.draw_goal(env_ids, self.episode[env_ids])
        self.pos[env_ids] = np.asarray(init_pos, dtype=np.float64)
        self.vel[env_ids] = 0.0
        self.steps[env_ids] = 0
        self.episode[env_ids] += 1
        return self.observe(env_ids)

    # -- step -------------------------------------------------------------------------------
    def step(self, act):
        """One parallel cycle (simple_env.step with all agents' actions -> _execute_world_step).
        act: (E, A) ints in 0..4, or (E, A, 5) floats (continuous_actions=True).  Returns obs (E,A,O) f32, rew (E,A) f64, terminated, truncated (E,A) bool."""
        act = np.asarray(act)
        A, L = self.A, self.L
        u = np.zeros((self.E, A, 2))
        if np.issubdtype(act.dtype, np.floating):
            # continuous_actions=True: Box(0, 1, (5,)) per agent, clipped to the box by ClipOutOfBoundsWrapper (mpe make_env wraps
            # continuous environments with it), then simple_env._set_action: u[0] += a[2] - a[1]; u[1] += a[4] - a[3] (float32
            # differences added to float64 zeros; the order keeps the directions of discrete actions 1..4)
            a = np.clip(act.reshape(self.E, A, 5).astype(np.float32), np.float32(0.0), np.float32(1.0))
            u[..., 0] = a[..., 2] - a[..., 1]
            u[..., 1] = a[..., 4] - a[..., 3]
        else:
            # simple_env._set_action (discrete): u in {-1,0,1}^2 times sensitivity
            act = act.reshape(self.E, A)
            u[..., 0] = np.where(act == 1, -1.0, np.where(act == 2, 1.0, 0.0))
            u[..., 1] = np.where(act == 3, -1.0, np.where(act == 4, 1.0, 0.0))
        u *= self.accel[None, :, None]
        # core.World.apply_action_force (u_noise None) then apply_environment_force
        force = np.zeros((self.E, A + L, 2))
        force[:, :A] = u
        for a in range(A + L):
            for b in range(a + 1, A + L):
                if not (self.collide[a] and self.collide[b]):
                    continue
                if a >= A and b >= A:
                    continue  # neither movable: forces are None
                delta = self.pos[:, a] - self.pos[:, b]
                dist = np.sqrt(np.sum(np.square(delta), axis=-1))
                dist_min = self.size[a] + self.size[b]
                k = CONTACT_MARGIN
                pen = np.logaddexp(0, -(dist - dist_min) / k) * k
                f = CONTACT_FORCE * delta / dist[:, None] * pen[:, None]
                if a < A:
                    force[:, a] = f + force[:, a]
                if b < A:
                    force[:, b] = -f + force[:, b]
        # core.World.integrate_state (mass = 1)
        p, v = self.pos[:, :A], self.vel
        if self.pos_first:
            p = p + v * DT
        v = v * (1 - DAMPING)
        v = v + (force[:, :A] / 1.0) * DT
        speed = np.sqrt(np.square(v[..., 0]) + np.square(v[..., 1]))
        over = speed > self.max_speed[None, :]
        with np.errstate(invalid="ignore", divide="ignore"):
            vc = v / speed[..., None] * self.max_speed[None, :, None]
        v = np.where(over[..., None], vc, v)
        if not self.pos_first:
            p = p + v * DT
        self.pos[:, :A], self.vel = p, v
        rew = self.reward()
        self.steps += 1
        trunc = np.repeat((self.steps >= self.max_cycles)[:, None], A, axis=1)
        term = np.zeros_like(trunc)
        return self.observe(), rew, term, trunc

    def _dist(self, i, j):
        d = self.pos[:, i] - self.pos[:, j]
        return np.sqrt(np.sum(np.square(d), axis=-1))

    def reward(self) -> np.ndarray:
        A, L = self.A, self.L
        rew = np.zeros((self.E, A))
        if self.scenario == SPREAD:
            # simple_spread.reward: -1 per other agent in collision; global_reward: -sum_l min_a dist
            for i in range(A):
                for j in range(A):
                    if i != j:
                        rew[:, i] -= 1.0 * (self._dist(j, i) < self.size[j] + self.size[i])
            g = np.zeros(self.E)
            for l in range(L):
                d = np.stack([self._dist(a, A + l) for a in range(A)], axis=0)
                g -= d.min(axis=0)
            # simple_env._execute_world_step: global*(1-lr) + local*lr
            return g[:, None] * (1 - self.local_ratio) + rew * self.local_ratio
        adv = np.nonzero(self.adversary)[0]
        good = np.nonzero(~self.adversary)[0]
        if self.scenario == ADVERSARY:
            # simple_adversary.reward, shaped: adversary_reward = -|p - goal|; agent_reward = sum_adv |p_adv - goal| - min_good |p_good - goal|
            gp = self.pos[np.arange(self.E), A + self.goal]
            dg = np.sqrt(np.sum(np.square(self.pos[:, :A] - gp[:, None]), axis=-1))  # (E, A)
            adv_rew = dg[:, adv].sum(axis=1)
            pos_rew = -dg[:, good].min(axis=1)
            rew[:, adv] = -dg[:, adv]
            rew[:, good] = (pos_rew + adv_rew)[:, None]
            return rew
        # adversary_reward: +10 per (good, adversary) pair in collision, shared
        ar = np.zeros(self.E)
        for ag in good:
            for ad in adv:
                ar += 10.0 * (self._dist(ag, ad) < self.size[ag] + self.size[ad])
        rew[:, adv] = ar[:, None]
        for ag in good:
            r = np.zeros(self.E)
            for ad in adv:
                r -= 10.0 * (self._dist(ad, ag) < self.size[ad] + self.size[ag])
            for p in range(2):
                x = np.abs(self.pos[:, ag, p])
                with np.errstate(over="ignore"):
                    b = np.where(x < 0.9, 0.0, np.where(x < 1.0, (x - 0.9) * 10, np.minimum(np.exp(2 * x - 2), 10)))
                r -= b
            rew[:, ag] = r
        return rew

    def observe(self, env_ids=None) -> np.ndarray:
        """scenario.observation, cast to float32 like simple_env.observe."""
        sl = slice(None) if env_ids is None else env_ids
        pos, vel = self.pos[sl], self.vel[sl]
        A, L = self.A, self.L
        n = pos.shape[0]
        obs = np.zeros((n, A, self.obs_dim))
        if self.scenario == ADVERSARY:
            # simple_adversary.observation: good [goal - p, landmarks - p, others - p]; adversary [landmarks - p, others - p] (zero-padded)
            goal = self.goal[sl]
            gp = pos[np.arange(n), A + goal]
            for i in range(A):
                parts = [] if self.adversary[i] else [gp - pos[:, i]]
                parts += [pos[:, A + l] - pos[:, i] for l in range(L)]
                parts += [pos[:, j] - pos[:, i] for j in range(A) if j != i]
                row = np.concatenate(parts, axis=-1)
                obs[:, i, : row.shape[1]] = row
            return obs.astype(np.float32)
        for i in range(A):
            parts = [vel[:, i], pos[:, i]]
            parts += [pos[:, A + l] - pos[:, i] for l in range(L)]
            others = [j for j in range(A) if j != i]
            parts += [pos[:, j] - pos[:, i] for j in others]
            if self.scenario == SPREAD:
                parts += [np.zeros((n, 2)) for _ in others]  # comm (silent agents, state.c == 0)
            else:
                parts += [vel[:, j] for j in others if not self.adversary[j]]
            row = np.concatenate(parts, axis=-1)
            obs[:, i, : row.shape[1]] = row  # tag: good agent's 14 dims zero-padded to 16
        return obs.astype(np.float32)
