"""K1 parity (GPU): csrc/mpe.cu through the C ABI vs the numpy float64 restatement in
oracle/mpe.py.  MPE parity is UNPINNED against PettingZoo itself (third-party, absent; see
oracle/mpe.py) -- these tests pin the kernel to the restatement.

Tolerances (float32 kernel vs float64 oracle):
 * teacher-forced single steps (kernel state := oracle state each step): rtol 1e-5, atol 1e-5
   on state, obs and reward (north_star: 1e-5 relative);
 * free-running 25-step episodes: contacts amplify rounding (contact stiffness 100/1e-3), so
   the check is median |err| < 1e-5 and 99.9th percentile < 1e-3.
Reset draws (Philox) are bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle.mpe import OracleMPE

pytestmark = pytest.mark.gpu

CASES = [("simple_spread_v3", 0, dict(n_agents=3)), ("simple_spread_v3", 0, dict(n_agents=10)),
         ("simple_tag_v3", 1, dict()), ("simple_spread_v3", 0, dict(n_agents=5, pos_first=False)),
         ("simple_adversary_v3", 2, dict(num_good=2, num_adversaries=1)), ("simple_adversary_v3", 2, dict(num_good=4, num_adversaries=1))]


def _mk(name, scen, kw, E, seed=0):
    from tianshou_marl_b200.env import MPEVectorEnv

    env = MPEVectorEnv(name, n_envs=E, seed=seed, **kw)
    orc = OracleMPE(scen, E, seed=seed, **kw)
    return env, orc


@pytest.mark.parametrize("name,scen,kw", CASES)
def test_reset_bit_exact(name, scen, kw):
    env, orc = _mk(name, scen, kw, 257, seed=1234567890123)
    for _ in range(3):  # three successive episodes: the Philox counter advances
        obs, _ = env.reset()
        want = orc.reset()
        assert np.array_equal(obs.cpu().numpy(), want.reshape(-1, orc.obs_dim))
        assert np.array_equal(env.pos.cpu().numpy().reshape(257, -1, 2), orc.pos[:, : orc.A].astype(np.float32))
    # partial reset by row ids of 3 worlds
    A = orc.A
    rows = np.concatenate([np.arange(e * A, (e + 1) * A) for e in (5, 17, 200)])
    obs, _ = env.reset(rows)
    want = orc.reset(np.array([5, 17, 200]))
    assert np.array_equal(obs.cpu().numpy(), want.reshape(-1, orc.obs_dim))
    with pytest.raises(ValueError):
        env.reset(rows[:-1])


@pytest.mark.parametrize("name,scen,kw", CASES)
def test_teacher_forced_steps(name, scen, kw):
    E = 4096
    env, orc = _mk(name, scen, kw, E)
    env.reset()
    orc.reset()
    rng = np.random.default_rng(5)
    A = orc.A
    for t in range(30):
        # crowd some worlds so that collisions actually happen
        if t % 5 == 0:
            orc.pos[: E // 4] *= 0.25
        env.pos.copy_(torch.as_tensor(orc.pos[:, :A].reshape(-1, 2), dtype=torch.float32))
        env.vel.copy_(torch.as_tensor(orc.vel.reshape(-1, 2), dtype=torch.float32))
        env.lm.copy_(torch.as_tensor(orc.pos[:, A:].reshape(-1, 2), dtype=torch.float32))
        orc.pos[:] = np.concatenate([env.pos.cpu().numpy().reshape(E, A, 2), env.lm.cpu().numpy().reshape(E, -1, 2)], 1)
        orc.vel[:] = env.vel.cpu().numpy().reshape(E, A, 2)
        act = rng.integers(0, 5, (E, A))
        obs, rew, term, trunc, _ = env.step(act.reshape(-1))
        o_w, r_w, te_w, tr_w = orc.step(act)
        assert np.allclose(env.pos.cpu().numpy().reshape(E, A, 2), orc.pos[:, :A], rtol=1e-5, atol=1e-5)
        assert np.allclose(env.vel.cpu().numpy().reshape(E, A, 2), orc.vel, rtol=1e-5, atol=1e-5)
        assert np.allclose(obs.cpu().numpy().reshape(E, A, -1), o_w, rtol=1e-5, atol=1e-5)
        rw = rew.cpu().numpy().reshape(E, A)
        # rewards contain collision counts (step functions of distance): allow a flip only where
        # a pair sits within 1e-5 of the contact threshold
        bad = ~np.isclose(rw, r_w, rtol=1e-5, atol=1e-5)
        assert bad.mean() < 1e-4, bad.mean()
        assert not term.any().item()
        assert np.array_equal(trunc.cpu().numpy().reshape(E, A), tr_w)
        if tr_w.any():
            env.reset()
            orc.reset()


@pytest.mark.parametrize("name,scen,kw", CASES[:3])
def test_free_running_episode(name, scen, kw):
    E = 4096
    env, orc = _mk(name, scen, kw, E, seed=3)
    o0, _ = env.reset()
    assert np.array_equal(o0.cpu().numpy(), orc.reset().reshape(-1, orc.obs_dim))
    rng = np.random.default_rng(9)
    A = orc.A
    for t in range(25):
        act = rng.integers(0, 5, (E, A))
        obs, rew, term, trunc, _ = env.step(act.reshape(-1))
        o_w, r_w, te_w, tr_w = orc.step(act)
    assert trunc.all().item() and tr_w.all()
    err = np.abs(obs.cpu().numpy().reshape(E, A, -1) - o_w)
    assert np.median(err) < 1e-5 and np.quantile(err, 0.999) < 1e-3, (np.median(err), np.quantile(err, 0.999), err.max())


def test_autoreset_carry():
    """tsm_mpe_step(autoreset=1): terminal obs goes to obs_next, the re-drawn obs to the carry
    (collector.py:935,1028), and the next episode's draws match the oracle's."""
    E, A = 300, 3
    env, orc = _mk("simple_spread_v3", 0, dict(n_agents=3, max_cycles=4), E, seed=11)
    orc.max_cycles = 4
    env.reset()
    orc.reset()
    rng = np.random.default_rng(1)
    carry = torch.zeros(E * A, 18, device="cuda")
    for t in range(9):
        act = rng.integers(0, 5, (E, A))
        obs, rew, term, trunc, _ = env.step(act.reshape(-1), autoreset=True, out={"obs_carry": carry})
        o_w, r_w, _, tr_w = orc.step(act)
        assert np.allclose(obs.cpu().numpy().reshape(E, A, -1), o_w, rtol=1e-4, atol=1e-4)
        if tr_w.all():
            o_reset = orc.reset()
            assert np.array_equal(carry.cpu().numpy().reshape(E, A, -1), o_reset)
            assert int(env.steps.max().item()) == 0
        else:
            assert torch.equal(carry, obs)
