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


CONT_CASES = [CASES[0], CASES[2], CASES[4]]


@pytest.mark.parametrize("name,scen,kw", CONT_CASES)
def test_continuous_actions_teacher_forced(name, scen, kw):
    """continuous_actions=True (tsm_mpe_step_continuous): Box(0, 1, (5,)) action vectors, values outside the box included."""
    from tianshou_marl_b200.env import MPEVectorEnv

    E = 2048
    env = MPEVectorEnv(name, n_envs=E, seed=3, continuous_actions=True, **kw)
    orc = OracleMPE(scen, E, seed=3, **kw)
    assert env.single_action_space.shape == (5,) and env.single_action_space.low == 0.0 and env.single_action_space.high == 1.0
    env.reset()
    orc.reset()
    rng = np.random.default_rng(11)
    A = orc.A
    for t in range(20):
        if t % 5 == 0:
            orc.pos[: E // 4] *= 0.25
        env.pos.copy_(torch.as_tensor(orc.pos[:, :A].reshape(-1, 2), dtype=torch.float32))
        env.vel.copy_(torch.as_tensor(orc.vel.reshape(-1, 2), dtype=torch.float32))
        env.lm.copy_(torch.as_tensor(orc.pos[:, A:].reshape(-1, 2), dtype=torch.float32))
        orc.pos[:] = np.concatenate([env.pos.cpu().numpy().reshape(E, A, 2), env.lm.cpu().numpy().reshape(E, -1, 2)], 1)
        orc.vel[:] = env.vel.cpu().numpy().reshape(E, A, 2)
        act = rng.uniform(-0.25, 1.25, (E, A, 5)).astype(np.float32)  # a quarter of the entries get clipped
        obs, rew, term, trunc, _ = env.step(act.reshape(-1, 5))
        o_w, r_w, _, tr_w = orc.step(act)
        assert np.allclose(env.pos.cpu().numpy().reshape(E, A, 2), orc.pos[:, :A], rtol=1e-5, atol=1e-5)
        assert np.allclose(env.vel.cpu().numpy().reshape(E, A, 2), orc.vel, rtol=1e-5, atol=1e-5)
        assert np.allclose(obs.cpu().numpy().reshape(E, A, -1), o_w, rtol=1e-5, atol=1e-5)
        bad = ~np.isclose(rew.cpu().numpy().reshape(E, A), r_w, rtol=1e-5, atol=1e-5)
        assert bad.mean() < 1e-4, bad.mean()
        assert np.array_equal(trunc.cpu().numpy().reshape(E, A), tr_w)
        if tr_w.any():
            env.reset()
            orc.reset()


@pytest.mark.parametrize("name,scen,kw", CONT_CASES)
def test_one_hot_continuous_equals_discrete(name, scen, kw):
    """simple_env._set_action: the continuous ordering keeps the directions of discrete actions 1..4, so a one-hot vector moves
    the world exactly like the discrete action of that index: bit-equal state, observations and rewards."""
    from tianshou_marl_b200.env import MPEVectorEnv

    E = 512
    d = MPEVectorEnv(name, n_envs=E, seed=9, **kw)
    c = MPEVectorEnv(name, n_envs=E, seed=9, continuous_actions=True, **kw)
    od, _ = d.reset()
    oc, _ = c.reset()
    assert torch.equal(od, oc)
    g = torch.Generator(device="cuda").manual_seed(2)
    for _ in range(25):
        a = torch.randint(0, 5, (d.n_rows,), device="cuda", generator=g)
        od, rd, _, td, _ = d.step(a)
        oc, rc, _, tc, _ = c.step(torch.nn.functional.one_hot(a, 5).to(torch.float32))
        assert torch.equal(od, oc) and torch.equal(rd, rc) and torch.equal(td, tc)
        assert torch.equal(d.pos, c.pos) and torch.equal(d.vel, c.vel)


def test_continuous_policy_collects_and_updates_on_mpe():
    """Box actions end to end: ContinuousActorPolicy (Gaussian head, Policy.map_action scaling [-1, 1] -> [0, 1]) acting in
    MPEVectorEnv(continuous_actions=True) through the Collector, float action vectors in the buffer, one PPO update."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, ContinuousActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    E, T = 64, 25
    env = MPEVectorEnv("simple_spread_v3", n_envs=E, seed=1, continuous_actions=True)
    net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, gaussian=True)
    before = net.flat.detach().clone()
    pol = ContinuousActorPolicy(actor=net, action_space=env.single_action_space)
    pol.is_within_training_step = True  # (what the trainer sets around collect + update)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5)
    buf = VectorReplayBuffer(T * env.n_rows, env.n_rows, index_type="torch")
    col = Collector(pol, env, buf)
    col.reset()
    st = col.collect(n_step=T * env.n_rows)
    assert st.n_collected_steps == T * env.n_rows and st.n_collected_episodes == env.n_rows
    act_f = buf._store["act_f"]
    assert act_f.shape == (T, env.n_rows, 5) and bool(torch.isfinite(act_f).all()) and float(act_f.std()) > 0.1
    ts = alg.update(buf, 1024, 2)
    assert ts.gradient_steps == 2 * len(alg.split_bounds(T * env.n_rows, 1024)) - 2
    assert np.isfinite(ts.loss.mean) and np.isfinite(ts.vf_loss.mean)
    assert not torch.equal(before, net.flat.detach())
    assert float((net.view("log_sigma").detach() - 0.0).abs().max()) > 0.0  # the scale parameters moved too


def test_kernel_contact_forces_cancel_at_full_size():
    """A property that needs no oracle (MPE parity is unpinned): contact forces come in opposite pairs, so with no-op actions the
    summed velocity of a world's agents only decays by the damping -- sum(v') = 0.75 * sum(v) -- however hard the agents overlap.
    65 536 simple_spread worlds (BASELINE.json configs[1] size), all agents piled up near the origin."""
    from tianshou_marl_b200.env import MPEVectorEnv

    E, A = 65536, 3
    env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=5)
    env.reset()
    g = torch.Generator(device="cuda").manual_seed(1)
    env.pos.copy_(torch.randn(E * A, 2, device="cuda", generator=g) * 0.1)
    env.vel.copy_(torch.randn(E * A, 2, device="cuda", generator=g))
    before = env.vel.view(E, A, 2).sum(1).double()
    speed_before = float(env.vel.abs().max())
    env.step(torch.zeros(E * A, dtype=torch.int32, device="cuda"))
    after = env.vel.view(E, A, 2).sum(1).double()
    assert float(env.vel.abs().max()) > 2.0 * speed_before / 3.0 and float(env.vel.abs().max()) > 5.0  # the contacts did push
    # pairs cancel exactly in IEEE arithmetic (same formula on negated differences); what is left is the rounding of each agent's sum
    assert float((after - 0.75 * before).abs().max()) < 2e-3
