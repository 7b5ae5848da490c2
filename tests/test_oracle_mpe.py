"""oracle/mpe.py on the CPU: properties the restatement must have whatever PettingZoo version it is checked against later
(MPE parity is unpinned: the package is absent, see the oracle's header)."""
import numpy as np

from oracle.mpe import OracleMPE


def test_one_hot_continuous_action_equals_discrete():
    # simple_env._set_action: u[0] += a[2] - a[1]; u[1] += a[4] - a[3] keeps the directions of discrete actions 1..4
    for scen, kw in ((0, dict(n_agents=3)), (1, dict()), (2, dict(num_good=2, num_adversaries=1))):
        d, c = OracleMPE(scen, 64, seed=5, **kw), OracleMPE(scen, 64, seed=5, **kw)
        assert np.array_equal(d.reset(), c.reset())
        rng = np.random.default_rng(0)
        for _ in range(10):
            a = rng.integers(0, 5, (64, d.A))
            od, rd, _, _ = d.step(a)
            oc, rc, _, _ = c.step(np.eye(5, dtype=np.float32)[a])
            assert np.array_equal(od, oc) and np.array_equal(rd, rc)


def test_continuous_actions_are_clipped_to_the_box():
    a, b = OracleMPE(0, 16, n_agents=3, seed=1), OracleMPE(0, 16, n_agents=3, seed=1)
    a.reset()
    b.reset()
    rng = np.random.default_rng(3)
    act = rng.uniform(-2.0, 3.0, (16, 3, 5)).astype(np.float32)
    oa, ra, _, _ = a.step(act)
    ob, rb, _, _ = b.step(np.clip(act, 0.0, 1.0))
    assert np.array_equal(oa, ob) and np.array_equal(ra, rb)


def test_contact_forces_cancel_between_two_agents():
    """core.World.apply_environment_force: the force on b is minus the force on a (mass 1), so with no actions the total momentum
    of the agents only decays by the damping: sum(v') = 0.75 * sum(v), however hard the agents press against each other."""
    o = OracleMPE(0, 32, n_agents=3, seed=2)
    o.reset()
    rng = np.random.default_rng(1)
    o.pos[:, :3] = rng.normal(size=(32, 3, 2)) * 0.1  # everybody overlaps everybody
    o.vel[:] = rng.normal(size=(32, 3, 2))
    before = o.vel.sum(axis=1)
    o.step(np.zeros((32, 3), np.int64))
    assert np.allclose(o.vel.sum(axis=1), 0.75 * before, rtol=1e-9, atol=1e-9)
    assert np.abs(o.vel).max() > 1.5  # the contacts did push


def test_spread_reward_mixes_global_and_local_terms():
    """simple_spread: reward = global * (1 - local_ratio) + local * local_ratio with global = -sum over landmarks of the nearest
    agent's distance and local = -1 per other agent in collision (scenarios/simple_spread.py, simple_env._execute_world_step)."""
    for lr in (0.0, 0.5, 1.0):
        o = OracleMPE(0, 1, n_agents=3, local_ratio=lr, seed=0)
        o.reset()
        o.pos[0, :3] = [[0.0, 0.0], [0.1, 0.0], [2.0, 2.0]]   # agents 0 and 1 collide (0.1 < 0.3), agent 2 is alone
        o.pos[0, 3:] = [[0.0, 1.0], [2.0, 3.0], [-3.0, 0.0]]  # landmarks
        r = o.reward()[0]
        glob = -(1.0 + 1.0 + 3.0)  # nearest agents: (0,0)->1.0, (2,2)->1.0, (0,0)->3.0
        local = np.array([-1.0, -1.0, 0.0])
        assert np.allclose(r, glob * (1 - lr) + local * lr, atol=1e-12)


def test_tag_speed_cap_and_boundary_penalty():
    """simple_tag: agents are capped at max_speed (adversaries 1.0, the good agent 1.3; core.World.integrate_state) and the good agent
    pays the boundary penalty outside |x| = 0.9 (scenarios/simple_tag.py bound())."""
    o = OracleMPE(1, 1, seed=0)
    o.reset()
    o.pos[0, :4] = [[-5.0, -5.0], [5.0, -5.0], [-5.0, 5.0], [0.95, 0.0]]  # nobody touches anybody
    o.pos[0, 4:] = [[9.0, 9.0], [-9.0, 9.0]]
    o.vel[0] = [[3.0, 4.0], [0.3, 0.4], [-6.0, 8.0], [0.0, 13.0]]
    _, rew, _, _ = o.step(np.zeros((1, 4), np.int64))
    speed = np.sqrt((o.vel[0] ** 2).sum(-1))
    assert np.allclose(speed, [1.0, 0.375, 1.0, 1.3], atol=1e-12)  # 0.75 * 0.5 = 0.375 stays below its cap
    x, y = abs(o.pos[0, 3, 0]), abs(o.pos[0, 3, 1])
    want = -sum((v - 0.9) * 10 if 0.9 <= v < 1.0 else (min(np.exp(2 * v - 2), 10) if v >= 1.0 else 0.0) for v in (x, y))
    assert np.isclose(rew[0, 3], want, atol=1e-12) and rew[0, 3] < 0 and np.all(rew[0, :3] == 0.0)


def test_truncation_and_reset_counter():
    o = OracleMPE(0, 4, n_agents=3, max_cycles=5, seed=7)
    first = o.reset().copy()
    for t in range(5):
        _, _, term, trunc = o.step(np.zeros((4, 3), np.int64))
        assert not term.any() and trunc.all() == (t == 4)
    second = o.reset()
    assert not np.array_equal(first, second)  # the episode counter keys the draws
    o2 = OracleMPE(0, 4, n_agents=3, max_cycles=5, seed=7)
    assert np.array_equal(o2.reset(), first)  # and the seed reproduces them
