"""Collector (GPU): the fused lock-step path vs the generic reference-style loop (same seeds ->
identical buffer contents), slot layout like the reference's collector tests
(test/base/test_collector.py:155-178: row r, step t -> flat index r*size + t), episode
accounting vs the MPE truncation horizon, and argument validation (collector.py:526-535, :827-831)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(E=64, A=3, T=12, max_cycles=5, seed=3, net_seed=1):
    from tianshou_marl_b200.algorithm import DiscreteActorPolicy
    from tianshou_marl_b200.data import VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, max_cycles=max_cycles, seed=seed)
    net = ActorCriticNet(env.obs_dim, 64, 5, seed=net_seed)
    pol = DiscreteActorPolicy(actor=net, seed=11)
    buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
    return env, net, pol, buf


def test_fused_equals_generic_loop():
    from tianshou_marl_b200.data import Collector

    E, A, T = 64, 3, 12
    env1, _, pol1, buf1 = _setup(E, A, T)
    env2, _, pol2, buf2 = _setup(E, A, T)
    c1 = Collector(pol1, env1, buf1)
    c2 = Collector(pol2, env2, buf2)
    c2._fused_ok = lambda random: False  # force the generic loop
    c1.reset()
    c2.reset()
    s1 = c1.collect(n_step=T * E * A)
    s2 = c2.collect(n_step=T * E * A)
    assert s1.n_collected_steps == s2.n_collected_steps == T * E * A
    assert s1.n_collected_episodes == s2.n_collected_episodes == (T // 5) * E * A
    for k in ("obs", "obs_next", "act", "rew", "terminated", "truncated", "done"):
        assert torch.equal(buf1._store[k], buf2._store[k]), k
    assert torch.allclose(buf1._store["logp"], buf2._store["logp"])
    assert np.allclose(np.sort(s1.returns), np.sort(s2.returns))
    assert (s1.lens == 5).all() and len(s1.lens) == s1.n_collected_episodes
    assert np.isclose(s1.returns_stat.mean, s1.returns.mean())
    # reference layout: sample(0) is row-major over rows, each row in time order
    idx = buf1.sample_indices(0)
    assert torch.equal(idx, torch.arange(T * E * A, device="cuda"))
    b = buf1[torch.tensor([0, 1, T, T + 1], device="cuda")]
    assert torch.equal(b.obs[0], buf1._store["obs"][0, 0]) and torch.equal(b.obs[1], buf1._store["obs"][1, 0])
    assert torch.equal(b.obs[2], buf1._store["obs"][0, 1])
    # obs[t+1] == obs_next[t] unless the row finished an episode at t (then it is the reset observation)
    st = buf1._store
    done = st["done"][:-1].bool()
    same = (st["obs"][1:] == st["obs_next"][:-1]).all(-1)
    assert same[~done].all() and not same[done].any()
    # a second collect continues from the carried observation and wraps the ring
    s3 = c1.collect(n_step=2 * E * A)
    assert len(buf1) == T * E * A and s3.n_collected_steps == 2 * E * A
    assert int(buf1._ins[0].item()) == 2


def test_collect_validation_and_random():
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer

    env, _, pol, buf = _setup(8, 3, 4)
    c = Collector(pol, env, buf)
    with pytest.raises(ValueError):
        c.collect(n_step=10)  # before reset
    c.reset()
    with pytest.raises(ValueError):
        c.collect(n_step=10, n_episode=1)
    with pytest.raises(ValueError):
        c.collect()
    with pytest.raises(ValueError):
        Collector(pol, env, VectorReplayBuffer(16, 4, index_type="torch"))  # fewer sub-buffers than rows
    s = c.collect(n_step=30, random=True)  # ceil(30/24) = 2 vector steps
    assert s.n_collected_steps == 48
    a = buf._store["act"][:2].cpu().numpy()
    assert a.min() >= 0 and a.max() <= 4 and len(np.unique(a)) == 5
    s = c.collect(n_episode=24)
    assert s.n_collected_episodes >= 24
    assert s.collect_speed > 0


def test_collect_then_update_learns():
    """collect -> PPO.update -> reset_buffer loop runs end to end and improves the simple_spread return."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory
    from tianshou_marl_b200.data import Collector

    E, A, T = 512, 3, 50
    env, net, pol, buf = _setup(E, A, T, max_cycles=25, seed=0)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5)
    c = Collector(alg, env, buf)
    c.reset()
    pol.is_within_training_step = True
    rets = []
    for it in range(12):
        s = c.collect(n_step=T * E * A)
        rets.append(s.returns_stat.mean)
        alg.update(buf, 8192, 4)
        c.reset_buffer(keep_statistics=True)
    assert np.isfinite(rets).all()
    assert np.mean(rets[-3:]) > np.mean(rets[:3]), rets


@pytest.mark.gpu
def test_parallel_stream_rollout_equals_single_stream():
    """tsm_collect_lockstep splits the environments into ranges that run on their own streams (TSM_COLLECT_STREAMS,
    read once per process): the rollout -- actions included, the sampling key is the global row -- must not depend on it."""
    import os
    import subprocess
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    lines = []
    for parts in ("1", "2", "3"):
        out = subprocess.run([sys.executable, os.path.join(here, "_split_rollout.py")], capture_output=True, text=True, timeout=300,
                             env={**os.environ, "TSM_COLLECT_STREAMS": parts})
        chk = [l for l in out.stdout.splitlines() if l.startswith("CHK")]
        assert out.returncode == 0 and chk, out.stdout[-2000:] + out.stderr[-2000:]
        lines.append(chk[0])
    assert lines[0] == lines[1] == lines[2], lines
