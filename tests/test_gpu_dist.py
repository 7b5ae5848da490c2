"""Multi-GPU check of the fused all-reduce + Adam path: spawns tests/dist_gpu_check.py under torchrun with 2 ranks
(skipped on a single-GPU box; the CPU-side logic is covered by tests/test_dist_cpu.py with gloo)."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
def test_peer_gradient_exchange_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(HERE, "dist_gpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "DIST_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.gpu
def test_mappo_peer_gradient_exchange_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29535", os.path.join(HERE, "dist_gpu_check_mappo.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "DIST_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.gpu
def test_update_is_undone_when_the_exchange_reports_a_missing_peer(monkeypatch):
    """The in-kernel exchange gives up on a peer after ~2 s and the epoch finishes with what it has; PPO.update then restores the
    parameters and optimizer state it started from and raises.  One GPU is enough to check the undo: the peer path is selected by
    hand (world size 2 reported, the epoch entry point forwarded to the single-GPU one, a peer object whose check() fails)."""
    import torch

    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    env = MPEVectorEnv("simple_spread_v3", n_envs=128, seed=1)
    net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, engine="tc")
    pol = DiscreteActorPolicy(actor=net, seed=3)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), seed=4)
    col = Collector(pol, env, VectorReplayBuffer(25 * env.n_rows, env.n_rows, index_type="torch"))
    col.reset()
    pol.is_within_training_step = True
    col.collect(n_step=25 * env.n_rows)
    alg.update(col.buffer, 1024, 1)  # one good update: non-trivial Adam state
    before = [t.detach().clone() for t in (net.flat, net.adam_m, net.adam_v, net.adam_step)]
    ku = net.kernel_updates

    class _Timeout:  # what PeerGradients.check() does after a timeout
        def check(self) -> None:
            raise _lib.TsmError("cross-rank gradient barrier timed out (a peer rank did not reach the optimizer step)")

    monkeypatch.setattr(torch.distributed, "is_initialized", lambda: True)
    monkeypatch.setattr(torch.distributed, "get_world_size", lambda *a, **k: 2)
    alg._peers = None
    real_epoch = _lib.lib().tsm_ppo_epoch

    def fake_dist(netd, pb, cfg, hb, n_mb, st, lr, b1, b2, eps, mgn, sums, peers, seq, stream):  # the epoch runs, the peers "time out"
        return real_epoch(netd, pb, cfg, hb, n_mb, st, lr, b1, b2, eps, mgn, sums, stream)

    class _Peers(_Timeout):
        seq = 0

        def byref(self):
            return None

    monkeypatch.setattr(_lib.lib(), "tsm_ppo_epoch_dist", fake_dist, raising=False)
    alg._peers = _Peers()
    with pytest.raises(_lib.TsmError, match="timed out"):
        alg.update(col.buffer, 1024, 2)
    after = (net.flat, net.adam_m, net.adam_v, net.adam_step)
    assert all(torch.equal(a, b.detach()) for a, b in zip(before, after)) and net.kernel_updates == ku
    assert float(net.grads.abs().max()) == 0.0
    # and the net is usable again (the fp16 image is rebuilt from the restored parameters)
    monkeypatch.undo()
    alg._peers = None
    ts = alg.update(col.buffer, 1024, 1)
    assert ts.gradient_steps > 0 and not torch.equal(before[0], net.flat.detach())
