"""BASELINE.json configs[1] at its FULL size (65 536 simple_spread worlds x 3 agents, T = 128: 25 165 824 rows per step) through
the public classes bench.py times, checked by properties that do not need an oracle of that size: ring / episode bookkeeping
identities of the rollout, the epoch's slot table being a permutation of the valid slots, minibatch accounting of the update,
and a zero learning rate leaving the parameters untouched while the loss sequences stay finite and reproducible."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _workload(lr: float):
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory
    from tianshou_marl_b200.bench_workload import Workload

    args = SimpleNamespace(envs=65536, agents=3, horizon=128, hidden=64, engine="tc", batch_size=65536, repeat=2, no_update=False,
                           reuse_rollout=False)
    wl = Workload(args, 0, 1)
    wl.alg.optim = AdamOptimizerFactory(lr=lr)
    return wl, args


def test_full_size_c2_rollout_and_update_properties():
    if torch.cuda.get_device_properties(0).total_memory < 60e9:
        pytest.skip("needs the memory of a B200")
    wl, args = _workload(lr=0.0)
    R, T = wl.R, wl.T
    B = R * T
    stats = wl.collector.collect(n_step=B)
    buf, st = wl.buf, wl.buf._store
    # ---- rollout bookkeeping ----
    assert stats.n_collected_steps == B and len(buf) == B
    done = st["done"].view(T, R)
    term, trunc = st["terminated"].view(T, R), st["truncated"].view(T, R)
    assert torch.equal(done, term | trunc)
    # simple_spread never terminates; every world truncates after max_cycles = 25 steps, all rows together (auto-reset)
    assert int(term.sum()) == 0
    want = torch.zeros(T, dtype=torch.uint8, device="cuda")
    want[24::25] = 1
    assert torch.equal(trunc, want[:, None].expand(T, R))
    n_ep = R * (T // 25)
    assert stats.n_collected_episodes == n_ep
    summ = wl.collector.last_episode_summary
    # sum of the episode returns == sum of the rewards of the steps inside completed episodes (float64 accumulators vs float32 rewards)
    rew_done = st["rew"].view(T, R)[: 25 * (T // 25)].double().sum().item()
    assert abs(summ["ret_sum"] - rew_done) <= 1e-6 * abs(rew_done) + 1e-3
    assert summ["len_sum"] == 25 * n_ep and summ["len_min"] == 25 and summ["len_max"] == 25
    # the three agents of a world share the reward (simple_spread, local_ratio 0.5 on a shared global term: rows differ, worlds do not repeat)
    assert torch.isfinite(st["rew"]).all() and torch.isfinite(st["obs"]).all()
    act = st["act"].view(T, R)
    assert int(act.min()) >= 0 and int(act.max()) <= 4
    counts = torch.bincount(act.view(-1).long(), minlength=5).double() / B
    assert float(counts.min()) > 0.1  # a fresh orthogonal policy samples every action
    # obs chain: obs[t + 1] == obs_next[t] wherever the episode went on
    if "obs_next" in st:
        on, ob = st["obs_next"].view(T, R, -1), st["obs"].view(T, R, -1)
        keep = (done[:-1] == 0)
        assert torch.equal(on[:-1][keep], ob[1:][keep])
    # ---- update: accounting, zero learning rate ----
    before = wl.net.flat.detach().clone()
    ts = wl.alg.update(buf, args.batch_size, args.repeat)
    n_mb = B // args.batch_size
    assert ts.gradient_steps == args.repeat * n_mb and int(wl.net.adam_step.item()) == args.repeat * n_mb
    assert torch.equal(wl.net.flat.detach(), before)  # lr = 0
    seq = wl.alg.last_sequences
    for k in ("loss", "clip", "vf", "ent"):
        assert len(seq[k]) == args.repeat * n_mb and np.isfinite(seq[k]).all()
    # with frozen parameters both epochs see the same data in a different order: the epoch means of every loss term agree
    # (value / entropy terms exactly up to summation order; the clipped surrogate too, since ratio == 1 +- rounding)
    for k in ("vf", "ent", "clip"):
        a, b = np.mean(seq[k][:n_mb]), np.mean(seq[k][n_mb:])
        assert abs(a - b) <= 2e-3 * max(1.0, abs(a)), (k, a, b)
    assert 1.55 < np.mean(seq["ent"]) <= np.log(5) + 1e-3  # near-uniform fresh policy
    # ---- the epoch's slot table is a permutation of the valid storage slots ----
    slots = wl.alg._work["slots"]
    assert slots.numel() == B
    s_sorted = torch.sort(slots.long()).values
    assert torch.equal(s_sorted, torch.arange(B, device="cuda"))
    # advantage statistics of the last epoch: every minibatch mean is the mean of its rows' advantages
    adv = wl.alg._work["adv"].view(-1)
    mbs = wl.alg._work["mb_stats"][(args.repeat - 1) * n_mb:]
    for m in (0, n_mb // 2, n_mb - 1):
        rows = slots[m * args.batch_size: (m + 1) * args.batch_size].long()
        a = adv[rows].double()
        assert abs(float(mbs[m, 0]) - a.mean().item()) < 1e-5
        assert abs(float(mbs[m, 1]) - 1.0 / (a.std().item() + 1e-8)) < 1e-4 * float(mbs[m, 1])
