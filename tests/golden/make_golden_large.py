"""C2-scale PPO fixtures from the UNMODIFIED reference (tests/golden/ppo_large.npz): R = 4096 rows, T = 32 steps, minibatches of
4096 rows, 2 epochs.  Inputs come from a frozen numpy RandomState stream (tests/_ppo_golden.py:large_inputs) and are NOT stored:
the file holds the initial / final parameters, every minibatch's losses and a strided sample of v_s / returns / adv / logp_old.
Purpose: measure the tcgen05 engine's error against the reference at a realistic minibatch size (tests/test_gpu_tc.py).

    python tests/golden/make_golden_large.py          # only where /root/reference exists; never on the GPU box
"""
from __future__ import annotations

import os
import sys

import numpy as np

from make_golden import save  # noqa: E402  (sets sys.path for the reference + shims)
from make_golden_more import _flat_params, _ref_ppo

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _ppo_golden import large_inputs  # noqa: E402

CASES = [
    dict(R=4096, T=32, O=18, H=64, nA=5, bs=4096, repeat=2, in_seed=11, kw=dict()),
    dict(R=4096, T=32, O=18, H=64, nA=5, bs=4096, repeat=2, in_seed=12,
         kw=dict(max_grad_norm=0.5, value_clip=True, dual_clip=3.0, advantage_normalization=False, lr=1e-3)),
]


def gen() -> None:
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.data.stats import SequenceSummaryStats
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    out: dict[str, np.ndarray] = {}
    for ci, c in enumerate(CASES):
        R, T, O, H, nA = c["R"], c["T"], c["O"], c["H"], c["nA"]
        alg, ac = _ref_ppo(O, H, nA, seed=40 + ci, softmax_probs=False, **dict(c["kw"]))
        pre = f"p{ci}_"
        for k, v in _flat_params(ac).items():
            out[pre + k] = v
        x = large_inputs(c["in_seed"], R, T, O, nA)
        out[pre + "checksum"] = x["checksum"]
        buf = VectorReplayBuffer(R * T, R)
        for t in range(T):
            buf.add(Batch(obs=x["obs"][t], act=x["act"][t], rew=x["rew"][t].astype(np.float64), terminated=x["term"][t], truncated=x["trunc"][t],
                          obs_next=x["obs_next"][t], info={}))
        seqs: list[list[float]] = []
        orig = SequenceSummaryStats.from_sequence.__func__

        def spy(cls, seq, _orig=orig, _seqs=seqs):
            _seqs.append([float(v) for v in seq])
            return _orig(cls, seq)

        SequenceSummaryStats.from_sequence = classmethod(spy)
        try:
            np.random.seed(70 + ci)
            with policy_within_training_step(alg.policy):
                batch, indices = buf.sample(0)
                assert np.array_equal(indices, np.arange(R * T))
                batch = alg._preprocess_batch(batch, buf, indices)
                sidx = np.arange(0, R * T, 97)
                out[pre + "sample_idx"] = sidx
                for k in ("v_s", "returns", "adv", "logp_old"):
                    out[pre + k] = batch[k].detach().cpu().numpy()[sidx].copy()
                with torch_train_mode(alg):
                    alg._update_with_batch(batch, c["bs"], c["repeat"])
        finally:
            SequenceSummaryStats.from_sequence = classmethod(orig)
        out[pre + "loss_seq"], out[pre + "clip_seq"], out[pre + "vf_seq"], out[pre + "ent_seq"] = (np.array(s) for s in seqs[:4])
        for k, v in _flat_params(ac).items():
            out[pre + "after_" + k] = v
        kw = c["kw"]
        out[pre + "cfg"] = np.array([R, T, O, H, nA, c["bs"], c["repeat"], 70 + ci, c["in_seed"]], np.int64)
        out[pre + "hyper"] = np.array([kw.get("eps_clip", 0.2), kw.get("dual_clip") or 0.0, kw.get("vf_coef", 0.5), kw.get("ent_coef", 0.01),
                                       float(kw.get("value_clip", False)), float(kw.get("advantage_normalization", True)),
                                       kw.get("max_grad_norm") or 0.0, kw.get("lr", 3e-4), 0.99, 0.95])
    out["n_cases"] = np.array(len(CASES))
    save("ppo_large.npz", **out)


if __name__ == "__main__":
    gen()
