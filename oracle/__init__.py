"""CPU oracle for the tianshou_marl on-policy multi-agent hot path.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
may import it.  Nothing under ``tianshou_marl_b200/`` imports it; the product path fails
loudly when the CUDA library is missing instead of falling back to this code.

It restates, in numpy / plain C, the algorithms of the reference (all ``file:line``
citations are relative to the reference checkout, eric-downes/tianshou_marl):

* ``oracle.gae``     -- ``_gae`` / ``Algorithm.compute_episodic_return`` /
                        ``episode_mc_return_to_go``  (tianshou/algorithm/algorithm_base.py)
* ``oracle.buffer``  -- ``ReplayBuffer`` / ``ReplayBufferManager`` / ``VectorReplayBuffer``
                        index state machine          (tianshou/data/buffer/*.py)
* ``oracle.mpe``     -- PettingZoo 1.24.2 MPE ``simple_spread`` / ``simple_tag`` physics
                        (third-party, NOT in the reference tree -> parity UNPINNED)
* ``oracle.ppo``     -- torch-fp32 restatement of ``PPO._preprocess_batch`` /
                        ``_update_with_batch`` / ``Algorithm.Optimizer.step``
* ``oracle.philox``  -- the counter-based RNG the CUDA kernels use (our own design; the
                        oracle restates it so resets/samples can be compared bit-for-bit)

Pinning status (see DESIGN.md section "Oracle"):
  gae / buffer / collector slots / PPO update / policy managers: PINNED -- checked against
  the reference's own known-answer tests and against fixtures produced by importing the
  unmodified reference in the build container (tests/golden/make_golden.py).
  mpe: PARITY UNPINNED -- pettingzoo is a third-party dependency absent from the reference
  tree and not installable offline; the restatement follows the published 1.24.2 algorithm.
"""
