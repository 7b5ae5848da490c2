"""Actor-critic MLP whose parameters live in ONE flat float32 device buffer consumed by the
libtsm_b200 kernels (csrc/mlp.cu).

Stands where the reference builds  Net(state_shape, hidden_sizes=[H, H])  shared between
DiscreteActor(preprocess_net=net, action_shape) and DiscreteCritic(preprocess_net=net)
(tianshou/utils/net/common.py:286-369, discrete.py:27-124, test/discrete/test_ppo_discrete.py:91-105),
or the separate DecentralizedActor / CentralizedCritic 3-layer MLPs of ctde.py:346-414.

Flat layout (torch.nn.Linear convention weight[out][in]):
    W1[H][in] b1[H] W2[H][H] b2[H] W3[out][H] b3[out],  out = n_act (+ 1 value row when shared).
"""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np
import torch
from torch import nn

from .. import _lib

MODES = {"raw": 0, "sample": 1, "mode": 2, "eval": 3}


class ActorCriticNet(nn.Module):
    def __init__(self, in_dim: int, hidden: int = 64, n_act: int = 5, with_value: bool = True, n_value_only: int = 0,
                 device: str | torch.device = "cuda", seed: int | None = None, orthogonal: bool = True,
                 engine: str = "fp32", softmax_output: bool = False, gaussian: bool = False, unbounded: bool = True,
                 max_action: float = 1.0, conditioned_sigma: bool = False) -> None:
        """engine: "fp32" = exact CUDA-core kernels (csrc/mlp.cu); "tc" = tcgen05 tensor-core kernels with fp16 operands and
        fp32 accumulation (csrc/mlp_tc.cu; hidden 64, in_dim <= 31, out_dim <= 8); "tcw" = the wide tcgen05 engine of the CTDE
        nets (csrc/mlp_tcw.cu; hidden 128, in_dim <= 1023, out_dim <= 16, separate actor / value-only critic)."""
        super().__init__()
        if engine not in ("fp32", "tc", "tcw"):
            raise ValueError("engine must be 'fp32', 'tc' or 'tcw'")
        # gaussian: the n_act outputs are the means of a diagonal Gaussian and `log_sigma` (n_act more parameters behind b3, zero
        # initialised) its state-independent log standard deviations: ContinuousActorProbabilistic(conditioned_sigma=False)
        # (utils/net/continuous.py:196-247; `unbounded` / `max_action` as there, except that unbounded defaults to True here).
        # Exact fp32 kernels only.
        if gaussian and (engine != "fp32" or softmax_output or n_value_only):
            raise ValueError("gaussian=True needs engine 'fp32', an actor head and no softmax_output")
        self.gaussian = bool(gaussian)
        # conditioned_sigma=True (continuous.py:228-233,242): n_act more head outputs are the per-row log standard deviations
        # (clamped to [-20, 2]); there is no log_sigma parameter then.  Head rows: [mu | log sigma | value].
        self.cond_sigma = bool(conditioned_sigma)
        if self.cond_sigma and not gaussian:
            raise ValueError("conditioned_sigma=True needs gaussian=True")
        # unbounded=False (the reference's default, continuous.py:222,239-241): mean = max_action * tanh(head output)
        self.mu_bound = float(max_action) if (self.gaussian and not unbounded) else 0.0
        if self.gaussian and not unbounded and not max_action > 0:
            raise ValueError("max_action must be > 0")
        if engine == "tcw" and (hidden != 128 or with_value and not n_value_only):
            raise ValueError("engine 'tcw' (csrc/mlp_tcw.cu) runs separate hidden-128 nets: an actor (with_value=False) or a value-only critic")
        self.engine = engine
        # DiscreteActor(softmax_output=True) + Categorical(probs=...) (utils/net/discrete.py:41,85-86): the kernels keep working on
        # logits and evaluate log-probabilities / entropy as torch does for the probs parameterisation
        self.softmax_output = bool(softmax_output)
        self.in_dim, self.hidden, self.n_act = int(in_dim), int(hidden), int(n_act)
        if n_value_only:  # critic-only net: out_dim value heads, no logits
            self.n_act, self.out_dim, self.value_index = 0, int(n_value_only), 0
        else:
            heads = self.n_act * (2 if conditioned_sigma else 1)
            self.out_dim = heads + (1 if with_value else 0)
            self.value_index = heads if with_value else -1
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.TsmError("ActorCriticNet needs a CUDA device (no CPU fallback)")
        L = _lib.lib()
        self.n_sigma = self.n_act if (self.gaussian and not self.cond_sigma) else 0
        n = int(L.tsm_net_param_count(self.in_dim, self.hidden, self.out_dim)) + self.n_sigma
        self.n_params = n
        self.flat = nn.Parameter(torch.zeros(n, dtype=torch.float32, device=self.device))
        self.grads = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.adam_m = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.adam_v = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.adam_step = torch.zeros(1, dtype=torch.int64, device=self.device)
        # engine "tc": fp16 operand image of the parameters, kept current by the optimizer kernels; any host-side
        # write to `flat` (tracked through its version counter) triggers tsm_net_pack before the next kernel call
        self.tc_image = torch.zeros(self._image_bytes(), dtype=torch.uint8, device=self.device) if engine != "fp32" else None
        self._ximg: torch.Tensor | None = None  # engine "tcw": packed fp16 input image of the last forward
        self._packed_version = -1
        self.kernel_updates = 0  # optimizer steps taken by kernels (they do not bump flat._version): see param_version
        H, I, O = self.hidden, self.in_dim, self.out_dim
        sizes = [("W1", (H, I)), ("b1", (H,)), ("W2", (H, H)), ("b2", (H,)), ("W3", (O, H)), ("b3", (O,))]
        if self.n_sigma:
            sizes.append(("log_sigma", (self.n_sigma,)))
        self._views: dict[str, tuple[int, tuple[int, ...]]] = {}
        off = 0
        for k, shp in sizes:
            self._views[k] = (off, shp)
            off += int(np.prod(shp))
        assert off == n
        self._desc = _lib.TsmNet()
        self._fill_desc()
        if orthogonal:
            g = torch.Generator().manual_seed(0 if seed is None else seed)
            with torch.no_grad():
                for k in ("W1", "W2", "W3"):
                    w = torch.empty(self._views[k][1])
                    nn.init.orthogonal_(w, generator=g)  # test_ppo_discrete.py:101-105
                    self.view(k).copy_(w)

    def gaussian_mean(self, raw: torch.Tensor) -> torch.Tensor:
        """Means of the Gaussian head from the raw outputs [n, >= n_act] of forward(mode="raw") (continuous.py:239-241)."""
        mu = raw[:, : self.n_act]
        return self.mu_bound * torch.tanh(mu) if self.mu_bound > 0 else mu

    def gaussian_log_sigma(self, raw: torch.Tensor) -> torch.Tensor:
        """Log standard deviations of the Gaussian head: the state-independent parameter [n_act], or per row [n, n_act] from the
        raw outputs with conditioned_sigma=True (clamped like continuous.py:242)."""
        if self.cond_sigma:
            return raw[:, self.n_act: 2 * self.n_act].clamp(-20.0, 2.0)
        return self.view("log_sigma")

    # -- plumbing ---------------------------------------------------------------------------
    ENGINES = {"fp32": 0, "tc": 1, "tcw": 2}

    def _image_bytes(self) -> int:
        return _lib.TC_IMAGE_BYTES if self.engine == "tc" else int(_lib.lib().tsm_tcw_image_bytes(self.in_dim))

    def _fill_desc(self) -> None:
        d = self._desc
        d.in_dim, d.hidden, d.out_dim, d.n_act, d.value_index, d.n_params = (self.in_dim, self.hidden, self.out_dim, self.n_act,
                                                                           self.value_index, self.n_params)
        d.engine = self.ENGINES[self.engine]
        d.params, d.grads = self.flat.data_ptr(), self.grads.data_ptr()
        d.adam_m, d.adam_v, d.adam_step = self.adam_m.data_ptr(), self.adam_v.data_ptr(), self.adam_step.data_ptr()
        d.tc_image = self.tc_image.data_ptr() if self.tc_image is not None else None
        xs = getattr(self, "_x_scratch", None)
        d.x_scratch, d.x_scratch_bytes = (xs.data_ptr(), xs.numel()) if xs is not None else (None, 0)
        d.probs_param = int(self.softmax_output)
        d.n_sigma = int(self.n_sigma)
        d.mu_bound = float(self.mu_bound)
        d.cond_sigma = int(self.cond_sigma)

    @property
    def desc(self) -> Any:
        if (self._desc.params != self.flat.data_ptr() or self._desc.grads != self.grads.data_ptr()
                or self._desc.engine != self.ENGINES[self.engine]):
            self._fill_desc()
        if self.engine != "fp32" and self.flat._version != self._packed_version:
            if self.tc_image is None:
                self.tc_image = torch.zeros(self._image_bytes(), dtype=torch.uint8, device=self.device)
                self._fill_desc()
            _lib.check(_lib.lib().tsm_net_pack(C.byref(self._desc), _lib.stream_ptr()))
            self._packed_version = self.flat._version
        return C.byref(self._desc)

    def snapshot(self) -> tuple:
        """Parameters and optimizer state as device copies (what restore() takes): data-parallel updates keep one so that an
        update whose gradient exchange lost a peer can be undone."""
        return (self.flat.detach().clone(), self.adam_m.clone(), self.adam_v.clone(), self.adam_step.clone(), self.kernel_updates)

    def restore(self, snap: tuple) -> None:
        with torch.no_grad():
            self.flat.copy_(snap[0])  # (bumps flat's version: the next kernel call rebuilds the fp16 operand image)
            self.adam_m.copy_(snap[1])
            self.adam_v.copy_(snap[2])
            self.adam_step.copy_(snap[3])
            self.grads.zero_()
        self.kernel_updates = snap[4]

    @property
    def param_version(self) -> tuple[int, int]:
        """Changes whenever the parameters do: host-side writes (torch's version counter) or kernel optimizer steps."""
        return (int(self.flat._version), int(self.kernel_updates))

    def view(self, name: str) -> torch.Tensor:
        off, shp = self._views[name]
        return self.flat.detach()[off: off + int(np.prod(shp))].view(shp)  # shares flat's version counter

    def named_views(self) -> dict[str, torch.Tensor]:
        return {k: self.view(k) for k in self._views}

    def load_arrays(self, W1: Any, b1: Any, W2: Any, b2: Any, W3: Any, b3: Any) -> "ActorCriticNet":
        with torch.no_grad():
            for k, a in zip(("W1", "b1", "W2", "b2", "W3", "b3"), (W1, b1, W2, b2, W3, b3)):
                self.view(k).copy_(torch.as_tensor(np.asarray(a), dtype=torch.float32).reshape(self._views[k][1]))
        return self

    def load_reference_state_dict(self, sd: dict[str, Any]) -> "ActorCriticNet":
        """Copy weights from a reference ActorCritic(actor, critic).state_dict() (shared Net trunk)."""
        pick = lambda suf, root: next(v for k, v in sd.items() if k.startswith(root) and k.endswith(suf))  # noqa: E731
        W3 = [pick("last.model.0.weight", "actor")] + ([pick("last.model.0.weight", "critic")] if self.value_index >= 0 else [])
        b3 = [pick("last.model.0.bias", "actor")] + ([pick("last.model.0.bias", "critic")] if self.value_index >= 0 else [])
        return self.load_arrays(pick("model.model.0.weight", "actor.preprocess"), pick("model.model.0.bias", "actor.preprocess"),
                                pick("model.model.2.weight", "actor.preprocess"), pick("model.model.2.bias", "actor.preprocess"),
                                np.concatenate([np.asarray(w) for w in W3], 0), np.concatenate([np.asarray(b) for b in b3], 0))

    def reset_optimizer(self) -> None:
        self.grads.zero_(); self.adam_m.zero_(); self.adam_v.zero_(); self.adam_step.zero_()

    def ensure_scratch(self, n_rows: int) -> None:
        """engine "tcw": scratch image for calls that hand the library float32 rows (tsm_collect_lockstep, tsm_net_forward)."""
        need = self.x_bytes(n_rows + 128)
        xs = getattr(self, "_x_scratch", None)
        if xs is None or xs.numel() < need:
            self._x_scratch = torch.zeros(need, dtype=torch.uint8, device=self.device)
            self._fill_desc()

    # -- engine "tcw": packed fp16 tile images of the inputs (tsm_tcw_pack_x) -------------------------------------------------
    def x_bytes(self, n_rows: int) -> int:
        return int(_lib.lib().tsm_tcw_x_bytes(self.in_dim, int(n_rows)))

    def pack_x(self, x: torch.Tensor, idx: torch.Tensor | None = None, out: torch.Tensor | None = None, n_rows: int | None = None) -> torch.Tensor:
        """x: float32 [*, ld] rows on the device (ld >= in_dim); idx: optional int64 source row per packed row."""
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        x2 = x.view(-1, x.shape[-1])
        n = int(n_rows if n_rows is not None else (idx.numel() if idx is not None else x2.shape[0]))
        need = self.x_bytes(n)
        if out is None:
            if self._ximg is None or self._ximg.numel() < need:
                self._ximg = torch.empty(need, dtype=torch.uint8, device=self.device)
            out = self._ximg
        assert out.numel() >= need
        _lib.check(_lib.lib().tsm_tcw_pack_x(_lib.ptr(x2), int(x2.shape[1]), self.in_dim, _lib.ptr(idx), n, _lib.ptr(out), _lib.stream_ptr()))
        return out

    # -- forward ----------------------------------------------------------------------------
    def forward(self, obs: torch.Tensor, mode: str = "raw", act: torch.Tensor | None = None, seed: int = 0, counter: int = 0,
                rows: torch.Tensor | None = None, scatter: bool = False, out: dict[str, torch.Tensor] | None = None,
                want_logits: bool = False, want_entropy: bool = False) -> dict[str, torch.Tensor]:
        """One kernel launch (tsm_net_forward).  obs: [n_total, in_dim] float32 on the device.
        Returns a dict with `out` (mode raw) or `act`, `logp`, `value` (+ `logits`, `entropy`)."""
        assert obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous()
        x = obs.view(-1, self.in_dim)
        n = int(rows.numel()) if rows is not None else x.shape[0]
        n_out = x.shape[0] if (scatter and rows is not None) else n
        m = MODES[mode]
        o = dict(out or {})
        dev = self.device
        if m == 0:
            if "out" not in o and not (self.value_index >= 0 and "value" in o):
                o["out"] = torch.empty(n_out, self.out_dim, dtype=torch.float32, device=dev)
        else:
            if m == 3:
                assert act is not None
                o["act"] = act.to(torch.int32).contiguous().view(-1)
            elif "act" not in o:
                o["act"] = torch.empty(n_out, dtype=torch.int32, device=dev)
            if "logp" not in o:
                o["logp"] = torch.empty(n_out, dtype=torch.float32, device=dev)
            if self.value_index >= 0 and "value" not in o:
                o["value"] = torch.empty(n_out, dtype=torch.float32, device=dev)
            if want_logits and "out" not in o:
                o["out"] = torch.empty(n_out, self.n_act, dtype=torch.float32, device=dev)
            if want_entropy and "entropy" not in o:
                o["entropy"] = torch.empty(n_out, dtype=torch.float32, device=dev)
        P = _lib.ptr
        if self.engine == "tcw":
            if rows is not None:
                raise NotImplementedError("engine 'tcw': row subsets are passed as the index list of pack_x")
            ximg = self.pack_x(x)
            _lib.check(_lib.lib().tsm_tcw_forward(self.desc, P(ximg), n, m, int(seed), int(counter), 0, P(o.get("out")), P(o.get("act")),
                                                  P(o.get("logp")), P(o.get("entropy")), _lib.stream_ptr()))
            if m != 0 and "out" in o:
                o["logits"] = o["out"]
            return o
        _lib.check(_lib.lib().tsm_net_forward(self.desc, P(x), P(rows), n, m, int(scatter), int(seed), int(counter),
                                              P(o.get("out")), P(o.get("act")), P(o.get("logp")), P(o.get("value")),
                                              P(o.get("entropy")), _lib.stream_ptr()))
        if m != 0 and "out" in o:
            o["logits"] = o["out"]
        return o
