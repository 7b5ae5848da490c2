"""Policy.map_action on the device (reference tianshou/algorithm/algorithm_base.py:255-288): bounding + scaling of raw network
outputs into a Box action space before they go to the environment (the buffer keeps the raw action).  Identity for Discrete."""
from __future__ import annotations

from typing import Any

import numpy as np
import torch

from .. import _lib

_BOUND = {None: 0, "": 0, "clip": 1, "tanh": 2}


class BoxActionMapper:
    """`action_scaling` / `action_bound_method` of the reference Policy constructor for a Box(low, high) action space."""

    def __init__(self, low: Any, high: Any, action_scaling: bool = True, action_bound_method: str | None = "clip",
                 device: str | torch.device = "cuda") -> None:
        if action_bound_method not in _BOUND:
            raise ValueError("action_bound_method must be None, 'clip' or 'tanh'")
        self.device = torch.device(device)
        self.low = torch.as_tensor(np.asarray(low, dtype=np.float32).reshape(-1), device=self.device)
        self.high = torch.as_tensor(np.asarray(high, dtype=np.float32).reshape(-1), device=self.device)
        self.action_scaling, self.action_bound_method = bool(action_scaling), action_bound_method
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)

    def map_action(self, act: Any, check: bool = True) -> Any:
        host = not (isinstance(act, torch.Tensor) and act.is_cuda)
        a = torch.as_tensor(np.asarray(act) if host else act, device=self.device).to(torch.float32).contiguous()
        dim = int(self.low.numel())
        x = a.view(-1, dim)
        out = torch.empty_like(x)
        self._flag.zero_()
        _lib.check(_lib.lib().tsm_map_action(_lib.ptr(x), x.shape[0], dim, _BOUND[self.action_bound_method], int(self.action_scaling),
                                             _lib.ptr(self.low), _lib.ptr(self.high), _lib.ptr(out), _lib.ptr(self._flag), _lib.stream_ptr()))
        if check and self.action_scaling and int(self._flag.item()):
            raise AssertionError("action scaling only accepts raw action range = [-1, 1]")  # algorithm_base.py:283-285
        out = out.view(a.shape)
        return out.cpu().numpy() if host else out
