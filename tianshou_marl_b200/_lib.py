"""ctypes binding of libtsm_b200.so (C ABI: include/tsm_b200.h).

The product path has NO CPU fallback: if the shared library is missing or a kernel call
fails, this module raises.  Build it with `python -c "import __graft_entry__ as g; g.build()"`
or `make -C tianshou_marl_b200/csrc`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TSM_B200_LIB") or os.path.join(_HERE, "libtsm_b200.so")  # override: instrumented builds (tools/)

c_i64, c_i32, c_f64, c_vp, c_sz = C.c_int64, C.c_int, C.c_double, C.c_void_p, C.c_size_t


class TsmError(RuntimeError):
    pass


class TsmBuf(C.Structure):
    """struct tsm_buf (include/tsm_b200.h)."""

    _fields_ = [
        ("n_bufs", c_i64), ("size", c_i64), ("obs_dim", c_i64),
        ("ins", c_vp), ("sz", c_vp), ("last_index", c_vp), ("ep_len", c_vp), ("ep_start", c_vp), ("ep_ret", c_vp),
        ("obs", c_vp), ("obs_next", c_vp), ("act", c_vp), ("rew", c_vp),
        ("terminated", c_vp), ("truncated", c_vp), ("done", c_vp), ("logp", c_vp), ("value", c_vp),
    ]


class TsmMpe(C.Structure):
    """struct tsm_mpe (include/tsm_b200.h)."""

    _fields_ = [
        ("scenario", C.c_int32), ("n_agents", C.c_int32), ("n_landmarks", C.c_int32), ("n_adversaries", C.c_int32),
        ("obs_dim", C.c_int32), ("max_cycles", C.c_int32), ("pos_first", C.c_int32), ("local_ratio", C.c_float),
        ("n_envs", c_i64), ("seed", C.c_uint64),
        ("pos", c_vp), ("vel", c_vp), ("lm", c_vp), ("steps", c_vp), ("episode", c_vp),
    ]


TC_IMAGE_BYTES = 123648  # TSM_TC_IMAGE_BYTES
PPO_RECORD_BYTES = 128  # TSM_PPO_RECORD_BYTES (96 used, 128-byte pitch)


class TsmNet(C.Structure):
    """struct tsm_net (include/tsm_b200.h)."""

    _fields_ = [
        ("in_dim", C.c_int32), ("hidden", C.c_int32), ("out_dim", C.c_int32), ("n_act", C.c_int32),
        ("value_index", C.c_int32), ("engine", C.c_int32), ("n_params", c_i64),
        ("params", c_vp), ("grads", c_vp), ("adam_m", c_vp), ("adam_v", c_vp), ("adam_step", c_vp),
        ("tc_image", c_vp), ("x_scratch", c_vp), ("x_scratch_bytes", c_i64), ("probs_param", C.c_int32), ("n_sigma", C.c_int32),
        ("mu_bound", C.c_float), ("cond_sigma", C.c_int32),
    ]


class TsmPeers(C.Structure):
    """struct tsm_peers (include/tsm_b200.h)."""

    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("recv", c_vp * 8), ("flags", c_vp * 8), ("status", c_vp), ("ll", c_vp * 8)]


class TsmRollout(C.Structure):
    """struct tsm_rollout (include/tsm_b200.h)."""

    _fields_ = [
        ("obs", c_vp), ("act", c_vp), ("rew", c_vp), ("terminated", c_vp), ("truncated", c_vp), ("obs_next", c_vp), ("logp", c_vp),
        ("value", c_vp), ("carry", c_vp), ("size", c_i64), ("n_rows", c_i64), ("obs_dim", c_i64), ("slot", c_i64),
        ("ep_ret", c_vp), ("ep_len", c_vp), ("acc", c_vp), ("list_ret", c_vp), ("list_len", c_vp), ("list_count", c_vp),
        ("list_cap", c_i64),
    ]


class TsmPpoBatch(C.Structure):
    """struct tsm_ppo_batch (include/tsm_b200.h)."""

    _fields_ = [
        ("obs", c_vp), ("act", c_vp), ("adv", c_vp), ("ret", c_vp), ("logp_old", c_vp), ("v_old", c_vp), ("indices", c_vp),
        ("size", c_i64), ("n_rows", c_i64), ("start", c_i64), ("len", c_i64), ("perm_key", C.c_uint64), ("perm_n", c_i64),
        ("records", c_vp), ("slots", c_vp), ("act_f", c_vp),
    ]


class TsmPpoCfg(C.Structure):
    """struct tsm_ppo_cfg (include/tsm_b200.h)."""

    _fields_ = [
        ("eps_clip", C.c_float), ("dual_clip", C.c_float), ("vf_coef", C.c_float), ("ent_coef", C.c_float),
        ("value_clip", C.c_int32), ("adv_norm", C.c_int32), ("plain_pg", C.c_int32), ("value_heads", C.c_int32),
    ]


class TsmTcwTiles(C.Structure):
    """struct tsm_tcw_tiles (include/tsm_b200.h)."""

    _fields_ = [("ximg", c_vp), ("n_rows", c_i64), ("super_ids", c_vp), ("sub", C.c_int32), ("tile0", c_i64), ("n_tiles", c_i64)]


class TsmTcwLoss(C.Structure):
    """struct tsm_tcw_loss (include/tsm_b200.h)."""

    _fields_ = [("kind", C.c_int32), ("act", c_vp), ("adv", c_vp), ("logp_old", c_vp), ("target", c_vp), ("v_old", c_vp), ("weight", c_vp),
                ("heads", C.c_int32)]


class TsmTcwEpoch(C.Structure):
    """struct tsm_tcw_epoch (include/tsm_b200.h)."""

    _fields_ = [("actor", c_vp), ("critic", c_vp), ("ximg_actor", c_vp), ("ximg_critic", c_vp), ("n_rows", c_i64), ("n_agents", C.c_int32),
                ("super_ids", c_vp), ("n_super", c_i64), ("mb_super", c_i64), ("act", c_vp), ("adv", c_vp), ("logp_old", c_vp), ("ret", c_vp),
                ("v_old", c_vp), ("h_mb_rows", c_vp), ("mb_stats", c_vp), ("sums_actor", c_vp), ("sums_critic", c_vp),
                ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("max_grad_norm", C.c_float),
                ("step0", c_i64), ("adam_ws", c_vp), ("adam_seq0", C.c_uint32), ("peers", c_vp), ("xseq0", C.c_uint32)]


class TsmTcwPeers(C.Structure):
    """struct tsm_tcw_peers (include/tsm_b200.h)."""

    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("arena", c_vp * 8), ("flags", c_vp * 8), ("stride", c_i64), ("net_off", c_i64 * 2),
                ("gsum", c_vp), ("status", c_vp)]


c_f32, c_u64 = C.c_float, C.c_uint64

_SIGS = {
    "tsm_abi_version": (c_i32, []),
    "tsm_last_error": (C.c_char_p, []),
    "tsm_device_info": (c_i32, [c_i32, c_vp, c_vp, c_vp]),
    "tsm_gae_flat_workspace_bytes": (c_sz, [c_i64]),
    "tsm_gae_flat": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_i64, c_i64, c_f64, c_f64,
                             c_vp, c_vp, c_vp, c_sz, c_vp]),
    "tsm_gae_flat_static_max_n": (c_i64, []),
    "tsm_gae_flat_static": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_i64, c_i64, c_f64, c_f64, c_vp, c_vp, c_vp]),
    "tsm_gae_tm": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_f64, c_f64, c_vp, c_vp, c_vp]),
    "tsm_gae_tm_scaled": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_f64, c_f64, c_f64, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "tsm_values_next_prepare": (c_i32, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "tsm_buf_reset": (c_i32, [c_vp, c_i32, c_vp]),
    "tsm_buf_add": (c_i32, [c_vp] * 10 + [c_i64] + [c_vp] * 5),
    "tsm_buf_indices_workspace_bytes": (c_sz, [c_i64]),
    "tsm_buf_indices_all": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "tsm_buf_prev": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "tsm_buf_next": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "tsm_buf_unfinished": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "tsm_mpe_reset": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "tsm_mpe_step": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "tsm_mpe_step_continuous": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "tsm_random_actions": (c_i32, [c_vp, c_i64, c_i32, C.c_uint64, C.c_uint64, c_vp]),
    "tsm_collect_lockstep": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, C.c_uint64, C.c_uint64, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "tsm_map_action": (c_i32, [c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tsm_collect_lockstep_groups": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_i64, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "tsm_adam_step_dist": (c_i32, [c_vp, c_vp, C.c_uint32, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, c_vp]),
    "tsm_ppo_epoch_dist": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, c_vp, C.c_uint32, c_vp]),
    "tsm_global_state": (c_i32, [c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "tsm_td_target": (c_i32, [c_vp, c_vp, c_i32, c_vp, c_vp, c_f32, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "tsm_net_param_count": (c_i64, [c_i32, c_i32, c_i32]),
    "tsm_net_pack": (c_i32, [c_vp, c_vp]),
    "tsm_ppo_pack_records": (c_i32, [c_vp, c_vp, c_vp, c_vp]),
    "tsm_ppo_pack_records_rows": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "tsm_net_forward": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_u64, c_u64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tsm_ppo_slot_table": (c_i32, [c_vp, c_i64, c_i64, c_vp]),
    "tsm_ppo_adv_stats": (c_i32, [c_vp, c_vp, c_i64, c_f32, c_vp, c_vp]),
    "tsm_ppo_minibatch": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_vp]),
    "tsm_adam_step": (c_i32, [c_vp, c_f32, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, c_vp]),
    "tsm_ppo_epoch": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, c_vp]),
    "tsm_episode_stats": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "tsm_rollout_step": (c_i32, [c_vp] * 14 + [c_i64, c_i64, c_i64, c_vp]),
    "tsm_net_forward_range": (c_i32, [c_vp, c_vp, c_i64, c_i64, c_i32, C.c_uint64, C.c_uint64, c_vp, c_vp, c_vp, c_vp]),
    "tsm_tc_selftest": (c_i32, [c_vp, c_i32, c_vp, c_i32, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "tsm_tcw_image_bytes": (c_i64, [c_i32]),
    "tsm_tcw_x_bytes": (c_i64, [c_i32, c_i64]),
    "tsm_tcw_pack_x": (c_i32, [c_vp, c_i64, c_i32, c_vp, c_i64, c_vp, c_vp]),
    "tsm_tcw_forward": (c_i32, [c_vp, c_vp, c_i64, c_i32, c_u64, c_u64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tsm_tcw_train": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_f32, c_f32, c_vp, c_vp, c_vp]),
    "tsm_tcw_adam": (c_i32, [c_vp, c_i32, c_i64, c_f32, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, C.c_uint32, c_vp, c_vp]),
    "tsm_tcw_adam_peers": (c_i32, [c_vp, c_i32, c_i64, c_f32, c_f32, c_f32, c_f32, c_f32, c_vp, C.c_uint32, c_vp, C.c_uint32, c_vp]),
    "tsm_tcw_ppo_epoch": (c_i32, [c_vp, c_vp, c_vp]),
    "tsm_tcw_adv_stats": (c_i32, [c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_f32, c_vp, c_vp]),
    "tsm_buf_gather": (c_i32, [c_vp, c_vp, c_i64] + [c_vp] * 9 + [c_vp, c_vp, c_i32, c_vp]),
}

_lib = None


def exported_symbols() -> list[str]:
    """Names include/tsm_b200.h declares (kept in sync by tests/test_abi.py)."""
    return list(_SIGS)


def build(verbose: bool = False) -> str:
    """Compile the library in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise TsmError("building libtsm_b200.so failed")
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TsmError(
                f"{LIB_PATH} is missing: the CUDA extension is required (no CPU fallback). "
                "Build it with `make -C tianshou_marl_b200/csrc` or __graft_entry__.build()."
            )
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        if l.tsm_abi_version() != 1:
            raise TsmError("libtsm_b200.so ABI version mismatch")
        _lib = l
    return _lib


def check(rc: int) -> None:
    """Translate a tsm_status into the reference's exception types."""
    if rc == 0:
        return
    msg = lib().tsm_last_error().decode(errors="replace")
    if rc == -1:
        raise ValueError(msg)
    if rc == -4:
        raise NotImplementedError(msg)
    raise TsmError(f"tsm error {rc}: {msg}")


def ptr(t) -> int | None:
    """Device/host pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream


# -- optional stage timer (bench.py installs one; None costs one attribute check per stage) ------
class StageTimer:
    """CUDA-event pairs around stages, recorded on the launching (current) stream."""

    def __init__(self) -> None:
        self.pairs: list = []

    def start(self, name: str):
        import torch

        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        self.pairs.append((name, a, b))
        return b

    def totals(self) -> dict:
        out: dict = {}
        for name, a, b in self.pairs:
            ms, n = out.get(name, (0.0, 0))
            out[name] = (ms + a.elapsed_time(b), n + 1)
        return out


timer: StageTimer | None = None


def t_start(name: str):
    return timer.start(name) if timer is not None else None


def t_stop(tok) -> None:
    if tok is not None:
        tok.record()
