"""ctypes loader of libasvrl.so (the C-ABI of include/asvrl.h) and torch-tensor call wrappers.

There is NO fallback: ``lib()`` raises if the shared library is missing or a symbol is absent,
and every compute wrapper raises if its tensors are not CUDA tensors.  PyTorch is used for
device memory and streams only; the kernels are the hand-written sm_100a ones in ``csrc/``.
"""
import ctypes as C
import os

import torch

from . import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libasvrl.so")
_LIB = None

u8p, i32p, u32p, u64p, i64p, f32p, vp = (C.c_void_p,) * 7

_SIGNATURES = {
    "asvrl_version": (C.c_int, []),
    "asvrl_last_error_string": (C.c_char_p, []),
    "asvrl_env_default_params": (None, [C.POINTER(_abi.EnvParams)]),
    "asvrl_reset_default_params": (None, [C.c_void_p]),
    "asvrl_env_step": (C.c_int, [C.POINTER(_abi.EnvParams), C.POINTER(_abi.EnvStatePtrs), vp, f32p,
                                 u8p, C.c_uint64, C.c_uint64, C.POINTER(_abi.EnvOutPtrs), C.c_int,
                                 C.c_int, C.c_int, C.c_int, C.c_int64, C.c_uint32, u64p, i32p, vp]),
    "asvrl_env_step_autoreset": (C.c_int, [C.POINTER(_abi.EnvParams), C.POINTER(_abi.EnvStatePtrs), vp,
                                           C.c_uint64, C.c_uint64, C.POINTER(_abi.EnvOutPtrs),
                                           C.POINTER(_abi.EnvAutoReset), C.c_int, C.c_int, C.c_int,
                                           C.c_int, C.c_int64, C.c_uint32, u64p, vp]),
    "asvrl_env_noise": (C.c_int, [C.POINTER(_abi.EnvParams), C.c_uint64, C.c_uint64, C.c_int,
                                  C.c_int, C.c_int, C.c_int64, f32p, vp]),
    "asvrl_env_reset": (C.c_int, [C.c_void_p, C.POINTER(_abi.EnvStatePtrs), u8p, i32p, i32p, i32p,
                                  f32p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int64, u64p, i32p, vp]),
    "asvrl_env_stage_refill": (C.c_int, [C.POINTER(_abi.EnvParams), C.c_void_p, C.POINTER(_abi.EnvStatePtrs),
                                         C.POINTER(_abi.EnvOutPtrs), i32p, i32p, i32p, f32p, i32p, i32p,
                                         u8p, i32p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_int64, vp]),
    "asvrl_replay_push": (C.c_int, [f32p, C.c_int64, u64p, f32p, f32p, f32p, vp, C.c_int, f32p,
                                    f32p, f32p, f32p, u32p, C.c_int, u32p, f32p, f32p, f32p, u8p,
                                    C.c_int, vp]),
    "asvrl_replay_sample": (C.c_int, [f32p, u64p, C.c_int64, i64p, C.c_uint64, C.c_uint64, C.c_int,
                                      f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p, i64p,
                                      u64p, vp]),
    "asvrl_iqn_tau_cos": (C.c_int, [f32p, C.c_uint64, C.c_uint64, C.c_float, C.c_int, C.c_int,
                                    C.c_int, f32p, f32p, u64p, vp]),
    "asvrl_qhuber_fwd_bwd": (C.c_int, [f32p, f32p, f32p, f32p, f32p, C.c_float, C.c_float, C.c_int,
                                       C.c_int, C.c_int, f32p, f32p, f32p, u32p, vp]),
    "asvrl_iqn_scale_relu_fwd": (C.c_int, [f32p, f32p, f32p, C.c_int, C.c_int, C.c_int, f32p, vp]),
    "asvrl_iqn_scale_relu_bwd": (C.c_int, [f32p, f32p, f32p, f32p, C.c_int, C.c_int, C.c_int, f32p, f32p,
                                           f32p, f32p, u32p, vp]),
    "asvrl_iqn_scale_relu_bwd_partials": (C.c_long, [C.c_int, C.c_int]),
    "asvrl_iqn_head_fwd": (C.c_int, [f32p, f32p, f32p, f32p, C.c_int, C.c_int, f32p, vp]),
    "asvrl_iqn_head_bwd": (C.c_int, [f32p, f32p, f32p, f32p, C.c_int, C.c_int, f32p, f32p, f32p, f32p,
                                     f32p, u32p, vp]),
    "asvrl_iqn_head_bwd_partials": (C.c_long, [C.c_int, C.c_int]),
    "asvrl_tower_fwd": (C.c_int, [f32p] * 7 + [C.c_int] * 6 + [f32p, vp]),
    "asvrl_tower_bwd": (C.c_int, [f32p] * 4 + [C.c_int] * 6 + [f32p] * 5 + [u32p, vp]),
    "asvrl_tower_bwd_partials": (C.c_long, [C.c_int] * 4),
    "asvrl_tower_bwd_static": (C.c_int, [f32p] * 4 + [C.c_int] * 6 + [f32p] * 4 + [vp]),
    "asvrl_linear_tf32x3": (C.c_int, [f32p, f32p, f32p, C.c_int, C.c_long, C.c_int, C.c_int, f32p, C.c_void_p]),
    "asvrl_linear_tf32x3_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_actor_head": (C.c_int, [f32p, f32p, f32p, f32p, C.c_float, f32p, C.c_float, C.POINTER(C.c_float * 2),
                                   C.POINTER(C.c_float * 2), C.c_uint64, C.c_uint64, u64p, C.c_int, C.c_int, f32p, vp]),
    "asvrl_policy_forward_workspace": (C.c_long, []),
    "asvrl_policy_forward": (C.c_int, [f32p] * 13 + [C.c_float, f32p, C.c_float, C.POINTER(C.c_float * 2),
                                       C.POINTER(C.c_float * 2), C.c_uint64, C.c_uint64, u64p, C.c_long, C.c_int,
                                       vp, f32p, vp]),
    "asvrl_policy_forward_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_policy_hidden": (C.c_int, [f32p] * 11 + [C.c_long, C.c_int, vp, f32p, vp, vp]),
    "asvrl_action_encoder_fwd": (C.c_int, [f32p, f32p, f32p, C.c_int, C.c_int, f32p, vp]),
    "asvrl_action_encoder_bwd": (C.c_int, [f32p, f32p, f32p, f32p, C.c_int, C.c_int, f32p, f32p, f32p, f32p, u32p, vp]),
    "asvrl_narrow_bwd_partials": (C.c_long, [C.c_int]),
    "asvrl_atan_head_fwd": (C.c_int, [f32p, f32p, f32p, C.c_float, C.c_int, C.c_int, f32p, f32p, vp]),
    "asvrl_atan_head_bwd": (C.c_int, [f32p, f32p, f32p, f32p, C.c_float, C.c_int, C.c_int, f32p, f32p, f32p, f32p, u32p, vp]),
    "asvrl_neg_mean": (C.c_int, [f32p, C.c_long, f32p, vp]),
    "asvrl_c51_head_workspace": (C.c_long, []),
    "asvrl_c51_head": (C.c_int, [vp, f32p, C.c_int, f32p, f32p, f32p, f32p, C.c_long, C.c_int, C.c_int, vp, i32p, f32p, vp]),
    "asvrl_c51_head_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_rainbow_append": (C.c_int, [f32p, f32p, f32p, i32p, f32p, u32p, C.c_int, C.c_long, C.c_long, f32p, i32p, f32p, i32p,
                                      f32p, i64p, i64p, i64p, u32p, vp]),
    "asvrl_c51_greedy": (C.c_int, [f32p] * 5 + [C.c_long, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_long, i32p, f32p, vp]),
    "asvrl_critic_forward_workspace": (C.c_long, []),
    "asvrl_critic_forward_chunk_bytes": (C.c_long, []),
    "asvrl_critic_forward": (C.c_int, [f32p, C.c_uint64, C.c_uint64, u64p, C.c_float] + [f32p] * 10
                             + [C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, f32p, f32p, vp, vp, vp, vp, vp]),
    "asvrl_critic_forward_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_critic_backward_workspace": (C.c_long, []),
    "asvrl_critic_backward": (C.c_int, [f32p, vp, vp, vp] + [f32p] * 6 + [C.c_int, C.c_int, C.c_int, vp, f32p, f32p,
                                                                          vp, vp, vp, vp, f32p, vp]),
    "asvrl_critic_backward_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_critic_wgrad_workspace": (C.c_long, [C.c_int]),
    "asvrl_critic_wgrad": (C.c_int, [vp] * 8 + [f32p, C.c_int, C.c_int] + [f32p] * 8 + [vp]),
    "asvrl_critic_wgrad_status": (C.c_int, [C.POINTER(C.c_int)]),
    "asvrl_clip_adam_step_p2p": (C.c_int, [C.POINTER(_abi.AdamTensors), C.POINTER(_abi.PeerBuckets), f32p, u32p]
                                 + [C.c_float] * 5 + [f32p, f32p, u32p, vp]),
    "asvrl_clip_adam_step": (C.c_int, [C.POINTER(_abi.AdamTensors)] + [C.c_float] * 6 + [f32p, f32p, u32p, vp]),
}
EXPORTED = tuple(_SIGNATURES)


ResetParams = _abi.ResetParams


def lib():
    """The loaded library; raises (never falls back) when it is missing or incomplete."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                f"{SO_PATH} is missing: build it with `python -m "
                "distributional_rl_decision_and_control_b200.csrc.build` (there is no CPU fallback)")
        so = C.CDLL(SO_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(so, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        if so.asvrl_version() != _abi.VERSION:
            raise RuntimeError(f"libasvrl.so version {so.asvrl_version()} != header {_abi.VERSION}")
        _LIB = so
    return _LIB


def check(rc, what):
    if rc != 0:
        msg = lib().asvrl_last_error_string().decode(errors="replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def default_reset_params():
    p = ResetParams()
    lib().asvrl_reset_default_params(C.byref(p))
    return p


def ptr(t, dtype=None):
    """Raw device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libasvrl works on CUDA tensors only (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("tensor must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected {dtype}, got {t.dtype}")
    return t.data_ptr()


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------ typed call wrappers
def state_ptrs(st):
    s = _abi.EnvStatePtrs()
    for n in _abi.STATE_FLOAT_FIELDS:
        setattr(s, n, ptr(st[n], torch.float32))
    s.flags = ptr(st["flags"], torch.int32)
    for n in _abi.STATE_ENV_INT_FIELDS:
        setattr(s, n, ptr(st[n], torch.int32))
    s.obstacles = ptr(st["obstacles"], torch.float32)
    s.cores = ptr(st["cores"], torch.float32)
    return s


def out_ptrs(out):
    o = _abi.EnvOutPtrs()
    o.obs_self = ptr(out["obs_self"], torch.float32)
    o.obs_objects = ptr(out["obs_objects"], torch.float32)
    o.obs_mask = ptr(out["obs_mask"], torch.float32)
    o.reward = ptr(out["reward"], torch.float32)
    o.status = ptr(out["status"], torch.int32)
    o.phi = ptr(out["phi"], torch.float32)
    o.top_ids = ptr(out.get("top_ids"), torch.int32)
    o.detected = ptr(out.get("detected"), torch.int64)
    o.env_status = ptr(out["env_status"], torch.int32)
    o.ended = ptr(out.get("ended"), torch.uint8)
    return o


def new_state(E, RP, OP, CP, device):
    """Zeroed SoA env state in HBM (layout of asvrl_env_state; flags/status are int32 views of
    the header's uint32 words, `detected` an int64 view of uint64)."""
    st = {n: torch.zeros((E, RP), dtype=torch.float32, device=device) for n in _abi.STATE_FLOAT_FIELDS}
    st["flags"] = torch.zeros((E, RP), dtype=torch.int32, device=device)
    for n in _abi.STATE_ENV_INT_FIELDS:
        st[n] = torch.zeros((E,), dtype=torch.int32, device=device)
    st["obstacles"] = torch.zeros((E, max(OP, 1), 4), dtype=torch.float32, device=device)
    st["cores"] = torch.zeros((E, max(CP, 1), 4), dtype=torch.float32, device=device)
    return st


# obs_mask comes last: a host-side caller can leave it on the device (mask[i] = objects[i][4] != 0; the
# reference's step() returns no mask either, RB:65 builds it) and copy only the first 34 words per ASV
HOST_BLOCK_FIELDS = (("obs_self", 7), ("obs_objects", 25), ("reward", 1), ("status", 1), ("obs_mask", 5))
HOST_BLOCK_NO_MASK = 34


def carve_host_block(flat, E, RP):
    """Views of one flat float32 buffer [E*RP*39]: what a host-side caller of step() reads
    (observations, rewards, status) is contiguous, so it crosses PCIe in ONE copy."""
    views, off, n = {}, 0, E * RP
    for name, w in HOST_BLOCK_FIELDS:
        v = flat[off:off + n * w]
        off += n * w
        if name == "status":
            v = v.view(torch.int32)
        shape = {"obs_self": (E, RP, 7), "obs_objects": (E, RP, 5, 5), "obs_mask": (E, RP, 5)}.get(name, (E, RP))
        views[name] = v.view(shape)
    return views


def new_out(E, RP, device, debug=True, pin=False):
    """Step outputs (layout of asvrl_env_out).  ``host_block`` is the flat buffer the
    observation / reward / status tensors are views of."""
    width = sum(w for _, w in HOST_BLOCK_FIELDS)
    kw = dict(pin_memory=True) if pin else dict(device=device)
    flat = torch.zeros((E * RP * width,), dtype=torch.float32, **kw)
    out = carve_host_block(flat, E, RP)
    out["host_block"] = flat
    if pin:
        return out
    out.update({
        "phi": torch.zeros((E, RP), dtype=torch.float32, device=device),
        "env_status": torch.zeros((E,), dtype=torch.int32, device=device),
        "ended": torch.zeros((E,), dtype=torch.uint8, device=device),
    })
    if debug:
        out["top_ids"] = torch.full((E, RP, 5), -1, dtype=torch.int32, device=device)
        out["detected"] = torch.zeros((E, RP), dtype=torch.int64, device=device)
    return out


def env_step(params, st, actions, noise, out, mode, max_obs, max_cores, seed=0, step=0,
             env_offset=0, env_mask=None, st_ptrs=None, o_ptrs=None, counter_dev=None,
             step_per_env=None):
    E, RP = st["x"].shape
    if actions is not None:
        want = torch.int32 if (mode & _abi.M_DISCRETE) else torch.float32
        a_ptr = ptr(actions, want)
    else:
        a_ptr = None
    sp = st_ptrs if st_ptrs is not None else state_ptrs(st)
    op = o_ptrs if o_ptrs is not None else out_ptrs(out)
    rc = lib().asvrl_env_step(C.byref(params), C.byref(sp), a_ptr, ptr(noise, torch.float32),
                              ptr(env_mask, torch.uint8), seed, step, C.byref(op), E, RP, max_obs,
                              max_cores, env_offset, mode, ptr(counter_dev, torch.int64),
                              ptr(step_per_env, torch.int32), stream_ptr())
    check(rc, "asvrl_env_step")
    return out


def env_step_autoreset(params, rparams, st, actions, out, mode, max_obs, max_cores, want_robots,
                       want_cores, want_obs, min_sg, episode_idx, seed=0, step=0, observe_seed=0,
                       term=None, stage_ptrs=None, stage_tag=None, stage_obs=None, refill_in_launch=True,
                       env_offset=0, st_ptrs=None, o_ptrs=None, counter_dev=None):
    """K1 + K1b + first observation in one launch (asvrl_env_step_autoreset).  ``episode_idx`` =
    int32 [E] per-env episode counters (in/out); ``term`` = dict with obs_self / obs_objects /
    obs_mask tensors that receive the terminal rows of ended envs; ``stage_ptrs`` (state_ptrs of a
    second state block) + ``stage_tag`` int32 [E] enable pre-generated next episodes."""
    E, RP = st["x"].shape
    want = torch.int32 if (mode & _abi.M_DISCRETE) else torch.float32
    sp = st_ptrs if st_ptrs is not None else state_ptrs(st)
    op = o_ptrs if o_ptrs is not None else out_ptrs(out)
    ar = _abi.EnvAutoReset()
    ar.reset_params = C.pointer(rparams)
    ar.want_robots = ptr(want_robots, torch.int32)
    ar.want_cores = ptr(want_cores, torch.int32)
    ar.want_obs = ptr(want_obs, torch.int32)
    ar.min_start_goal_dis = ptr(min_sg, torch.float32)
    ar.episode_idx = ptr(episode_idx, torch.int32)
    ar.observe_seed = observe_seed & 0xFFFFFFFFFFFFFFFF
    if stage_ptrs is not None:
        ar.stage = C.pointer(stage_ptrs)
        ar.stage_tag = ptr(stage_tag, torch.int32)
        ar.refill_in_launch = 1 if refill_in_launch else 0
        if stage_obs is not None:  # the staged first observation (asvrl_env_stage_refill's output)
            ar.stage_self = ptr(stage_obs["obs_self"], torch.float32)
            ar.stage_objects = ptr(stage_obs["obs_objects"], torch.float32)
            ar.stage_mask = ptr(stage_obs["obs_mask"], torch.float32)
    if term is not None:
        ar.term_self = ptr(term["obs_self"], torch.float32)
        ar.term_objects = ptr(term["obs_objects"], torch.float32)
        ar.term_mask = ptr(term["obs_mask"], torch.float32)
    rc = lib().asvrl_env_step_autoreset(C.byref(params), C.byref(sp), ptr(actions, want), seed, step,
                                        C.byref(op), C.byref(ar), E, RP, max_obs, max_cores, env_offset,
                                        mode, ptr(counter_dev, torch.int64), stream_ptr())
    check(rc, "asvrl_env_step_autoreset")
    return out


def env_noise(params, seed, step, E, RP, OP, device, env_offset=0):
    out = torch.empty((E, RP, OP + RP, 5), dtype=torch.float32, device=device)
    rc = lib().asvrl_env_noise(C.byref(params), seed, step, E, RP, OP, env_offset, ptr(out),
                               stream_ptr())
    check(rc, "asvrl_env_noise")
    return out


def env_reset(rparams, st, reset_mask, want_robots, want_cores, want_obs, min_sg, seed, episode,
              max_obs, max_cores, env_offset=0, st_ptrs=None, counter_dev=None, episode_per_env=None):
    E, RP = st["x"].shape
    sp = st_ptrs if st_ptrs is not None else state_ptrs(st)
    rc = lib().asvrl_env_reset(C.byref(rparams), C.byref(sp), ptr(reset_mask, torch.uint8),
                               ptr(want_robots, torch.int32), ptr(want_cores, torch.int32),
                               ptr(want_obs, torch.int32), ptr(min_sg, torch.float32), seed, episode,
                               E, RP, max_obs, max_cores, env_offset, ptr(counter_dev, torch.int64),
                               ptr(episode_per_env, torch.int32), stream_ptr())
    check(rc, "asvrl_env_reset")


def env_stage_refill(params, rparams, stage_ptrs, stage_out_ptrs, want_robots, want_cores, want_obs, min_sg,
                     episode_idx, stage_tag, need_ws, episode_ws, seed, observe_seed, E, RP, max_obs,
                     max_cores, env_offset=0, stream=None):
    """asvrl_env_stage_refill on ``stream`` (a torch.cuda.Stream; default: the current one)."""
    sp = stream.cuda_stream if stream is not None else stream_ptr()
    rc = lib().asvrl_env_stage_refill(C.byref(params), C.byref(rparams), C.byref(stage_ptrs),
                                      C.byref(stage_out_ptrs), ptr(want_robots, torch.int32),
                                      ptr(want_cores, torch.int32), ptr(want_obs, torch.int32),
                                      ptr(min_sg, torch.float32), ptr(episode_idx, torch.int32),
                                      ptr(stage_tag, torch.int32), ptr(need_ws, torch.uint8),
                                      ptr(episode_ws, torch.int32), seed, observe_seed & 0xFFFFFFFFFFFFFFFF,
                                      E, RP, max_obs, max_cores, env_offset, sp)
    check(rc, "asvrl_env_stage_refill")


def qhuber_fwd_bwd(z, tau, z_next, r, done, gamma, kappa, workspace, lps=None, grad=None, loss=None):
    B, N = z.shape
    Np = z_next.shape[1]
    dev = z.device
    lps = torch.empty((B,), dtype=torch.float32, device=dev) if lps is None else lps
    grad = torch.empty((B, N), dtype=torch.float32, device=dev) if grad is None else grad
    loss = torch.empty((1,), dtype=torch.float32, device=dev) if loss is None else loss
    rc = lib().asvrl_qhuber_fwd_bwd(ptr(z, torch.float32), ptr(tau, torch.float32),
                                    ptr(z_next, torch.float32), ptr(r, torch.float32),
                                    ptr(done, torch.float32), gamma, kappa, B, N, Np, ptr(lps),
                                    ptr(grad), ptr(loss), ptr(workspace, torch.int32), stream_ptr())
    check(rc, "asvrl_qhuber_fwd_bwd")
    return loss, lps, grad


def iqn_tau_cos(B, N, n_cos, device, tau_in=None, seed=0, counter=0, cvar=1.0, tau_out=None,
                cos_out=None, counter_dev=None):
    tau_out = torch.empty((B, N), dtype=torch.float32, device=device) if tau_out is None else tau_out
    cos_out = torch.empty((B, N, n_cos), dtype=torch.float32, device=device) if cos_out is None else cos_out
    rc = lib().asvrl_iqn_tau_cos(ptr(tau_in, torch.float32), seed, counter, cvar, B, N, n_cos,
                                 ptr(tau_out), ptr(cos_out), ptr(counter_dev, torch.int64), stream_ptr())
    check(rc, "asvrl_iqn_tau_cos")
    return tau_out, cos_out


# ------------------------------------------------------------------ K5: critic trunk glue
_TRUNK_WS = {}


def _trunk_ws(device, n_floats):
    """Per (device, stream) scratch for the deterministic cross-CTA reductions of K5."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _TRUNK_WS.get(key)
    if ws is None or ws[0].numel() < n_floats:
        ws = (torch.empty((max(int(n_floats), 1 << 16),), dtype=torch.float32, device=device),
              ws[1] if ws is not None else torch.zeros((1,), dtype=torch.int32, device=device))
        _TRUNK_WS[key] = ws
    return ws


def iqn_scale_relu_fwd(pre, bias, scale, n_tau):
    M, D = pre.shape
    B = M // n_tau
    out = torch.empty_like(pre)
    rc = lib().asvrl_iqn_scale_relu_fwd(ptr(pre, torch.float32), ptr(bias, torch.float32),
                                        ptr(scale, torch.float32), B, n_tau, D, ptr(out), stream_ptr())
    check(rc, "asvrl_iqn_scale_relu_fwd")
    return out


def iqn_scale_relu_bwd(dout, pre, bias, scale, n_tau, want_dpre=True, want_dscale=True, want_dbias=True):
    M, D = pre.shape
    B = M // n_tau
    dev = pre.device
    dpre = torch.empty_like(pre) if want_dpre else None
    dscale = torch.empty((B, D), dtype=torch.float32, device=dev) if want_dscale else None
    dbias = torch.empty((D,), dtype=torch.float32, device=dev) if want_dbias else None
    part, cnt = _trunk_ws(dev, lib().asvrl_iqn_scale_relu_bwd_partials(B, D)) if want_dbias else (None, None)
    rc = lib().asvrl_iqn_scale_relu_bwd(ptr(dout, torch.float32), ptr(pre, torch.float32),
                                        ptr(bias, torch.float32), ptr(scale, torch.float32), B, n_tau, D,
                                        ptr(dpre), ptr(dscale), ptr(dbias), ptr(part), ptr(cnt), stream_ptr())
    check(rc, "asvrl_iqn_scale_relu_bwd")
    return dpre, dscale, dbias


def iqn_head_fwd(pre, bias, w, b_out):
    M, D = pre.shape
    z = torch.empty((M,), dtype=torch.float32, device=pre.device)
    rc = lib().asvrl_iqn_head_fwd(ptr(pre, torch.float32), ptr(bias, torch.float32), ptr(w, torch.float32),
                                  ptr(b_out, torch.float32), M, D, ptr(z), stream_ptr())
    check(rc, "asvrl_iqn_head_fwd")
    return z


def iqn_head_bwd(dz, pre, bias, w, want_params=True):
    M, D = pre.shape
    dev = pre.device
    dpre = torch.empty_like(pre)
    dw = dbias = dbo = part = cnt = None
    if want_params:
        dw = torch.empty((D,), dtype=torch.float32, device=dev)
        dbias = torch.empty((D,), dtype=torch.float32, device=dev)
        dbo = torch.empty((1,), dtype=torch.float32, device=dev)
        part, cnt = _trunk_ws(dev, lib().asvrl_iqn_head_bwd_partials(M, D))
    rc = lib().asvrl_iqn_head_bwd(ptr(dz, torch.float32), ptr(pre, torch.float32), ptr(bias, torch.float32),
                                  ptr(w, torch.float32), M, D, ptr(dpre), ptr(dw), ptr(dbias), ptr(dbo),
                                  ptr(part), ptr(cnt), stream_ptr())
    check(rc, "asvrl_iqn_head_bwd")
    return dpre, dw, dbias, dbo


def actor_head(pre, bias, w, b_out, atan_scale, eps, lo, hi, seed, counter, counter_dev=None):
    """-> (rows, 2) epsilon-greedy actions; ``eps`` a float or a 0-d CUDA tensor."""
    M, D = pre.shape
    actions = torch.empty((M, 2), dtype=torch.float32, device=pre.device)
    eps_dev, eps_f = (eps, 0.0) if isinstance(eps, torch.Tensor) else (None, float(eps))
    f = torch.float32
    rc = lib().asvrl_actor_head(ptr(pre, f), ptr(bias, f), ptr(w, f), ptr(b_out, f), float(atan_scale),
                                ptr(eps_dev, f), eps_f, C.byref((C.c_float * 2)(*lo)), C.byref((C.c_float * 2)(*hi)),
                                int(seed) & 0xFFFFFFFFFFFFFFFF, int(counter), ptr(counter_dev, torch.int64), M, D,
                                ptr(actions), stream_ptr())
    check(rc, "asvrl_actor_head")
    return actions


# ------------------------------------------------------------------ K8: dense layer on tcgen05 (3 x TF32)
def linear_tf32x3(x, w, bias=None, relu=False, out=None):
    """[relu](x @ w.T [+ bias]) at fp32-GEMM accuracy on the tensor cores; forward only."""
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=x.device)
    f = torch.float32
    rc = lib().asvrl_linear_tf32x3(ptr(x, f), ptr(w, f), ptr(bias, f), int(bool(relu)), M, K, N, ptr(out, f),
                                   stream_ptr())
    check(rc, "asvrl_linear_tf32x3")
    return out


# ------------------------------------------------------------------ K9: fused collection policy forward
_POLICY_WS = {}


def policy_forward(self_obs, objects, mask, w_self, b_self, w_obj, b_obj, w1, b1, w2, b2, w_out, b_out,
                   atan_scale, eps, lo, hi, seed, counter=0, counter_dev=None, repack=True, out=None):
    """(rows, 2) epsilon-greedy actions of the AC-IQN / DDPG actor for every row in one persistent tcgen05
    launch (+ one small weight-pack launch when ``repack``).  ``eps``: float or 0-d CUDA tensor."""
    M = self_obs.shape[0]
    dev = self_obs.device
    ws = _POLICY_WS.get(dev)
    if ws is None:
        ws = _POLICY_WS[dev] = torch.empty((lib().asvrl_policy_forward_workspace(),), dtype=torch.uint8, device=dev)
        repack = True
    if out is None:
        out = torch.empty((M, 2), dtype=torch.float32, device=dev)
    eps_dev, eps_f = (eps, 0.0) if isinstance(eps, torch.Tensor) else (None, float(eps))
    f = torch.float32
    rc = lib().asvrl_policy_forward(
        ptr(self_obs, f), ptr(objects, f), ptr(mask, f), ptr(w_self, f), ptr(b_self, f), ptr(w_obj, f), ptr(b_obj, f),
        ptr(w1, f), ptr(b1, f), ptr(w2, f), ptr(b2, f), ptr(w_out, f), ptr(b_out, f), float(atan_scale),
        ptr(eps_dev, f), eps_f, C.byref((C.c_float * 2)(*lo)), C.byref((C.c_float * 2)(*hi)),
        int(seed) & 0xFFFFFFFFFFFFFFFF, int(counter), ptr(counter_dev, torch.int64), M, int(bool(repack)),
        ws.data_ptr(), ptr(out, f), stream_ptr())
    check(rc, "asvrl_policy_forward")
    return out


def policy_hidden(self_obs, objects, mask, w_self, b_self, w_obj, b_obj, w1, b1, w2, b2, image=False):
    """(rows, 128): tower -> 256 -> 128 -> 128 (ReLU) of the policy family in K9's persistent tcgen05 launch.
    ``image``: return the operand image (tiles, 2, 32768) uint8 instead (what c51_head reads)."""
    M = self_obs.shape[0]
    dev = self_obs.device
    ws = _POLICY_WS.get(dev)
    if ws is None:
        ws = _POLICY_WS[dev] = torch.empty((lib().asvrl_policy_forward_workspace(),), dtype=torch.uint8, device=dev)
    f = torch.float32
    if image:
        out = torch.empty(((M + 127) // 128, 2, 32768), dtype=torch.uint8, device=dev)
    else:
        out = torch.empty((M, 128), dtype=f, device=dev)
    rc = lib().asvrl_policy_hidden(
        ptr(self_obs, f), ptr(objects, f), ptr(mask, f), ptr(w_self, f), ptr(b_self, f), ptr(w_obj, f), ptr(b_obj, f),
        ptr(w1, f), ptr(b1, f), ptr(w2, f), ptr(b2, f), M, 1, ws.data_ptr(), None if image else ptr(out, f),
        out.data_ptr() if image else None, stream_ptr())
    check(rc, "asvrl_policy_hidden")
    return out


def policy_forward_status():
    """Synchronous read of K9's device status word; raises when a bounded wait of the pipeline expired."""
    v = C.c_int(0)
    check(lib().asvrl_policy_forward_status(C.byref(v)), "asvrl_policy_forward_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_policy_forward: device pipeline fault, status {v.value}")
    return 0


# ------------------------------------------------------------------ the two narrow layers of the update
def action_encoder_fwd(actions, wa, ba):
    B, D = actions.shape[0], wa.shape[0]
    f = torch.float32
    aenc = torch.empty((B, D), dtype=f, device=actions.device)
    check(lib().asvrl_action_encoder_fwd(ptr(actions, f), ptr(wa, f), ptr(ba, f), B, D, ptr(aenc), stream_ptr()),
          "asvrl_action_encoder_fwd")
    return aenc


def action_encoder_bwd(daenc, aenc, actions, wa, want_dact=True):
    B, D = aenc.shape
    f = torch.float32
    dev = aenc.device
    dwa = torch.empty((D, 2), dtype=f, device=dev)
    dba = torch.empty((D,), dtype=f, device=dev)
    dact = torch.empty((B, 2), dtype=f, device=dev) if want_dact else None
    part, cnt = _trunk_ws(dev, lib().asvrl_narrow_bwd_partials(D))
    check(lib().asvrl_action_encoder_bwd(ptr(daenc, f), ptr(aenc, f), ptr(actions, f), ptr(wa, f), B, D, ptr(dwa), ptr(dba),
                                         ptr(dact), ptr(part), ptr(cnt), stream_ptr()), "asvrl_action_encoder_bwd")
    return dwa, dba, dact


def atan_head_fwd(h, w3, b3, scale):
    B, D = h.shape
    f = torch.float32
    y = torch.empty((B, 2), dtype=f, device=h.device)
    act = torch.empty((B, 2), dtype=f, device=h.device)
    check(lib().asvrl_atan_head_fwd(ptr(h, f), ptr(w3, f), ptr(b3, f), float(scale), B, D, ptr(y), ptr(act), stream_ptr()),
          "asvrl_atan_head_fwd")
    return y, act


def atan_head_bwd(dact, y, h, w3, scale, want_dh=True):
    B, D = h.shape
    f = torch.float32
    dev = h.device
    dh = torch.empty((B, D), dtype=f, device=dev) if want_dh else None
    dw3 = torch.empty((2, D), dtype=f, device=dev)
    db3 = torch.empty((2,), dtype=f, device=dev)
    part, cnt = _trunk_ws(dev, lib().asvrl_narrow_bwd_partials(D))
    check(lib().asvrl_atan_head_bwd(ptr(dact, f), ptr(y, f), ptr(h, f), ptr(w3, f), float(scale), B, D, ptr(dh), ptr(dw3),
                                    ptr(db3), ptr(part), ptr(cnt), stream_ptr()), "asvrl_atan_head_bwd")
    return dh, dw3, db3


_C51_HEAD_WS = {}


def c51_head(h_img, zv, wa, ba, bv, support, rows, n_actions, want_q=False):
    """K13: greedy actions (rows,) int32 [and expected values (rows, A)] of Rainbow's dueling C51 head from the advantage
    stream's hidden-layer image (policy_hidden(..., image=True)) and the value stream's bias-free product zv (rows, >= 64)."""
    dev = zv.device
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _C51_HEAD_WS.get(key)
    if ws is None:
        ws = _C51_HEAD_WS[key] = torch.empty((lib().asvrl_c51_head_workspace(),), dtype=torch.uint8, device=dev)
    f = torch.float32
    atoms = support.numel()
    action = torch.empty((rows,), dtype=torch.int32, device=dev)
    q = torch.empty((rows, n_actions), dtype=f, device=dev) if want_q else None
    rc = lib().asvrl_c51_head(h_img.data_ptr(), ptr(zv, f), int(zv.shape[1]), ptr(wa, f), ptr(ba, f), ptr(bv, f),
                              ptr(support, f), int(rows), int(n_actions), int(atoms), ws.data_ptr(),
                              ptr(action, torch.int32), ptr(q, f), stream_ptr())
    check(rc, "asvrl_c51_head")
    return (action, q) if want_q else action


def c51_head_status():
    v = C.c_int(0)
    check(lib().asvrl_c51_head_status(C.byref(v)), "asvrl_c51_head_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_c51_head: device pipeline fault, status {v.value}")
    return 0


def neg_mean(x):
    """-mean(x) as a 0-d tensor: one launch, fixed summation order."""
    x = x.contiguous()
    out = torch.empty((), dtype=torch.float32, device=x.device)
    check(lib().asvrl_neg_mean(ptr(x, torch.float32), x.numel(), ptr(out), stream_ptr()), "asvrl_neg_mean")
    return out


def c51_greedy(zv, za, bv, ba, support, n_actions, want_q=False, actions_per_block=0):
    """Greedy actions (rows,) int32 [and expected values (rows, A)] of a dueling C51 head from its bias-free products.
    zv (rows, >= atoms); za (rows, A * atoms), or with ``actions_per_block`` (n_blocks, rows, ld) blocks."""
    rows = zv.shape[0]
    atoms = support.numel()
    f = torch.float32
    action = torch.empty((rows,), dtype=torch.int32, device=zv.device)
    q = torch.empty((rows, n_actions), dtype=f, device=zv.device) if want_q else None
    za_ld, za_bs = (za.shape[2], za.stride(0)) if actions_per_block else (0, 0)
    rc = lib().asvrl_c51_greedy(ptr(zv, f), ptr(za, f), ptr(bv, f), ptr(ba, f), ptr(support, f), rows, int(n_actions),
                                int(atoms), int(zv.shape[1]), int(actions_per_block), int(za_ld), int(za_bs),
                                ptr(action, torch.int32), ptr(q, f), stream_ptr())
    check(rc, "asvrl_c51_greedy")
    return (action, q) if want_q else action


# ------------------------------------------------------------------ K10: critic trunk forward on tcgen05
_CRITIC_WS = {}


def critic_workspace(key, device):
    """The 256 KB packed weight image of one network (``key``: any hashable, e.g. id of the module)."""
    ws = _CRITIC_WS.get((key, device))
    if ws is None:
        ws = _CRITIC_WS[(key, device)] = torch.empty((lib().asvrl_critic_forward_workspace(),), dtype=torch.uint8,
                                                     device=device)
    return ws


def critic_forward(feats, aenc, wc, bc, wh, bh, wh2, bh2, wo, bo, n_tau, workspace, tau_in=None, seed=0, counter=0,
                   counter_dev=None, cvar=1.0, repack=True, save=False, mean_hidden=False, pack_backward=False):
    """z (B, N), tau (B, N) [, images] of the IQN critic trunk in one persistent tcgen05 launch (K10).
    ``save``: also return the operand images (cos, h1, h2, r3) the backward kernels read.  ``mean_hidden`` (n_tau = 32,
    no aenc / wo / bo): z is (B, 128), the mean of the last hidden layer over the quantile rows (IQN's act).
    ``pack_backward``: a critic_backward_workspace() tensor — the weight-pack launch also writes K11's image into it
    (critic_backward(..., workspace=that tensor, repack=False) may follow)."""
    B = feats.shape[0]
    dev = feats.device
    f = torch.float32
    z = torch.empty((B, 128 if mean_hidden else n_tau), dtype=f, device=dev)
    tau = torch.empty((B, n_tau), dtype=f, device=dev) if tau_in is None else tau_in
    imgs = (None, None, None, None)
    if save:
        tiles = (B * n_tau + 127) // 128
        cb = lib().asvrl_critic_forward_chunk_bytes()
        imgs = tuple(torch.empty((tiles, n, cb), dtype=torch.uint8, device=dev) for n in (1, 4, 2, 2))
    rc = lib().asvrl_critic_forward(
        ptr(tau_in, f), int(seed) & 0xFFFFFFFFFFFFFFFF, int(counter), ptr(counter_dev, torch.int64), float(cvar),
        ptr(feats, f), ptr(aenc, f), ptr(wc, f), ptr(bc, f), ptr(wh, f), ptr(bh, f), ptr(wh2, f), ptr(bh2, f),
        ptr(wo, f), ptr(bo, f), B, int(n_tau), int(bool(mean_hidden)), int(bool(repack)), workspace.data_ptr(),
        None if pack_backward is None or pack_backward is False else pack_backward.data_ptr(),
        None if tau_in is not None else ptr(tau, f), ptr(z, f),
        *[None if t is None else t.data_ptr() for t in imgs], stream_ptr())
    check(rc, "asvrl_critic_forward")
    return (z, tau, imgs) if save else (z, tau)


_CRITIC_BWD_WS = {}


def critic_backward_workspace(dev, shared=True):
    """Buffer for K11's packed weight image: the (device, stream)'s shared one, or a fresh one (``shared=False``: one per
    differentiable forward, so interleaved forward / backward pairs cannot overwrite each other's image)."""
    n = lib().asvrl_critic_backward_workspace()
    if not shared:
        return torch.empty((n,), dtype=torch.uint8, device=dev)
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _CRITIC_BWD_WS.get(key)
    if ws is None:
        ws = _CRITIC_BWD_WS[key] = torch.empty((n,), dtype=torch.uint8, device=dev)
    return ws


def critic_backward(dz, r3_img, h2_img, h1_img, feats, aenc, wh, wh2, wo, n_tau, full=True, repack=True, fmax=None,
                    workspace=None):
    """K11.  full: -> (daenc, dfeats, (dp3, dp2, dp1, aux) images, scales); else -> daenc (the actor update's pass).
    ``fmax``: 1-element tensor, max of feats (sets the power-of-two scale of the dP1 image)."""
    B = feats.shape[0]
    dev = feats.device
    f = torch.float32
    ws = critic_backward_workspace(dev) if workspace is None else workspace
    daenc = torch.empty((B, 128), dtype=f, device=dev)
    dfeats = scales = None
    imgs = (None, None, None, None)
    if full:
        tiles = (B * n_tau + 127) // 128
        cb = lib().asvrl_critic_forward_chunk_bytes()
        dfeats = torch.empty((B, 256), dtype=f, device=dev)
        imgs = tuple(torch.empty((tiles, n, cb), dtype=torch.uint8, device=dev) for n in (2, 2, 4, 1))
        scales = torch.empty((3,), dtype=f, device=dev)   # 1 / S, 1 / S1 (for K12), sum dz
    rc = lib().asvrl_critic_backward(
        ptr(dz, f), r3_img.data_ptr(), h2_img.data_ptr(), h1_img.data_ptr() if full else None, ptr(feats, f), ptr(aenc, f),
        ptr(wh, f), ptr(wh2, f), ptr(wo, f), ptr(fmax, f), B, int(n_tau), int(bool(repack)), ws.data_ptr(), ptr(daenc, f),
        ptr(dfeats, f), *[None if t is None else t.data_ptr() for t in imgs], ptr(scales, f), stream_ptr())
    check(rc, "asvrl_critic_backward")
    return (daenc, dfeats, imgs, scales) if full else daenc


def critic_backward_status():
    v = C.c_int(0)
    check(lib().asvrl_critic_backward_status(C.byref(v)), "asvrl_critic_backward_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_critic_backward: device pipeline fault, status {v.value}")
    return 0


_WGRAD_WS = {}


def critic_wgrad(dp3, dp2, dp1, aux, cos_img, h1_img, h2_img, r3_img, scales, B, n_tau):
    """K12: (d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo) from the operand images of K10 / K11."""
    dev = dp3.device
    tiles = (B * n_tau + 127) // 128
    need = lib().asvrl_critic_wgrad_workspace(tiles)
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _WGRAD_WS.get(key)
    if ws is None or ws.numel() < need:
        ws = _WGRAD_WS[key] = torch.empty((need,), dtype=torch.float32, device=dev)
    new = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
    d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo = new(256, 64), new(256), new(128, 256), new(128), new(128, 128), new(128), new(128)
    rc = lib().asvrl_critic_wgrad(dp3.data_ptr(), dp2.data_ptr(), dp1.data_ptr(), aux.data_ptr(), cos_img.data_ptr(),
                                  h1_img.data_ptr(), h2_img.data_ptr(), r3_img.data_ptr(), ptr(scales, torch.float32),
                                  B, int(n_tau), ptr(ws), ptr(d_wc), ptr(d_bc), ptr(d_wh), ptr(d_bh), ptr(d_wh2),
                                  ptr(d_bh2), ptr(d_wo), stream_ptr())
    check(rc, "asvrl_critic_wgrad")
    return d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo


def critic_wgrad_status():
    v = C.c_int(0)
    check(lib().asvrl_critic_wgrad_status(C.byref(v)), "asvrl_critic_wgrad_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_critic_wgrad: device pipeline fault, status {v.value}")
    return 0


def critic_forward_status():
    """Synchronous read of K10's device status word; raises when a bounded wait of the pipeline expired."""
    v = C.c_int(0)
    check(lib().asvrl_critic_forward_status(C.byref(v)), "asvrl_critic_forward_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_critic_forward: device pipeline fault, status {v.value}")
    return 0


def linear_tf32x3_status():
    """Synchronous read of K8's device status word; raises when a bounded wait on the tensor pipe expired."""
    v = C.c_int(0)
    check(lib().asvrl_linear_tf32x3_status(C.byref(v)), "asvrl_linear_tf32x3_status")
    if v.value != 0:
        raise RuntimeError(f"asvrl_linear_tf32x3: device pipeline fault, status {v.value}")
    return 0


# ------------------------------------------------------------------ K6: observation tower
def tower_fwd(self_obs, objects, mask, w_self, b_self, w_obj, b_obj):
    B, SD = self_obs.shape
    K, OD = objects.shape[1], objects.shape[2]
    SF, OF = w_self.shape[0], w_obj.shape[0]
    feats = torch.empty((B, SF + K * OF), dtype=torch.float32, device=self_obs.device)
    f = torch.float32
    rc = lib().asvrl_tower_fwd(ptr(self_obs, f), ptr(objects, f), ptr(mask, f), ptr(w_self, f), ptr(b_self, f),
                               ptr(w_obj, f), ptr(b_obj, f), B, SD, OD, K, SF, OF, ptr(feats), stream_ptr())
    check(rc, "asvrl_tower_fwd")
    return feats


def tower_bwd(dfeats, feats, self_obs, objects, SF, OF):
    B, SD = self_obs.shape
    K, OD = objects.shape[1], objects.shape[2]
    dev = self_obs.device
    new = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
    dws, dbs, dwo, dbo = new(SF, SD), new(SF), new(OF, OD), new(OF)
    part, cnt = _trunk_ws(dev, lib().asvrl_tower_bwd_partials(B, K, SF, OF))
    f = torch.float32
    rc = lib().asvrl_tower_bwd(ptr(dfeats, f), ptr(feats, f), ptr(self_obs, f), ptr(objects, f), B, SD, OD, K,
                               SF, OF, ptr(dws), ptr(dbs), ptr(dwo), ptr(dbo), ptr(part), ptr(cnt), stream_ptr())
    check(rc, "asvrl_tower_bwd")
    return dws, dbs, dwo, dbo


# ------------------------------------------------------------------ K7: clip + Adam
def clip_adam_step(params, grads, exp_avgs, exp_avg_sqs, steps, max_norm, lr, beta1, beta2, eps, grad_scale,
                   norm_out, scratch, counter, peers=None, grad_sum=None, epoch=None):
    """One optimiser step over the tensor lists (<= 64 tensors per call is the ABI limit)."""
    if len(params) > _abi.OPT_MAX_TENSORS:
        raise RuntimeError(f"clip_adam_step: more than {_abi.OPT_MAX_TENSORS} tensors")
    t = _abi.AdamTensors()
    t.n = len(params)
    f = torch.float32
    for i, (p, g, m, v, st) in enumerate(zip(params, grads, exp_avgs, exp_avg_sqs, steps)):
        if g.numel() != p.numel():
            raise RuntimeError("clip_adam_step: gradient / parameter size mismatch")
        t.param[i], t.grad[i], t.exp_avg[i], t.exp_avg_sq[i] = ptr(p, f), ptr(g, f), ptr(m, f), ptr(v, f)
        t.step[i] = ptr(st, f)
        t.numel[i] = p.numel()
    if peers is not None:  # all-reduce over NVLink peer memory inside the kernels (grad_scale = 1 / world)
        rc = lib().asvrl_clip_adam_step_p2p(C.byref(t), C.byref(peers), ptr(grad_sum, f), ptr(epoch, torch.int32),
                                            max_norm, lr, beta1, beta2, eps, ptr(norm_out, f), ptr(scratch, f),
                                            ptr(counter, torch.int32), stream_ptr())
        check(rc, "asvrl_clip_adam_step_p2p")
        return
    rc = lib().asvrl_clip_adam_step(C.byref(t), max_norm, lr, beta1, beta2, eps, grad_scale,
                                    ptr(norm_out, f), ptr(scratch, f), ptr(counter, torch.int32), stream_ptr())
    check(rc, "asvrl_clip_adam_step")


# ------------------------------------------------------------------ K2: device replay ring
def new_replay(capacity, n_asv_max, device):
    """Ring of float32[70] rows + device counters (head_size[0] rows ever written,
    head_size[1] rows valid) + the push workspace (include/asvrl.h)."""
    return {
        "rows": torch.zeros((int(capacity), _abi.ROW), dtype=torch.float32, device=device),
        "head_size": torch.zeros((2,), dtype=torch.int64, device=device),
        "workspace": torch.zeros(((int(n_asv_max) + 63) // 64 + 4,), dtype=torch.int32, device=device),
        "capacity": int(capacity),
    }


def replay_push(rb, obs_self, obs_objects, obs_mask, actions, discrete, reward, next_self,
                next_objects, next_mask, status, term=None, ended=None):
    """``term`` / ``ended``: after a fused auto-reset step the next state of ASVs whose env ended
    is read from term["obs_*"] (out's rows already hold the fresh episode's observation)."""
    n_asv = status.numel()
    if (term is None) != (ended is None):
        raise RuntimeError("term and ended go together")
    rpe = status.shape[-1] if status.dim() >= 2 else 1
    if (n_asv + 63) // 64 + 4 > rb["workspace"].numel():
        raise RuntimeError("replay workspace too small for this many ASVs")
    a_dtype = torch.int32 if discrete else torch.float32
    rc = lib().asvrl_replay_push(ptr(rb["rows"], torch.float32), rb["capacity"],
                                 ptr(rb["head_size"], torch.int64), ptr(obs_self, torch.float32),
                                 ptr(obs_objects, torch.float32), ptr(obs_mask, torch.float32),
                                 ptr(actions, a_dtype), 1 if discrete else 0,
                                 ptr(reward, torch.float32), ptr(next_self, torch.float32),
                                 ptr(next_objects, torch.float32), ptr(next_mask, torch.float32),
                                 ptr(status, torch.int32), n_asv, ptr(rb["workspace"], torch.int32),
                                 ptr(term["obs_self"], torch.float32) if term else None,
                                 ptr(term["obs_objects"], torch.float32) if term else None,
                                 ptr(term["obs_mask"], torch.float32) if term else None,
                                 ptr(ended, torch.uint8), rpe, stream_ptr())
    check(rc, "asvrl_replay_push")


def new_batch(B, device):
    f = dict(dtype=torch.float32, device=device)
    return {"self": torch.empty((B, 7), **f), "obj": torch.empty((B, 5, 5), **f),
            "mask": torch.empty((B, 5), **f), "act": torch.empty((B, 2), **f),
            "rew": torch.empty((B,), **f), "nself": torch.empty((B, 7), **f),
            "nobj": torch.empty((B, 5, 5), **f), "nmask": torch.empty((B, 5), **f),
            "done": torch.empty((B,), **f),
            "indices": torch.empty((B,), dtype=torch.int64, device=device)}


def replay_sample(rb, batch, seed=0, counter=0, indices=None, counter_dev=None):
    """Fills the dense batch tensors of ``batch`` (new_batch) from the ring."""
    B = batch["rew"].numel()
    rc = lib().asvrl_replay_sample(ptr(rb["rows"], torch.float32), ptr(rb["head_size"], torch.int64),
                                   rb["capacity"], ptr(indices, torch.int64), seed, counter, B,
                                   ptr(batch["self"]), ptr(batch["obj"]), ptr(batch["mask"]),
                                   ptr(batch["act"]), ptr(batch["rew"]), ptr(batch["nself"]),
                                   ptr(batch["nobj"]), ptr(batch["nmask"]), ptr(batch["done"]),
                                   ptr(batch["indices"], torch.int64), ptr(counter_dev, torch.int64),
                                   stream_ptr())
    check(rc, "asvrl_replay_sample")
    return batch
