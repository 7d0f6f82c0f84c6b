"""ctypes binding of libacx.so (the C ABI declared in include/acx.h).

There is no CPU fallback: if the CUDA library is missing or a call fails, this raises.
PyTorch is used only for device memory, streams and torch.distributed.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# ACX_LIB_PATH: development override (A/B builds of the same ABI, e.g. `make prof`); the product loads the in-tree library
LIB_PATH = os.environ.get("ACX_LIB_PATH") or os.path.join(_HERE, "libacx.so")
CSRC = os.path.join(_HERE, "csrc")

ENV_PENDULUM, ENV_CARTPOLE, ENV_ACROBOT = 0, 1, 2
SCAN_AUTO, SCAN_COLS, SCAN_WARP = 0, 1, 2
OPT_SIZE, STATS_SIZE = 16, 16
SCAN_STATIC_MAX_COLUMNS = 131072
(OPT_ACTOR_LR, OPT_CRITIC_LR, OPT_BETA_KL, OPT_GAMMA_ENT, OPT_BETA1_POW, OPT_BETA2_POW, OPT_CLIP, OPT_DESIRED_KL,
 OPT_MIN_LR, OPT_MAX_LR, OPT_ITER, OPT_SCHEDULE, OPT_ADAM_T) = range(13)
(STAT_ADV_MEAN, STAT_ADV_INV_STD, STAT_ADV_STD, STAT_ACTOR_LOSS, STAT_CRITIC_LOSS, STAT_KL_PRE, STAT_ENT_MEAN,
 STAT_KL_POST, STAT_REW_LOSS, STAT_ACTOR_LR_USED, STAT_EV_BEFORE, STAT_EV_AFTER, STAT_EP_REW_MEAN,
 STAT_EP_COUNT) = range(14)


class AcxError(RuntimeError):
    pass


def build(verbose=False):
    """Compile libacx.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", CSRC, "-j4"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode != 0:
        raise AcxError("building libacx.so failed (see output above)")
    return LIB_PATH


_p, _i, _f, _d, _u64, _u32, _sz = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_uint64, C.c_uint32, C.c_size_t
_ip = C.POINTER(C.c_int)

# name -> (restype, argtypes); must list every symbol include/acx.h declares (tests check this)
SIGNATURES = {
    "acx_abi_version": (_i, []),
    "acx_last_error": (C.c_char_p, []),
    "acx_params_layout": (_i, [_i, _i, _i, _i, _ip, _ip]),
    "acx_env_dims": (_i, [_i, _ip, _ip, _ip, _ip]),
    "acx_env_reset": (_i, [_i, _i, _p, _p, _p, _p, _u64, _u64, _u32, _p]),
    "acx_env_step": (_i, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _u64, _u64, _u32, _p, _i, _p]),
    "acx_env_step_pendulum": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _u64, _u64, _u32, _p, _i, _p]),
    "acx_env_step_cartpole": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _u64, _u64, _u32, _p, _i, _p]),
    "acx_env_step_acrobot": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _u64, _u64, _u32, _p, _i, _p]),
    "acx_policy_forward_sample_gauss": (_i, [_p, _i, _i, _f, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p]),
    "acx_policy_forward_sample_cat": (_i, [_p, _i, _i, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p]),
    "acx_kl_workspace_bytes": (_sz, []),
    "acx_policy_logp_gauss": (_i, [_p, _i, _i, _f, _i, _i, _p, _p, _p, _p, _p, _p, _p, _ip, _p]),
    "acx_policy_logp_cat": (_i, [_p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _ip, _p]),
    "acx_value_forward": (_i, [_p, _i, _i, _i, _p, _p, _i, _p]),
    "acx_policy_forward_sample_gauss_cached": (_i, [_p, _i, _i, _f, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p, _p, _p]),
    "acx_policy_forward_sample_cat_cached": (_i, [_p, _i, _i, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p, _p, _p]),
    "acx_value_forward_cached": (_i, [_p, _i, _i, _i, _p, _p, _i, _p, _p, _p]),
    "acx_loss_backward_gauss_cached": (_i, [_p, _p, _i, _i, _f, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_loss_backward_cat_cached": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_frame_push": (_i, [_i, _i, _i, _p, _p, _p, _p, _p]),
    "acx_activation_cache_floats": (_sz, [_i, _i, _i]),
    "acx_act_mask_words": (_sz, [_i, _i, _i]),
    "acx_tc_packed_bwd_floats": (_sz, [_i, _i, _i, _i]),
    "acx_loss_backward_tc_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "acx_tc_pack_weights_bwd": (_i, [_i, _p, _i, _i, _i, _p, _p]),
    "acx_loss_backward_gauss_tc": (_i, [_p, _p, _p, _p, _i, _i, _f, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_loss_backward_cat_tc": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_tc_packed_floats": (_sz, [_i, _i, _i, _i]),
    "acx_tc_pack_weights": (_i, [_i, _p, _i, _i, _i, _p, _p]),
    "acx_policy_forward_sample_gauss_tc": (_i, [_p, _p, _i, _i, _f, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p, _p, _p]),
    "acx_policy_forward_sample_cat_tc": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _u64, _u64, _u32, _p, _p, _p, _p, _p, _p, _p, _p]),
    "acx_rollout_tc": (_i, [_i, _p, _p, _f, _i, _i, _p, _p, _p, _p, _p, _u64, _u64, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "acx_policy_logp_gauss_tc": (_i, [_p, _p, _i, _i, _f, _i, _i, _p, _p, _p, _p, _p, _p, _p, _ip, _p]),
    "acx_policy_logp_cat_tc": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _ip, _p]),
    "acx_value_forward_tc": (_i, [_p, _p, _i, _i, _i, _p, _p, _i, _p, _p, _p, _p]),
    "acx_returns_adv_workspace_bytes": (_sz, [_i, _i]),
    "acx_returns_adv_scan": (_i, [_i, _i, _p, _p, _p, _d, _i, _p, _p, _p, _p, _sz, _i, _p]),
    "acx_returns_adv_scan_static": (_i, [_i, _i, _p, _p, _p, _d, _i, _p, _p, _p, _i, _p]),
    "acx_diag_workspace_bytes": (_sz, []),
    "acx_episode_returns": (_i, [_i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_explained_variance": (_i, [C.c_longlong, _p, _p, _p, _p, _sz, _p]),
    "acx_log_push": (_i, [_p, _i, _p, _p, _p]),
    "acx_env_stagger": (_i, [_i, _i, _p, _u64, _u64, _i, _p]),
    "acx_loss_backward_workspace_bytes": (_sz, [_i, _i, _i]),
    "acx_loss_backward_gauss": (_i, [_p, _p, _i, _i, _f, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_loss_backward_cat": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_loss_backward_basic_critic": (_i, [_p, _i, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "acx_opt_init": (_i, [_p, _f, _f, _f, _f, _f, _f, _f, _f, _i, _p]),
    "acx_adam_apply": (_i, [_p, _p, _p, _p, _i, _p, _p, _p, _p, _i, _p, _f, _p]),
    "acx_adam_tick": (_i, [_p, _p]),
    "acx_counter_add": (_i, [_p, _u32, _p]),
    "acx_kl_finalize": (_i, [_p, _i, _d, _p, _p]),
    "acx_kl_lr_update": (_i, [_p, _p, _p]),
    "acx_kl_lr_update_split": (_i, [_p, _p, _p, _f, _i, _p]),
    "acx_xgpu_flag_offset_bytes": (_sz, [_i]),
    "acx_xgpu_buffer_bytes": (_sz, [_i]),
    "acx_allreduce_adam": (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _i, _p]),
}

_lib = None


def load():
    """dlopen libacx.so and bind every entry point; raises AcxError if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AcxError("%s not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.acx_abi_version() != 1:
        raise AcxError("libacx ABI version mismatch")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().acx_last_error().decode(errors="replace")
        raise AcxError("%s failed (%d): %s" % (what or "libacx call", rc, msg))


def call(name, *args):
    """Call an int-returning entry point and raise AcxError on a non-zero code."""
    check(getattr(load(), name)(*args), name)


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def cur_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def params_layout(net, ob_dim, num_ac=1, cont=True):
    """-> (dict name -> (offset, size), total padded floats)."""
    offs, sizes = (C.c_int * 9)(), (C.c_int * 9)()
    total = load().acx_params_layout(net, ob_dim, num_ac, int(cont), offs, sizes)
    if total < 0:
        check(total, "acx_params_layout")
    names = ["W1", "b1", "W2", "b2", "W3", "b3", "Wh", "bh", "log_std"]
    return {n: (offs[i], sizes[i]) for i, n in enumerate(names) if offs[i] >= 0}, total
