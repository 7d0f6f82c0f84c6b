"""The kernels write nothing outside the buffers the caller hands them (include/acx.h: "no allocation inside; the caller
owns every buffer").  compute-sanitizer is not available on the GPU pool, so this is the bounds check: every device
buffer of a trainer is re-homed into the middle of a larger allocation whose 4 KB margins hold a byte pattern, two
iterations run on ragged sizes (env counts that are not a multiple of the 128-sample tile), and the margins must still
hold the pattern.  An out-of-bounds write of up to 4 KB on either side of any buffer fails the test and names the buffer."""
import pytest
import torch

pytestmark = pytest.mark.gpu
GUARD, PATTERN = 4096, 0xA5

# every tensor attribute of A2CTrainer that a kernel may write (None-valued ones are skipped)
BUFFERS = ["params", "m", "v", "_gradbuf", "state", "elapsed", "obs_buf", "raw_obs", "acs", "logps", "ent", "rew", "done",
           "V", "G", "adv", "stats", "opt", "step_ctr", "ep_ret", "ws_diag", "V_after", "log_ring", "ws_scan", "ws_bwd",
           "ws_bwd2", "kl_part", "ax2", "ax3", "cx2", "cx3", "amask", "cmask", "packed_a", "packed_c", "packed_bwd_a",
           "packed_bwd_c", "old_params"]


def rehome(t):
    """Copy t into the middle of a pattern-filled allocation; returns (view with t's dtype / shape / contents, raw)."""
    assert t.is_contiguous()
    nbytes = t.numel() * t.element_size()
    body = (nbytes + 255) // 256 * 256                      # keep the 256-byte alignment torch's allocator gives
    raw = torch.full((GUARD + body + GUARD,), PATTERN, device=t.device, dtype=torch.uint8)
    mid = raw[GUARD:GUARD + nbytes]
    mid.copy_(t.reshape(-1).view(torch.uint8))
    return mid.view(t.dtype).reshape(t.shape), raw, nbytes


def check(name, raw, nbytes):
    lo = raw[:GUARD]
    hi = raw[GUARD + nbytes:]                               # the alignment slack after the body counts as margin too
    bad_lo = int((lo != PATTERN).sum())
    bad_hi = int((hi != PATTERN).sum())
    assert bad_lo == 0 and bad_hi == 0, "%s: %d bytes written before it, %d after it" % (name, bad_lo, bad_hi)


CASES = [
    dict(env_id="Pendulum-v0", n_envs=300, horizon=5),
    dict(env_id="Pendulum-v0", n_envs=257, horizon=4),
    dict(env_id="CartPole-v0", n_envs=391, horizon=6, diagnostics=True),
    dict(env_id="Acrobot-v1", n_envs=200, horizon=5),
    dict(env_id="Pendulum-v0", n_envs=261, horizon=5, tensor_cores=False),
    dict(env_id="CartPole-v0", n_envs=130, horizon=5, frames=3),
    dict(env_id="CartPole-v0", n_envs=250, horizon=8, basic=True, clip=1.0, max_lr=1.0),
    dict(env_id="Pendulum-v0", n_envs=300, horizon=5, fused_rollout=False),
]


@pytest.mark.parametrize("kw", CASES, ids=lambda kw: "-".join("%s" % v for v in kw.values()))
def test_no_write_outside_the_callers_buffers(kw):
    from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer
    tr = A2CTrainer(seed=11, use_graph=False, **kw)
    torch.cuda.synchronize()
    raws = {}
    for name in BUFFERS:
        t = getattr(tr, name, None)
        if t is None:
            continue
        new, raw, nbytes = rehome(t)
        setattr(tr, name, new)
        raws[name] = (raw, nbytes)
    # the views the trainer derived from the re-homed tensors
    P = tr.Pa + tr.Pc
    tr.actor, tr.critic = tr.params[:tr.Pa], tr.params[tr.Pa:]
    tr.grads = tr._gradbuf[:P]
    for _ in range(2):
        tr.iteration()
    torch.cuda.synchronize()
    st = tr.read_stats()
    assert st["iteration"] == 2 and st["crit_loss"] == st["crit_loss"]
    for name, (raw, nbytes) in raws.items():
        check(name, raw, nbytes)
