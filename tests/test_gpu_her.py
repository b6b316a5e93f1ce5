"""Hindsight experience replay on device storage against the reference's own buffers
(tests/golden/her.npz, make_golden.gen_her): seeded sample_indices re-write the same episodes
with the same goals and rewards, and _restore_cache puts the storage back.  Bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _np(x):
    return x.cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


@pytest.mark.parametrize("name", ["ring", "ring_wrapped", "ring_no_obs_next", "vector"])
def test_her_rewrite_and_restore_vs_reference(name):
    import tianshou_fork_b200 as ts
    from helpers import HER_SPECS, her_reward, her_stream
    g = load_golden("her")[name]
    _, nb, total, steps, ignore, horizon, k = next(s for s in HER_SPECS if s[0] == name)
    kw = dict(compute_reward_fn=her_reward, horizon=horizon, future_k=k, ignore_obs_next=ignore)
    buf = ts.HERVectorReplayBuffer(total, nb, **kw) if nb else ts.HERReplayBuffer(total, **kw)
    rng = np.random.default_rng(970)
    for step in her_stream(rng, steps, max(nb, 1)):
        if nb:
            buf.add(ts.Batch(**step), buffer_ids=np.arange(nb))
        else:
            buf.add(ts.Batch(**{k2: (v[0] if not isinstance(v, dict) else
                                     {a: b[0] for a, b in v.items()})
                                for k2, v in step.items()}))

    def check(tag):
        assert np.array_equal(_np(buf._meta.obs.desired_goal), g[f"{tag}/goal"]), tag
        assert np.array_equal(_np(buf._meta.rew), g[f"{tag}/rew"]), tag
        if not ignore:
            assert np.array_equal(_np(buf._meta.obs_next.desired_goal), g[f"{tag}/goal_next"]), tag

    check("orig")
    np.random.seed(5)
    for r, bsz in enumerate((6, 11)):
        idx = buf.sample_indices(bsz)
        assert np.array_equal(_np(idx), g[f"draw{r}/indices"]), r
        check(f"draw{r}")
    buf._restore_cache()
    check("restored")
    # writes restore first: an add after a sample leaves no re-written goal behind
    buf.sample_indices(6)
    step = next(her_stream(np.random.default_rng(1), 1, max(nb, 1)))
    if nb:
        buf.add(ts.Batch(**step), buffer_ids=np.arange(nb))
    else:
        buf.add(ts.Batch(**{k2: (v[0] if not isinstance(v, dict) else
                                 {a: b[0] for a, b in v.items()}) for k2, v in step.items()}))
    dg = buf._meta.obs.desired_goal  # every filled slot holds the environment's goal again
    assert bool(((dg == 5) | (dg == 0)).all())
