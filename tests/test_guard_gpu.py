"""Step guard of the fused engine (include/radnerf_b200.h: RNB_GRAD_TAIL, rnb_engine_adam_guarded, rnb_engine_grad_check):
the GradScaler equivalent of the reference's Lightning precision=16 run (/root/reference/train_ml.py:304) and the
handling of a step whose samples do not fit the workspace (the reference sizes for N_rays * 1024 and never truncates:
raymarching.cu:298-305).

  * a step whose gradient seeds saturate fp16 / whose gradients hold an inf or nan is SKIPPED on the device: parameters,
    Adam moments and the applied-step count stay, the gradients are cleared, the host halves the loss scale;
  * after `growth_interval` applied steps the scale doubles;
  * a step whose samples exceed sample_capacity is skipped the same way and the host raises SampleCapacityError
    (on_overflow="raise") or re-allocates the workspace and carries on (on_overflow="grow");
  * the applied-step count used for Adam's bias corrections lives on the device: a skipped step does not advance it.
"""
import numpy as np
import pytest
import torch

from radnerf_b200 import synthetic as syn
from test_e2e_gpu import build

pytestmark = pytest.mark.gpu


def _batch(n, seed=42):
    cam = syn.make_cameras(100, 1.5)
    o, d, _, _ = syn.train_batch(cam, n, seed)
    tgt = np.random.RandomState(seed).rand(n, 3).astype(np.float32)
    return torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), torch.from_numpy(tgt).cuda()


def _engine(n=1024, K=2, **kw):
    from radnerf_b200.engine import Engine
    P, model, gate, bitfields = build(0.5, K, 0.2)
    return Engine(model, gate, n_rays=n, exp_step_factor=0.0, **kw)


def test_saturated_seed_skips_the_step_and_halves_the_scale():
    eng = _engine(loss_scale=2.0 ** 40)            # every gradient seed saturates fp16 at this scale
    o, d, t = _batch(1024)
    before = eng.flat.clone(); m0 = eng.exp_avg.clone()
    eng.step(o, d, t)
    st = eng.guard_status()
    assert st["skipped_nonfinite"] == 1 and st["applied"] == 0
    assert torch.equal(eng.flat, before) and torch.equal(eng.exp_avg, m0)          # parameters and moments untouched
    assert float(eng.grads.abs().max()) == 0.0                                     # gradients cleared for the next step
    eng._poll_guard(wait=True)
    assert eng.cfg.loss_scale == 2.0 ** 39                                         # backoff 0.5
    # keeps halving until the seeds fit, then trains
    for _ in range(40):
        eng.step(o, d, t)
        eng._poll_guard(wait=True)
    st = eng.guard_status()
    assert st["applied"] >= 1 and st["skipped_nonfinite"] < 40
    assert not torch.equal(eng.flat, before)
    assert bool(torch.isfinite(eng.flat).all()) and bool(torch.isfinite(eng.exp_avg_sq).all())


@pytest.mark.parametrize("where", ["table", "geo", "gate"])
def test_injected_nonfinite_gradient_is_skipped(where):
    """An inf planted in the gradient buffer (as an overflowing fp16 chain would leave it) must not reach Adam's moments."""
    eng = _engine()
    o, d, t = _batch(1024)
    eng.step(o, d, t)                                # one clean step
    assert eng.guard_status()["applied"] == 1
    snap = (eng.flat.clone(), eng.exp_avg.clone(), eng.exp_avg_sq.clone())
    eng.forward_backward(o, d, t, None, eng.background(), True)
    off = {"table": eng.offsets["table"] + 12345, "geo": eng.offsets["geo0"] + 7, "gate": eng.offsets["gate"] + 3}[where]
    eng.grads[off] = float("inf") if where != "gate" else float("nan")
    if where == "table":
        # the table gradient is checked where it is produced (d(encoding) in the field backward); a planted value stands for
        # that detection having fired
        eng._grad_buf[eng.n_params:].view(torch.int32)[0] |= 1
    eng.adam_step()
    st = eng.guard_status()
    assert st["skipped_nonfinite"] == 1 and st["applied"] == 1
    for a, b in zip(snap, (eng.flat, eng.exp_avg, eng.exp_avg_sq)):
        assert torch.equal(a, b)
    assert float(eng.grads.abs().max()) == 0.0
    eng.step(o, d, t)                                # the next step applies again, with the bias corrections of step 2
    st = eng.guard_status()
    assert st["applied"] == 2 and bool(torch.isfinite(eng.flat).all())


def test_overflowing_fp16_chain_is_skipped():
    """Weights arranged so that the FORWARD is unchanged but the backward chain overflows fp16 between two layers: the
    geometry feature is scaled down by f where the geo network emits it and scaled back up by f where the rgb network reads
    it, so d(feature) = dZ W_rgb1[:, 16:32] f leaves fp16's range.  The inf lands in the geo weight gradients and in
    d(encoding); the field backward's own check and rnb_engine_grad_check raise the flag (no planted values), the step is
    skipped.  The same loss scale with the original weights is applied cleanly first: it is not the seeds that saturate."""
    o, d, t = _batch(1024)
    eng = _engine(loss_scale=2.0 ** 24, dynamic_loss_scale=False)
    eng.step(o, d, t)
    assert eng.guard_status()["applied"] == 1
    loss_ref = float(eng.loss[0])
    skipped = None
    for f in (2.0 ** 12, 2.0 ** 14, 2.0 ** 16):
        eng = _engine(loss_scale=2.0 ** 24, dynamic_loss_scale=False)
        with torch.no_grad():
            for i in range(2):
                geo = getattr(eng.model, f"geo_net_{i}").params          # [W1 (64,32) | W2 (32,64)]: rows 1..16 of W2 emit the feature
                geo[64 * 32 + 64:64 * 32 + 17 * 64].mul_(1.0 / f)
                rgb = getattr(eng.model, f"rgb_net_{i}").params.view(-1)[:64 * 32].view(64, 32)   # W1 (64,32): columns 16..31 read it
                rgb[:, 16:].mul_(f)
        before = eng.flat.clone()
        eng.step(o, d, t)
        st = eng.guard_status()
        assert abs(float(eng.loss[0]) - loss_ref) < 2e-2 * abs(loss_ref), (f, float(eng.loss[0]), loss_ref)   # same forward (fp16 rounding aside)
        if st["skipped_nonfinite"] == 1:
            skipped = f
            assert st["applied"] == 0 and torch.equal(eng.flat, before) and float(eng.exp_avg.abs().max()) == 0.0
            break
    assert skipped is not None, "no factor made the backward chain overflow"


def test_loss_scale_grows_after_clean_steps():
    eng = _engine(growth_interval=4)
    o, d, t = _batch(1024)
    s0 = eng.cfg.loss_scale
    for _ in range(6):
        eng.step(o, d, t)
        eng._poll_guard(wait=True)
    assert eng.cfg.loss_scale == 2 * s0 and eng.guard_status()["applied"] == 6


def test_sample_capacity_overflow_raises():
    from radnerf_b200.engine import SampleCapacityError
    eng = _engine(samples_per_ray=4)                 # the scene needs ~30 samples per ray
    o, d, t = _batch(1024)
    before = eng.flat.clone()
    eng.step(o, d, t)                                # skipped on the device ...
    st = eng.guard_status()
    assert st["skipped_capacity"] == 1 and st["applied"] == 0 and st["max_samples_needed"] > eng.cfg.sample_capacity
    assert torch.equal(eng.flat, before)
    with pytest.raises(SampleCapacityError):         # ... and reported at the next step boundary
        eng._poll_guard(wait=True)


def test_sample_capacity_overflow_grows_and_matches_a_large_workspace():
    eng = _engine(samples_per_ray=4, on_overflow="grow")
    ref = _engine(samples_per_ray=256)
    o, d, t = _batch(1024)
    torch.manual_seed(3)
    noise = torch.rand(2, 1024, device="cuda")
    eng.step(o, d, t, noise=noise)                   # skipped, then the workspace grows
    eng._poll_guard(wait=True)
    assert eng.cfg.sample_capacity >= eng.guard_status()["max_samples_needed"]
    eng.step(o, d, t, noise=noise)
    ref.step(o, d, t, noise=noise)
    a, b = eng.guard_status(), ref.guard_status()
    assert a["applied"] == 1 and a["skipped_capacity"] == 1 and b["applied"] == 1
    assert abs(float(eng.loss[0]) - float(ref.loss[0])) < 1e-6
    assert float((eng.flat - ref.flat).abs().max()) < 1e-6


def test_trainer_grows_by_default():
    from radnerf_b200.trainer import Trainer
    tr = Trainer(scale=0.5, K=2, freeze_occupancy=True, device="cuda", engine=True, batch_rays=1024, samples_per_ray=4)
    from conftest import get_scene
    tr.set_occupancy([get_scene(0.5, jitter=i if i else None)[1] for i in range(2)])
    o, d, t = _batch(1024)
    for _ in range(3):
        tr.step(o, d, t)
        tr.engine._poll_guard(wait=True)
    st = tr.engine.guard_status()
    assert st["skipped_capacity"] >= 1 and st["applied"] >= 1
