"""Checkpoint compatibility with the reference (utils/util.py:8-43).

The drop-in modules keep tinycudann's flat fp32 `params` layout and the reference's module / buffer names, so a
state-dict written by the reference's trainer (PyTorch-Lightning `.ckpt` with keys `model.*`, `ray_gate.*`, or the
`slim_ckpt` form) loads directly, and what `save_ckpt` writes can be read back by the reference's `load_ckpt`."""
from __future__ import annotations

import torch


def extract_model_state_dict(ckpt_path, model_name="model", prefixes_to_ignore=()):
    """utils/util.py:8-23."""
    checkpoint = torch.load(ckpt_path, map_location="cpu", weights_only=False)
    if "state_dict" in checkpoint:          # a pytorch-lightning checkpoint
        checkpoint = checkpoint["state_dict"]
    out = {}
    for k, v in checkpoint.items():
        if not k.startswith(model_name):
            continue
        k = k[len(model_name) + 1:]
        if any(k.startswith(p) for p in prefixes_to_ignore):
            continue
        out[k] = v
    return out


def load_ckpt(model, ckpt_path, model_name="model", prefixes_to_ignore=()):
    """utils/util.py:26-31: keys present in the file overwrite the module's state, the rest keep their values."""
    if not ckpt_path:
        return
    state = model.state_dict()
    state.update(extract_model_state_dict(ckpt_path, model_name, prefixes_to_ignore))
    model.load_state_dict(state)


def slim_ckpt(ckpt_path, save_poses=False):
    """utils/util.py:34-43: drop what inference does not need."""
    ckpt = torch.load(ckpt_path, map_location="cpu", weights_only=False)
    sd = ckpt["state_dict"] if "state_dict" in ckpt else ckpt
    drop = ["directions", "model.density_grid", "model.grid_coords"] + ([] if save_poses else ["poses"])
    drop += [k for k in sd if k.startswith("val_lpips")]
    for k in drop:
        sd.pop(k, None)
    return sd


def save_ckpt(path, model, gate=None, extra=None, engine=None):
    """Writes {'state_dict': {'model.*': ..., 'ray_gate.*': ...}} -- the key layout of the reference's Lightning module
    (train_ml.py: self.model, self.ray_gate).

    Multi-GPU (sharded optimizer): the fp32 masters are current only inside each rank's shard, so they are gathered first --
    a COLLECTIVE.  Call save_ckpt on EVERY rank (all of them enter the gather); only rank 0 writes the file.  Calling it on
    rank 0 alone would deadlock in the gather, so that is refused when it can be detected (path=None on the other ranks is
    the explicit way to say "I only take part in the gather")."""
    import torch.distributed as dist
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    if engine is not None:
        engine.gather_master()
    if multi and (path is None or dist.get_rank() != 0):
        return
    sd = {f"model.{k}": v.detach().cpu() for k, v in model.state_dict().items()}
    if gate is not None:
        sd.update({f"ray_gate.{k}": v.detach().cpu() for k, v in gate.state_dict().items()})
    out = {"state_dict": sd}
    if extra:
        out.update(extra)
    torch.save(out, path)
