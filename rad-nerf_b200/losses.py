"""NeRFLoss with the reference's terms and keyword interface (losses.py:40-76):
rgb MSE, opacity entropy, CV^2 of the gate importance, depth-mutual.  The distortion term needs
results['ws_i'...] keys that ml_render never sets in the reference either."""
from __future__ import annotations

import torch
from torch import nn


class NeRFLoss(nn.Module):
    def __init__(self, lambda_opacity=1e-3):
        super().__init__()

    def forward(self, results, target, lambda_opacity=1e-3, lambda_distortion=0, lambda_disp=0,
                lambda_cv_importance=0, lambda_depth_mutual=0):
        loss = {}
        loss["rgb"] = (results["rgb"] - target["rgb"]) ** 2
        o = results["opacity"] + 1e-10
        loss["opacity"] = lambda_opacity * (-o * torch.log(o))      # pushes opacity to 0 or 1
        if lambda_disp > 0:
            loss["disp"] = lambda_disp * results["disp"] ** 2
        if lambda_distortion > 0:
            raise NotImplementedError("distortion loss is outside the Rad-NeRF hot path (SURVEY.md 8f, f3)")
        K = results["gating_code"].shape[-1] if "gating_code" in results else 1
        if lambda_cv_importance > 0 and K > 1:
            imp = results["gating_importance"].float()
            loss["cv_importance"] = lambda_cv_importance * (imp.var() / (imp.mean() ** 2 + 1e-10))
        if lambda_depth_mutual > 0 and K > 1:
            fused = torch.sum(results["depth"] * results["gating_code"], 1, keepdim=True).detach()
            loss["depth_mutual"] = lambda_depth_mutual * ((results["depth"] - fused) ** 2)
        return loss
