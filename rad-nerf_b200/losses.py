"""NeRFLoss with the reference's terms and keyword interface (losses.py:40-76): rgb MSE, opacity entropy, CV^2 of
the gate importance, depth-mutual, and DistortionLoss (losses.py:6-37) over results['ws_i' / 'deltas_i' / 'ts_i' /
'rays_a_i'] -- keys that ml_render itself never sets in the reference, so the term is reachable
only for callers that add them."""
from __future__ import annotations

import torch
from torch import nn

from . import vren


class DistortionLoss(torch.autograd.Function):
    """Mip-NeRF 360 distortion loss, DVGO-v2 formulation (reference losses.py:6-37).
    ws, deltas, ts (N); rays_a (N_rays,3) [ray_idx, start_idx, N_samples] -> loss (N_rays)."""

    @staticmethod
    def forward(ctx, ws, deltas, ts, rays_a):
        loss, ws_inclusive_scan, wts_inclusive_scan = vren.distortion_loss_fw(ws.contiguous(), deltas, ts, rays_a)
        ctx.save_for_backward(ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts, rays_a)
        return loss

    @staticmethod
    def backward(ctx, dL_dloss):
        ws_inclusive_scan, wts_inclusive_scan, ws, deltas, ts, rays_a = ctx.saved_tensors
        dL_dws = vren.distortion_loss_bw(dL_dloss.contiguous(), ws_inclusive_scan, wts_inclusive_scan, ws.contiguous(), deltas, ts, rays_a)
        return dL_dws, None, None, None


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
            loss["distortion"] = 0
            for i in range(results["gating_code"].shape[-1]):
                loss["distortion"] = loss["distortion"] + lambda_distortion * DistortionLoss.apply(
                    results[f"ws_{i}"], results[f"deltas_{i}"], results[f"ts_{i}"], results[f"rays_a_{i}"]).mean()
        K = results["gating_code"].shape[-1] if "gating_code" in results else 1
        if lambda_cv_importance > 0 and K > 1:
            imp = results["gating_importance"].float()
            loss["cv_importance"] = lambda_cv_importance * (imp.var() / (imp.mean() ** 2 + 1e-10))
        if lambda_depth_mutual > 0 and K > 1:
            fused = torch.sum(results["depth"] * results["gating_code"], 1, keepdim=True).detach()
            loss["depth_mutual"] = lambda_depth_mutual * ((results["depth"] - fused) ** 2)
        return loss
