"""Generates tests/golden/vren_*.npz from the REFERENCE's own CUDA kernels.

Run on a GPU box (`gpurun -- python tests/golden/make_golden_vren.py`): it imports
oracle/_ref/vren_ref*.so -- the unmodified /root/reference/models/csrc sources compiled for
sm_100a by oracle/build_ref.sh -- on small seeded inputs and stores inputs + outputs.
tests/test_golden.py then pins the CPU oracle (and the product kernels) to these vectors.
Output goes to gpurun_out/golden/ (copied into tests/golden/ and committed).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import vren_ref  # noqa: E402
from conftest import CONFIGS, get_scene  # noqa: E402
from radnerf_b200 import synthetic as syn  # noqa: E402


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def main():
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    n_rays = 192
    for name, (scale, esf) in CONFIGS.items():
        sc, bf = get_scene(scale)
        cam = syn.make_cameras(100, 1.5 if scale <= 0.5 else 1.0)
        o, d, _, _ = syn.train_batch(cam, n_rays, 123)
        d[:3, 1] = 0.0
        o[3:8] *= 20.0
        center = np.zeros((1, 3), np.float32); half = np.full((1, 3), scale, np.float32)
        g = {"scale": np.float32(scale), "esf": np.float32(esf), "cascades": np.int32(sc.cascades),
             "rays_o": o, "rays_d": d, "scene_seed": np.int32(1234)}
        cnt, ht, hi = vren_ref.ray_aabb_intersect(cu(o), cu(d), cu(center), cu(half), 1)
        g["aabb_cnt"], g["aabb_t"], g["aabb_idx"] = cnt.cpu().numpy(), ht.cpu().numpy(), hi.cpu().numpy()
        h = ht[:, 0].clone()
        h[(h[:, 0] >= 0) & (h[:, 0] < syn.NEAR_DISTANCE), 0] = syn.NEAR_DISTANCE
        g["hits_t"] = h.cpu().numpy()
        # packbits digest of the full grid (the grid itself is regenerated from the seed)
        pb = torch.zeros(len(bf), dtype=torch.uint8, device="cuda")
        vren_ref.packbits(cu(sc.density_grid), float(sc.threshold), pb)
        pbn = pb.cpu().numpy()
        g["packbits_popcount"] = np.int64(np.unpackbits(pbn).sum())
        g["packbits_head"] = pbn[:4096].copy()
        g["packbits_xor32"] = np.bitwise_xor.reduce(pbn.view(np.uint32))
        g["threshold"] = np.float32(sc.threshold)
        # morton
        coords = np.random.RandomState(1).randint(0, 128, (1000, 3)).astype(np.int32)
        g["morton_coords"] = coords
        g["morton_idx"] = vren_ref.morton3D(cu(coords)).cpu().numpy()
        g["morton_inv"] = vren_ref.morton3D_invert(cu(g["morton_idx"])).cpu().numpy()
        # train marching
        noise = np.random.RandomState(7).rand(n_rays).astype(np.float32)
        g["noise"] = noise
        ra, x, dd, dl, ts, c = vren_ref.raymarching_train(cu(o), cu(d), h.contiguous(), cu(bf), sc.cascades, scale, esf,
                                                          cu(noise), 128, 1024)
        tot = int(c[0])
        ra = ra.cpu().numpy(); order = np.argsort(ra[:, 0]); ra = ra[order]
        gather = np.concatenate([np.arange(s, s + n) for s, n in zip(ra[:, 1], ra[:, 2])]).astype(np.int64)
        g["train_counts"] = ra[:, 2].astype(np.int32)
        g["train_total"] = np.int32(tot)
        g["train_xyzs"] = x.cpu().numpy()[gather]; g["train_dirs_first"] = dd.cpu().numpy()[gather][:64]
        g["train_deltas"] = dl.cpu().numpy()[gather]; g["train_ts"] = ts.cpu().numpy()[gather]
        # composite on the canonical (ascending-ray) layout
        rays_a = np.stack([ra[:, 0], np.concatenate([[0], np.cumsum(ra[:, 2])[:-1]]), ra[:, 2]], 1).astype(np.int64)
        rng = np.random.RandomState(0)
        sig = np.minimum(np.exp(rng.normal(1.5, 2.0, tot)), np.exp(15)).astype(np.float32)
        rgbs = rng.uniform(0, 1, (tot, 3)).astype(np.float32)
        g["sigmas"], g["rgbs"] = sig, rgbs
        T, op, dp, col, ws = vren_ref.composite_train_fw(cu(sig), cu(rgbs), cu(g["train_deltas"]), cu(g["train_ts"]),
                                                         cu(rays_a), 1e-4)
        g["fw_total"], g["fw_opacity"], g["fw_depth"], g["fw_rgb"], g["fw_ws"] = (t.cpu().numpy() for t in (T, op, dp, col, ws))
        gO = rng.normal(size=n_rays).astype(np.float32); gD = rng.normal(size=n_rays).astype(np.float32)
        gC = rng.normal(size=(n_rays, 3)).astype(np.float32); gW = rng.normal(size=tot).astype(np.float32)
        g["gO"], g["gD"], g["gC"], g["gW"] = gO, gD, gC, gW
        gs, gc = vren_ref.composite_train_bw(cu(gO), cu(gD), cu(gC), cu(gW), cu(sig), cu(rgbs), ws, cu(g["train_deltas"]),
                                             cu(g["train_ts"]), cu(rays_a), op, dp, col, 1e-4)
        g["bw_dsigmas"], g["bw_drgbs"] = gs.cpu().numpy(), gc.cpu().numpy()
        # test-time marching: 3 rounds with N_samples = 4, then composite_test_fw
        ht2 = h.clone().contiguous()
        alive = torch.arange(n_rays, device="cuda")
        opac = torch.zeros(n_rays, device="cuda"); dep = torch.zeros(n_rays, device="cuda"); col2 = torch.zeros(n_rays, 3, device="cuda")
        for rnd in range(3):
            x, dd, dl, ts, ne = vren_ref.raymarching_test(cu(o), cu(d), ht2, alive, cu(bf), sc.cascades, scale, esf, 128, 1024, 4)
            g[f"test{rnd}_alive_in"] = alive.cpu().numpy()
            g[f"test{rnd}_xyzs"], g[f"test{rnd}_deltas"], g[f"test{rnd}_ts"], g[f"test{rnd}_neff"] = (
                t.cpu().numpy() for t in (x, dl, ts, ne))
            g[f"test{rnd}_hits_t"] = ht2.cpu().numpy()
            s2 = torch.from_numpy(np.exp(np.random.RandomState(rnd).normal(3.0, 2.0, tuple(dl.shape))).astype(np.float32)).cuda()
            c2 = torch.from_numpy(np.random.RandomState(10 + rnd).uniform(0, 1, tuple(x.shape)).astype(np.float32)).cuda()
            g[f"test{rnd}_sigmas"], g[f"test{rnd}_rgbs"] = s2.cpu().numpy(), c2.cpu().numpy()
            vren_ref.composite_test_fw(s2, c2, dl, ts, ht2, alive, 1e-4, ne, opac, dep, col2)
            g[f"test{rnd}_alive_out"] = alive.cpu().numpy()
            g[f"test{rnd}_opacity"], g[f"test{rnd}_depth"], g[f"test{rnd}_rgb"] = opac.cpu().numpy(), dep.cpu().numpy(), col2.cpu().numpy()
            alive = alive[alive >= 0]
        path = os.path.join(out_dir, f"vren_{name}.npz")
        np.savez_compressed(path, **g)
        print("wrote", path, os.path.getsize(path), "bytes; total samples", tot)


if __name__ == "__main__":
    main()
