"""Generate the golden fixtures in this directory from the REFERENCE's own code.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):   python tests/golden/make_golden.py

What is imported from the reference (read-only mount), unmodified:
  * submodules/gsplat/gsplat/cuda/_torch_impl.py -- gsplat's pure-PyTorch twin
    that its own tests compare the CUDA kernels against (tests/test_basic.py):
    _quat_scale_to_covar_preci :9, _fully_fused_projection :130, _isect_tiles :201,
    _isect_offset_encode :273, _spherical_harmonics :574
  * utils/loss_utils.py -- l1_loss :18, ssim :46 (the unfused loss the fused
    kernels of clm_kernels/ssim.cu replace)
The twin's _rasterize_to_pixels needs a CUDA op + nerfacc and cannot run here
(SURVEY.md 8c); rasterisation is pinned in tests/test_oracle_pin.py against a
torch-autograd restatement of `accumulate` (_torch_impl.py:360-371) instead.

Outputs: proj.npz, isect.npz, sh.npz, loss.npz  (small, committed).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "clm-gs_b200"))
sys.path.insert(0, "/root/reference/submodules/gsplat")
sys.path.insert(0, "/root/reference")

from gsplat.cuda import _torch_impl as T  # noqa: E402  (the reference)
from utils.loss_utils import l1_loss, ssim  # noqa: E402  (the reference)
from clmgs import scene as S  # noqa: E402


def small_scene(n=3000, W=96, H=64, ncam=3):
    spec = S.SceneSpec("golden", n, W, H, ncam, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, ncam, g["extent"])
    scales = np.exp(g["log_scale"]).astype(np.float32)
    quats = g["raw_quat"] / np.linalg.norm(g["raw_quat"], axis=1, keepdims=True)
    return spec, g, V, K, P, scales, quats.astype(np.float32)


def main(out_dir=HERE):
    torch.manual_seed(0)
    spec, g, V, K, P, scales, quats = small_scene()
    W, H = spec.width, spec.height
    means = torch.from_numpy(g["xyz"]).requires_grad_()
    q = torch.from_numpy(quats).requires_grad_()
    s = torch.from_numpy(scales).requires_grad_()
    viewmats, Ks = torch.from_numpy(V), torch.from_numpy(K)

    # ---- projection fwd + bwd --------------------------------------------
    covars, _ = T._quat_scale_to_covar_preci(q, s, True, False, triu=False)
    radii, means2d, depths, conics, _ = T._fully_fused_projection(means, covars, viewmats, Ks, W, H)
    valid = radii > 0
    v_means2d = torch.randn_like(means2d) * valid[..., None]
    v_depths = torch.randn_like(depths) * valid
    v_conics = torch.randn_like(conics) * valid[..., None]
    loss = (means2d * v_means2d).sum() + (depths * v_depths).sum() + (conics * v_conics).sum()
    v_means, v_quats, v_scales = torch.autograd.grad(loss, [means, q, s])
    np.savez_compressed(
        os.path.join(out_dir, "proj.npz"), means=g["xyz"], quats=quats, scales=scales, viewmats=V, Ks=K,
        width=W, height=H, radii=radii.numpy(), means2d=means2d.detach().numpy(),
        depths=depths.detach().numpy(), conics=conics.detach().numpy(), v_means2d=v_means2d.numpy(),
        v_depths=v_depths.numpy(), v_conics=v_conics.numpy(), v_means=v_means.numpy(),
        v_quats=v_quats.numpy(), v_scales=v_scales.numpy())

    # ---- tile intersection (exact) -----------------------------------------
    tw, th = (W + 15) // 16, (H + 15) // 16
    m2d, dep = means2d.detach(), depths.detach()
    tpg, isect_ids, flatten_ids = T._isect_tiles(m2d, radii, dep, 16, tw, th, sort=True)
    offsets = T._isect_offset_encode(isect_ids, len(V), tw, th)
    np.savez_compressed(
        os.path.join(out_dir, "isect.npz"), means2d=m2d.numpy(), radii=radii.numpy(), depths=dep.numpy(),
        tile_size=16, tile_width=tw, tile_height=th, tiles_per_gauss=tpg.numpy(),
        isect_ids=isect_ids.numpy(), flatten_ids=flatten_ids.numpy(), isect_offsets=offsets.numpy())

    # ---- spherical harmonics fwd + grads, degrees 0..4 ------------------------
    n_sh = 500
    out = {}
    dirs = torch.randn(n_sh, 3)
    for deg in range(5):
        Kb = 25 if deg == 4 else 16
        coeffs = torch.randn(n_sh, Kb, 3, requires_grad=True)
        d = dirs.clone().requires_grad_()
        col = T._spherical_harmonics(deg, d, coeffs)
        v_col = torch.randn_like(col)
        gc, gd = torch.autograd.grad((col * v_col).sum(), [coeffs, d], allow_unused=True)
        out[f"coeffs{deg}"] = coeffs.detach().numpy()
        out[f"colors{deg}"] = col.detach().numpy()
        out[f"v_colors{deg}"] = v_col.numpy()
        out[f"v_coeffs{deg}"] = gc.numpy()
        out[f"v_dirs{deg}"] = (gd if gd is not None else torch.zeros_like(d)).numpy()
    np.savez_compressed(os.path.join(out_dir, "sh.npz"), dirs=dirs.numpy(), **out)

    # ---- loss: 0.8*L1 + 0.2*(1-ssim) and its gradient -----------------------
    Hh, Ww = 45, 70  # not multiples of 32/16 on purpose
    img = torch.rand(3, Hh, Ww, requires_grad=True)
    gt_u8 = (torch.rand(3, Hh, Ww) * 255).to(torch.uint8)
    gt = torch.clamp(gt_u8 / 255.0, 0.0, 1.0)
    Ll1 = l1_loss(img, gt)
    ss = ssim(img.unsqueeze(0), gt.unsqueeze(0))
    L = 0.8 * Ll1 + 0.2 * (1.0 - ss)
    (gi,) = torch.autograd.grad(L, [img])
    np.savez_compressed(os.path.join(out_dir, "loss.npz"), image=img.detach().numpy(), gt_u8=gt_u8.numpy(),
                        loss=L.item(), ssim=ss.item(), l1=Ll1.item(), v_image=gi.numpy())
    for f in ("proj", "isect", "sh", "loss"):
        print(f, os.path.getsize(os.path.join(out_dir, f + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
