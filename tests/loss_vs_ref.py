"""Debug helper: SSIM gradient of ours vs the reference's fusedssim CUDA kernels vs torch autograd."""
import os, sys, importlib.util
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch, torch.nn.functional as F
from clmgs import ops
spec = importlib.util.spec_from_file_location("clm_kernels_ref_C", os.path.join(ROOT, "oracle/_ref/clm_kernels/clm_kernels_ref_C.so"))
CK = importlib.util.module_from_spec(spec); spec.loader.exec_module(CK)
torch.manual_seed(0)
for H, W in ((64, 64), (822, 1237)):
    a = torch.rand(1, 3, H, W, device="cuda"); b = torch.rand(1, 3, H, W, device="cuda")
    mp, d1, d2, d3 = CK.fusedssim(1e-4, 9e-4, a, b, True)
    dl = torch.full_like(a, 1.0 / a.numel())
    g_ref = CK.fusedssim_backward(1e-4, 9e-4, a, b, dl, d1, d2, d3)
    s, g_our = ops.fused_ssim_fwd_bwd(a, b)
    w1 = torch.tensor([0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331, 0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331, 0.036000773310661316, 0.0075987582094967365, 0.001028380123898387], dtype=torch.float64, device="cuda")
    win = (w1[:, None] * w1[None, :]).expand(3, 1, 11, 11)
    ta = a.double().requires_grad_(); tb = b.double()
    mu1 = F.conv2d(ta, win, padding=5, groups=3); mu2 = F.conv2d(tb, win, padding=5, groups=3)
    s11 = F.conv2d(ta * ta, win, padding=5, groups=3) - mu1 ** 2; s22 = F.conv2d(tb * tb, win, padding=5, groups=3) - mu2 ** 2
    s12 = F.conv2d(ta * tb, win, padding=5, groups=3) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 ** 2 + mu2 ** 2 + 1e-4) * (s11 + s22 + 9e-4))
    (g_true,) = torch.autograd.grad(m.mean(), [ta])
    rel = lambda x, y: float((x - y).abs().max() / y.abs().max())
    print(H, W, "ssim ref %.7f ours %.7f true %.7f" % (mp.mean().item(), s.item(), m.mean().item()))
    print("   grad: ours vs true %.2e | ref-cuda vs true %.2e | ours vs ref-cuda %.2e" % (rel(g_our.double(), g_true), rel(g_ref.double(), g_true), rel(g_our, g_ref)))
    d = (g_ref.double() - g_true).abs()[0, 0]
    iy, ix = np.unravel_index(int(d.argmax()), d.shape)
    print("   ref-cuda worst pixel at (y=%d, x=%d)" % (iy, ix))
