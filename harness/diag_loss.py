"""Diagnostic: time of the fused L1 + SSIM loss (forward + backward) against the reference's torch formulas at
1920x1080 on the GPU."""
import sys
import torch
from . import add_paths
add_paths()


def main():
    from utils.loss_utils import fused_l1_ssim_loss
    from oracle import loss_reference as lr
    dev = "cuda"
    g = torch.Generator().manual_seed(0)
    gt = torch.rand(3, 1080, 1920, generator=g).to(dev)
    img = (gt + 0.1 * torch.randn(3, 1080, 1920, generator=g).to(dev)).clamp(0, 1)

    def run(fn, n=20):
        x = img.clone().requires_grad_(True)
        for _ in range(3):
            x.grad = None
            fn(x, gt).backward()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            x.grad = None
            l = fn(x, gt)
            l.backward()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, float(l), x.grad

    t_o, l_o, g_o = run(lambda a, b: fused_l1_ssim_loss(a, b, 0.2)[0])
    t_r, l_r, g_r = run(lambda a, b: lr.l1_ssim_loss(a, b, 0.2)[0])
    print(f"fused: {t_o:.3f} ms  loss {l_o:.7f}   reference torch ops: {t_r:.3f} ms  loss {l_r:.7f}   "
          f"grad relerr {float((g_o - g_r).abs().max() / g_r.abs().max()):.2e}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
