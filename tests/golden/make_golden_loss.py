"""Generates tests/golden/loss_l1_ssim.npz by running the REFERENCE's own utils/loss_utils.py (pure PyTorch, imported
from /root/reference in the build container, CPU) on seeded images.  Run once:  python tests/golden/make_golden_loss.py"""
import importlib.util
import os

import numpy as np
import torch

REF = os.environ.get("H3DGS_REFERENCE_ROOT", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(REF, "utils", "loss_utils.py"))
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)


def case(seed, c, h, w, lam):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(c, h, w, generator=g)
    # a "render": the ground truth blurred by noise, clamped like gaussian_renderer does
    img = (gt + 0.15 * torch.randn(c, h, w, generator=g)).clamp(0, 1).requires_grad_(True)
    ll1 = ref.l1_loss(img, gt)
    s = ref.ssim(img, gt)
    loss = (1.0 - lam) * ll1 + lam * (1.0 - s)
    loss.backward()
    return dict(image=img.detach().numpy(), gt=gt.numpy(), lam=np.float32(lam), l1=ll1.item(), ssim=s.item(),
                loss=loss.item(), grad=img.grad.numpy())


def masked_case(seed, c, h, w, lam):
    """train_single.py:116-125 with an alpha mask: `image *= alpha_mask` (in place on the render), then the
    reference's l1_loss / ssim; and the depth loss of :133-139, torch.abs((invDepth - mono) * depth_mask).mean()."""
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(c, h, w, generator=g)
    img = (gt + 0.15 * torch.randn(c, h, w, generator=g)).clamp(0, 1).requires_grad_(True)
    mask = (torch.rand(1, h, w, generator=g) > 0.3).float()
    mask[:, : h // 3, : w // 4] = 0.5            # soft values too
    image = img * mask                              # out-of-place equivalent of `image *= alpha_mask`
    ll1 = ref.l1_loss(image, gt)
    s = ref.ssim(image, gt)
    loss = (1.0 - lam) * ll1 + lam * (1.0 - s)
    loss.backward()
    inv = torch.rand(1, h, w, generator=g).requires_grad_(True)
    mono = torch.rand(1, h, w, generator=g)
    dmask = (torch.rand(1, h, w, generator=g) > 0.5).float()
    ld = torch.abs((inv - mono) * dmask).mean()
    ld.backward()
    return dict(image=img.detach().numpy(), gt=gt.numpy(), mask=mask.numpy(), lam=np.float32(lam), l1=ll1.item(),
                ssim=s.item(), loss=loss.item(), grad=img.grad.numpy(), inv=inv.detach().numpy(), mono=mono.numpy(),
                dmask=dmask.numpy(), depth_l1=ld.item(), depth_grad=inv.grad.numpy())


out = {}
for name, args in {"m1": (10, 3, 40, 52, 0.2), "m2": (11, 3, 17, 33, 0.2)}.items():
    for k, v in masked_case(*args).items():
        out[f"{name}_{k}"] = v
for name, args in {"a": (0, 3, 45, 70, 0.2), "b": (1, 3, 16, 16, 0.2), "c": (2, 1, 7, 5, 0.5), "d": (3, 3, 33, 64, 0.0)}.items():
    for k, v in case(*args).items():
        out[f"{name}_{k}"] = v
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_l1_ssim.npz"), **out)
print({k: (v.shape if hasattr(v, "shape") and v.shape else float(v)) for k, v in out.items() if not k.endswith(("image", "gt", "grad", "mask", "inv", "mono"))})
