"""Generates tests/golden/loss_ssim_variants.npz with the REFERENCE's utils/loss_utils.py (imported from /root/reference,
CPU): ssim at another window size and per-batch (size_average=False), and the evaluation-only ssim_masked.
Run once:  python tests/golden/make_golden_ssim_variants.py"""
import importlib.util
import os

import numpy as np
import torch

REF = os.environ.get("H3DGS_REFERENCE_ROOT", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(REF, "utils", "loss_utils.py"))
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

g = torch.Generator().manual_seed(21)
a = torch.rand(2, 3, 36, 44, generator=g)
b = (a + 0.2 * torch.randn(2, 3, 36, 44, generator=g)).clamp(0, 1)
mask = (torch.rand(2, 1, 36, 44, generator=g) > 0.4).float()
mask[:, :, :14, :18] = 0            # a region without any support: invalid windows
out = dict(a=a.numpy(), b=b.numpy(), mask=mask.numpy(),
           ssim_w7=ref.ssim(a, b, 7).item(), ssim_w11_batch=ref.ssim(a, b, 11, False).numpy(),
           masked_mean=ref.ssim_masked(a, b, mask).item(),
           masked_map=ref.ssim_masked(a, b, mask, size_average=False).numpy(),
           masked_mean_3d=ref.ssim_masked(a, b, mask[:, 0]).item())
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_ssim_variants.npz"), **out)
print({k: (v.shape if hasattr(v, "shape") and v.shape else float(v)) for k, v in out.items()})
