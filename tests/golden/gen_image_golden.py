"""Generate tests/golden/image_ref.npz by running the reference's OWN image.img_gradient
(mast3r_slam/image.py:5-38, pure torch, CPU) on a seeded input.  Build container only
(needs /root/reference).

    python tests/golden/gen_image_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_image", os.path.join(REF, "mast3r_slam", "image.py"))
ref_image = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_image)

g = torch.Generator().manual_seed(4242)
out = {}
for name, shape in (("a", (2, 3, 13, 17)), ("b", (1, 5, 2, 2)), ("c", (3, 1, 40, 33))):
    img = torch.randn(shape, generator=g)
    gx, gy = ref_image.img_gradient(img)
    out[f"{name}_img"], out[f"{name}_gx"], out[f"{name}_gy"] = img.numpy(), gx.numpy(), gy.numpy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "image_ref.npz"), **out)
print("wrote image_ref.npz", {k: v.shape for k, v in out.items()})
