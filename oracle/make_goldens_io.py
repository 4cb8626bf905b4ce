"""Pin oracle/io_oracle.py against the reference's own datasets/ray_utils.py and datasets/depth_utils.py (loaded by file
path; kornia.create_meshgrid stubbed, see io_oracle's header) and emit tests/golden/g11_io.npz.  Build container only."""
import importlib.util
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
REF = os.environ.get("NERF_REFERENCE", "/root/reference")
from oracle import io_oracle as IO  # noqa: E402


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    kornia = types.ModuleType("kornia")

    def create_meshgrid(H, W, normalized_coordinates=False):
        assert not normalized_coordinates
        ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
        return torch.stack([xs, ys], -1)[None]
    kornia.create_meshgrid = create_meshgrid
    sys.modules["kornia"] = kornia
    ru = load(os.path.join(REF, "datasets", "ray_utils.py"), "ref_ray_utils")
    du = load(os.path.join(REF, "datasets", "depth_utils.py"), "ref_depth_utils")
    g = torch.Generator().manual_seed(0)
    out = {}
    for tag, (H, W, focal, ndc) in {"blender": (100, 120, 166.7, False), "llff": (63, 84, 67.9, True)}.items():
        a = torch.randn(3, 3, generator=g)
        q, _ = torch.linalg.qr(a)
        c2w = torch.cat([q, torch.tensor([[0.3], [-0.2], [4.0 if not ndc else 0.1]])], 1)
        if ndc:
            c2w[:, 2] = torch.tensor([0.05, -0.03, 1.0]); c2w[:, :3] = torch.linalg.qr(c2w[:, :3])[0]
        d = ru.get_ray_directions(H, W, focal)
        assert torch.equal(d, IO.get_ray_directions(H, W, focal))
        o_ref, d_ref = ru.get_rays(d, c2w)
        near, far = 2.0, 6.0
        if ndc:
            near, far = 0, 1
            o_ref, d_ref = ru.get_ndc_rays(H, W, focal, 1.0, o_ref, d_ref)
        rays_ref = torch.cat([o_ref, d_ref, near * torch.ones_like(o_ref[:, :1]), far * torch.ones_like(o_ref[:, :1])], 1)
        assert torch.equal(rays_ref, IO.image_rays(c2w, H, W, focal, 2.0, 6.0, ndc)), tag
        out[f"{tag}:c2w"], out[f"{tag}:rays"] = c2w.numpy(), rays_ref.numpy()
        out[f"{tag}:hwf"] = np.array([H, W, focal])
    depth = torch.rand(37, 53, generator=g).numpy() * 4 + 2
    depth[3, 5] = np.nan; depth[7, 9] = np.inf
    dn = np.nan_to_num(depth)
    with tempfile.NamedTemporaryFile(suffix=".pfm") as f:
        du.save_pfm(f.name, dn)
        ref_bytes = open(f.name, "rb").read()
    assert ref_bytes == IO.pfm_bytes(dn)
    out["pfm:depth"], out["pfm:bytes"] = depth, np.frombuffer(ref_bytes, dtype=np.uint8)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "g11_io.npz"), **out)
    print("ok", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
