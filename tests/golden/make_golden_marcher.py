"""Golden vectors for data/ray_utils.py:152-197 ray_marcher, produced by EXECUTING THE REFERENCE (build container only).

Usage:  python tests/golden/make_golden_marcher.py      (rewrites tests/golden/ray_marcher.npz)
Inputs are regenerated from the seed by tests (marcher_inputs below is imported by them)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {"lin": dict(S=16, lindisp=False, perturb=0.0), "lin_jit": dict(S=21, lindisp=False, perturb=1.0),
         "disp_jit": dict(S=64, lindisp=True, perturb=0.5)}


def marcher_inputs(name, R=37):
    c = CASES[name]
    rng = np.random.default_rng(sum(map(ord, name)))
    rays = np.empty((R, 8), np.float32)
    rays[:, 0:3] = rng.normal(size=(R, 3)).astype(np.float32)
    rays[:, 3:6] = rng.normal(size=(R, 3)).astype(np.float32)
    rays[:, 6] = (0.5 + rng.random(R)).astype(np.float32)
    rays[:, 7] = (2.5 + 3 * rng.random(R)).astype(np.float32)
    rnd = rng.random((R, c["S"]), dtype=np.float32)
    return rays, rnd, c


if __name__ == "__main__":
    import torch
    sys.path.insert(0, HERE)
    from make_golden import import_reference
    import_reference()
    import importlib.util, types
    REF = os.environ.get("UCNERF_REFERENCE", "/root/reference")
    pkg = types.ModuleType("data")     # data/__init__.py imports the dataset classes (imageio, ...): load the one file
    pkg.__path__ = [os.path.join(REF, "data")]
    sys.modules["data"] = pkg
    spec = importlib.util.spec_from_file_location("data.ray_utils", os.path.join(REF, "data", "ray_utils.py"))
    RU = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(RU)
    out = {}
    real_rand = torch.rand
    for name in CASES:
        rays, rnd, c = marcher_inputs(name)
        torch.rand = lambda *a, **k: torch.from_numpy(rnd)             # ray_utils.py:190 draws torch.rand(z_vals.shape)
        xyz, o, d, z = RU.ray_marcher(torch.from_numpy(rays), N_samples=c["S"], lindisp=c["lindisp"], perturb=c["perturb"])
        torch.rand = real_rand
        out[name + "_xyz"], out[name + "_z"] = xyz.numpy(), z.numpy()
    np.savez_compressed(os.path.join(HERE, "ray_marcher.npz"), **out)
    print({k: v.shape for k, v in out.items()})
