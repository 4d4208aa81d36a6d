"""Golden vectors of the latent head from the UNMODIFIED reference (run in the build container only):

    python -m oracle.make_golden_latent

Runs HierVAE.rsample (ggpm/hpgnn.py:34-41) in fp64 on seeded inputs, captures the noise it drew (same seed, same
``randn_like`` call), asserts the oracle restatement agrees to 1e-12 on outputs and gradients, and writes
tests/golden/latent_rsample.npz."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import latent_oracle as lo                                                 # noqa: E402
from oracle.ref_loader import load_reference                                            # noqa: E402


def main():
    load_reference()
    hpgnn = importlib.import_module("ggpm.hpgnn")
    out = {}
    torch.set_default_dtype(torch.float64)
    for case, (B, H, L, seed) in {"a": (20, 250, 24, 1), "b": (7, 48, 5, 2), "c": (33, 64, 12, 3)}.items():
        g = torch.Generator().manual_seed(seed)
        z = torch.randn(B, H, generator=g)
        wm, wv = torch.randn(L, H, generator=g) / H ** 0.5, torch.randn(L, H, generator=g) / H ** 0.5
        bm, bv = torch.randn(L, generator=g) * 0.1, torch.randn(L, generator=g) * 0.1
        cot = torch.randn(B, L, generator=g)
        for perturb in (True, False):
            Wm, Wv = torch.nn.Linear(H, L), torch.nn.Linear(H, L)
            with torch.no_grad():
                Wm.weight.copy_(wm); Wm.bias.copy_(bm); Wv.weight.copy_(wv); Wv.bias.copy_(bv)
            zin = z.clone().requires_grad_(True)
            torch.manual_seed(100 + seed)
            eps = torch.randn(B, L)                                   # what randn_like(z_mean) yields under this seed
            torch.manual_seed(100 + seed)
            zs, kl = hpgnn.HierVAE.rsample(None, zin, Wm, Wv, perturb)
            ((zs * cot).sum() + 0.7 * kl).backward()
            # oracle pin
            leaves = [t.clone().requires_grad_(True) for t in (z, wm, bm, wv, bv)]
            zo, klo = lo.rsample(*leaves, eps, perturb)
            ((zo * cot).sum() + 0.7 * klo).backward()
            ref_grads = [zin.grad, Wm.weight.grad, Wm.bias.grad, Wv.weight.grad, Wv.bias.grad]
            for a, b in [(zs, zo), (kl, klo)] + list(zip(ref_grads, [t.grad for t in leaves])):
                assert float((a - b).abs().max()) <= 1e-12 * max(1.0, float(b.abs().max())), case
            tag = "%s_%d_" % (case, int(perturb))
            out.update({tag + "z": zs.detach().numpy(), tag + "kl": kl.detach().numpy(), tag + "dz_in": zin.grad.numpy(),
                        tag + "dwm": Wm.weight.grad.numpy(), tag + "dbm": Wm.bias.grad.numpy(),
                        tag + "dwv": Wv.weight.grad.numpy(), tag + "dbv": Wv.bias.grad.numpy()})
        out.update({case + "_in_z": z.numpy(), case + "_wm": wm.numpy(), case + "_bm": bm.numpy(), case + "_wv": wv.numpy(),
                    case + "_bv": bv.numpy(), case + "_eps": eps.numpy(), case + "_cot": cot.numpy()})
    path = os.path.join(ROOT, "tests", "golden_latent", "latent_rsample.npz")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
