"""Golden vectors of the property heads from the UNMODIFIED reference (run in the build container only):

    python -m oracle.make_golden_property

Runs ggpm.property_optimizer.PropertyOptimizer (ggpm/property_optimizer.py:5-51) in fp64 on seeded inputs — eval-style
(dropout 0) and train mode with dropout 0.25, whose multipliers are captured by forward hooks on the reference's own
nn.Dropout modules — asserts the oracle restatement agrees to 1e-12 on outputs, losses and every gradient, and writes
tests/golden_property/property_heads.npz."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import property_oracle as po                                               # noqa: E402
from oracle.ref_loader import load_reference_module                                     # noqa: E402

CASES = {"a": (256, 12, 128, 0.0, 1), "b": (37, 12, 128, 0.25, 2), "c": (20, 8, [48], 0.25, 3), "d": (9, 6, [16, 10], 0.0, 4)}


def main():
    mod = load_reference_module("property_optimizer")
    torch.set_default_dtype(torch.float64)
    out = {}
    for case, (B, In, hidden, p, seed) in CASES.items():
        torch.manual_seed(seed)
        model = mod.PropertyOptimizer(In, hidden, p).train()
        for q in model.parameters():
            torch.nn.init.normal_(q, std=0.4)
        g = torch.Generator().manual_seed(100 + seed)
        xh, xl = torch.randn(B, In, generator=g), torch.randn(B, In, generator=g)
        th, tl = torch.randn(B, generator=g), torch.randn(B, generator=g)
        coth, cotl = torch.randn(B, generator=g), torch.randn(B, generator=g)
        drop, hooks = {}, []
        for prefix, reg in (("homo_linear.", model.homo_linear), ("lumo_linear.", model.lumo_linear)):
            for i, layer in enumerate(reg.linear):
                if isinstance(layer, torch.nn.Dropout):
                    def hook(m, inp, o, key=prefix + str(i)):
                        x = inp[0]
                        drop[key] = torch.where(x != 0, o / torch.where(x != 0, x, torch.ones_like(x)), torch.zeros_like(x)).detach()
                    hooks.append(layer.register_forward_hook(hook))
        a, b = xh.clone().requires_grad_(True), xl.clone().requires_grad_(True)
        lh, ll, yh, yl = model(a, b, (th, tl))
        (1.3 * lh + 0.7 * ll + (yh * coth).sum() + (yl * cotl).sum()).backward()
        for h in hooks:
            h.remove()
        sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        ao, bo = xh.clone().requires_grad_(True), xl.clone().requires_grad_(True)
        olh, oll, oyh, oyl = po.forward(leaves, ao, bo, (th, tl), drop)
        (1.3 * olh + 0.7 * oll + (oyh * coth).sum() + (oyl * cotl).sum()).backward()
        pairs = [(lh, olh), (ll, oll), (yh, oyh), (yl, oyl), (a.grad, ao.grad), (b.grad, bo.grad)]
        pairs += [(q.grad, leaves[k].grad) for k, q in model.named_parameters()]
        for u, v in pairs:
            assert float((u - v).abs().max()) <= 1e-12 * max(1.0, float(v.abs().max())), case
        out.update({case + "_xh": xh.numpy(), case + "_xl": xl.numpy(), case + "_th": th.numpy(), case + "_tl": tl.numpy(),
                    case + "_coth": coth.numpy(), case + "_cotl": cotl.numpy(), case + "_lh": lh.detach().numpy(),
                    case + "_ll": ll.detach().numpy(), case + "_yh": yh.detach().numpy(), case + "_yl": yl.detach().numpy(),
                    case + "_dxh": a.grad.numpy(), case + "_dxl": b.grad.numpy()})
        out.update({case + "_sd_" + k: v.numpy() for k, v in sd.items()})
        out.update({case + "_g_" + k: q.grad.numpy() for k, q in model.named_parameters()})
        out.update({case + "_drop_" + k: v.numpy() for k, v in drop.items()})
    path = os.path.join(ROOT, "tests", "golden_property", "property_heads.npz")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
