"""TEST INFRASTRUCTURE ONLY — CPU restatement of the property heads.

Follows PropertyRegressor.forward (ggpm/property_optimizer.py:63-66: Linear, ReLU, Dropout repeated, then Linear(., 1))
and PropertyOptimizer.forward / predict / compute_loss (:21-51) over a reference-keyed ``state_dict``; dropout is a fixed
multiplier per site (``drop[prefix + str(layer_index)]``), identity when absent.  Pinned to the live reference by
oracle/make_golden_property.py (<= 1e-12 in fp64) and tests/test_property_heads.py.
Only tests/ and __graft_entry__.smoke() may import this; the product never does.
"""
import torch


def regressor(sd, prefix, x, drop=None):
    """prefix e.g. 'homo_linear.'; layers are linear.0, linear.3, ... (ReLU / Dropout sit between, :57-60)"""
    idx = sorted({int(k[len(prefix) + len("linear."):].split(".")[0]) for k in sd if k.startswith(prefix + "linear.")})
    h = x
    for n, i in enumerate(idx):
        h = h @ sd["%slinear.%d.weight" % (prefix, i)].t() + sd["%slinear.%d.bias" % (prefix, i)]
        if n < len(idx) - 1:
            h = torch.relu(h)
            if drop and (prefix + str(i + 2)) in drop:
                h = h * drop[prefix + str(i + 2)].to(h.dtype)
    return h


def _flatten(x):                                                                       # :40-46
    if x.shape[-1] == 1:
        x = x[-1] if x.dim() == 1 else x[:, -1]
    return x


def predict(sd, homo_vecs, lumo_vecs, drop=None):
    return _flatten(regressor(sd, "homo_linear.", homo_vecs, drop)), _flatten(regressor(sd, "lumo_linear.", lumo_vecs, drop))


def forward(sd, homo_vecs, lumo_vecs, targets, drop=None):
    """-> (homo_loss, lumo_loss, homo_outputs, lumo_outputs), :25-37 with MSELoss(reduction='mean') (:21-23)"""
    yh, yl = predict(sd, homo_vecs, lumo_vecs, drop)
    return ((yh - targets[0]) ** 2).mean(), ((yl - targets[1]) ** 2).mean(), yh, yl
