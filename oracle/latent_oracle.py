"""TEST INFRASTRUCTURE ONLY — CPU restatement of the latent head.

Follows HierVAE.rsample (ggpm/hpgnn.py:34-41) = PropertyVAE.rsample (ggpm/property_vae.py:26-33), with the noise passed
in instead of drawn (the reference draws ``torch.randn_like(z_mean)`` at :39 whether or not ``perturb`` is set).
Pinned to the live reference by oracle/make_golden_latent.py (<= 1e-12 in fp64) and tests/test_latent.py.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this; the product never does.
"""
import torch


def rsample(z_vecs, w_mean, b_mean, w_var, b_var, epsilon, perturb=True):
    batch_size = z_vecs.size(0)                                                        # :35
    z_mean = z_vecs @ w_mean.t() + b_mean                                              # :36
    z_log_var = -torch.abs(z_vecs @ w_var.t() + b_var)                                 # :37
    kl_loss = -0.5 * torch.sum(1.0 + z_log_var - z_mean * z_mean - torch.exp(z_log_var)) / batch_size   # :38
    z = z_mean + torch.exp(z_log_var / 2) * epsilon if perturb else z_mean             # :40
    return z, kl_loss
