"""Drop-in for the reference's utils/vae_utils.py (same names / argument meaning), on sm_100a kernels.
Additive: every sampler takes an optional `eps=` so noise can be injected."""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops


def sample_from_latent(mu: torch.Tensor, logvar: torch.Tensor, temperature: float = 1.0,
                       eps: Optional[torch.Tensor] = None) -> torch.Tensor:
    """utils/vae_utils.py:5-10: z = mu + eps * exp(0.5 logvar) * temperature."""
    if eps is None:
        eps = torch.randn_like(mu, dtype=torch.float32)
    return ops.reparameterize(mu.float(), logvar.float(), eps.float(), temperature)


def encode_images(model, images):
    """utils/vae_utils.py:13-25 (leaves the model in eval mode, like the reference)."""
    model.eval()
    with torch.no_grad():
        return model.encode(images)


def generate_predictions(model, images, temperature=1.0, num_samples=3, eps: Optional[torch.Tensor] = None):
    """utils/vae_utils.py:28-76: one encoder pass, `num_samples` decodes, running mean of the LOGITS
    at H/2 (no sigmoid, no final upsample). The N decodes run as one batched decoder pass.
    eps [N,B,L] multiplies exp(.5 logvar) * temperature (the reference scales eps by T, :49)."""
    model.eval()
    eng = model.engine()
    sample = getattr(model, "latent_injection", "all") != "none"       # :43
    with torch.no_grad():
        lg, mu, lv = eng.mc_logits_honour_skip(images.float().contiguous(), temperature, num_samples, eps, sample)
        B, Ns, h, w = lg.shape
        if Ns == 1:
            return lg.reshape(B, 1, h, w)
        mean = ops.uncertainty_reduce(lg.contiguous(), h, w, input_is_prob=True, outs=("mean",))["mean"]
        return mean.unsqueeze(1)


def calculate_latent_stats(mu, logvar):
    """utils/vae_utils.py:79-103 (host-side scalars for logging)."""
    mean_mu = mu.float().mean(dim=0)
    var = torch.exp(logvar.float())
    mean_var = var.mean(dim=0)
    active_dims = ((mean_mu.abs() > 0.1) | (mean_var < 0.9) | (mean_var > 1.1)).sum().item()
    total_dims = mu.shape[1]
    kl_per_dim = 0.5 * (mean_mu.pow(2) + mean_var - logvar.float().mean(dim=0) - 1)
    return {'active_dims': active_dims, 'total_dims': total_dims, 'activity_ratio': active_dims / total_dims,
            'total_kl': kl_per_dim.sum().item(), 'mean_mu_abs': mean_mu.abs().mean().item(),
            'mean_var': mean_var.mean().item()}
