"""utils/functions.py:4-15 of the reference."""
import torch


def UnionPhiObs(phi_obs1, phi_obs2, alpha=0.0):
    """Union of two signed-distance fields: exact min for alpha == 1, else the smooth R-function
    (a + b - sqrt(a^2 + b^2 - 2 alpha a b)) / (1 + alpha)."""
    if alpha < 0.0 or alpha > 1.0:
        raise ValueError("alpha should be in the range [0,1].")
    if alpha == 1.0:
        return torch.min(phi_obs1, phi_obs2)
    a, b = phi_obs1, phi_obs2
    return 1.0 / (1.0 + alpha) * (a + b - torch.sqrt(a * a + b * b - 2 * alpha * a * b))
