"""activation.py of the reference: trunc_exp (fp32 exp forward; the clamped backward belongs to training)."""
import torch


def trunc_exp(x):
    """activation.py:5-10: `custom_fwd(cast_inputs=torch.float32)` then exp."""
    return torch.exp(x.float())
