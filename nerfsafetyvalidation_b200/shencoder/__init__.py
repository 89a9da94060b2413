"""Drop-in for the reference's `shencoder` package (shencoder/sphere_harmonics.py): `sh_encode`, `SHEncoder`.
Forward only, degree <= 4 (network_ff.py uses 4); always fp32 like `custom_fwd(cast_inputs=torch.float32)`."""
import torch
import torch.nn as nn

from .. import _cabi


def sh_encode(inputs, degree, calc_grad_inputs=False):
    lib = _cabi.load()
    if calc_grad_inputs or (torch.is_grad_enabled() and inputs.requires_grad):
        raise NotImplementedError("sh_encode: backward is outside the inference hot path built here")
    _cabi.require_cuda(inputs)
    inputs = inputs.float().contiguous()
    B, input_dim = inputs.shape
    if input_dim != 3:
        raise RuntimeError("SH encoder only support input dim == 3")
    outputs = torch.empty(B, degree ** 2, dtype=torch.float32, device=inputs.device)
    _cabi.check(lib.ngp_sh_encode_forward(_cabi.ptr(inputs), _cabi.ptr(outputs), B, int(degree), _cabi.stream()), "sh_encode_forward")
    return outputs


class SHEncoder(nn.Module):
    """shencoder/sphere_harmonics.py:60-86"""

    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        self.input_dim = input_dim
        self.degree = degree
        self.output_dim = degree ** 2
        assert self.input_dim == 3, "SH encoder only support input dim == 3"
        assert self.degree > 0 and self.degree <= 8, "SH encoder only supports degree in [1, 8]"

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        inputs = inputs / size
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.reshape(-1, self.input_dim)
        outputs = sh_encode(inputs, self.degree, inputs.requires_grad)
        return outputs.reshape(prefix_shape + [self.output_dim])
