"""Drop-in for the reference's `ffmlp` package (ffmlp/ffmlp.py): `ffmlp_forward`, `FFMLP`.

Inference only.  fp16 operands, fp32 accumulation on tensor cores (the reference accumulates in fp16).
Superset behaviour kept from SURVEY.md section 7.2: inputs / weights are cast to half unconditionally (the reference only
casts under autocast and would fail CHECK_IS_HALF otherwise), and the physical 128-row zero padding of
ffmlp.py:156-165 is not materialised (the kernel masks the tail rows instead)."""
import math

import torch
import torch.nn as nn

from .. import _cabi


def convert_activation(act):
    return {"relu": 0, "exponential": 1, "sine": 2, "sigmoid": 3, "squareplus": 4, "softplus": 5}.get(act, 6)


def ffmlp_forward(inputs, weights, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                  inference=False, calc_grad_inputs=False):
    """ffmlp.py:15-48 (forward).  inputs [B,input_dim], weights flat; returns half [B, output_dim] (padded dim, 16)."""
    lib = _cabi.load()
    if not inference or calc_grad_inputs or (torch.is_grad_enabled() and (inputs.requires_grad or weights.requires_grad)):
        raise NotImplementedError("ffmlp_forward: only inference (module.eval() under torch.no_grad()) is built here")
    _cabi.require_cuda(inputs, weights)
    inputs = inputs.to(torch.half).contiguous()
    weights = weights.to(torch.half).contiguous()
    B = inputs.shape[0]
    outputs = torch.empty(B, output_dim, device=inputs.device, dtype=torch.half)
    _cabi.check(lib.ngp_ffmlp_inference(_cabi.ptr(inputs), _cabi.ptr(weights), B, int(input_dim), int(output_dim),
                                        int(hidden_dim), int(num_layers), int(activation), int(output_activation),
                                        _cabi.ptr(outputs), _cabi.stream()), "ffmlp_inference")
    return outputs


class FFMLP(nn.Module):
    """ffmlp/ffmlp.py:99-169: one flat fp32 parameter `weights`, layers [hidden][in], (num_layers-1) x [hidden][hidden],
    [16][hidden] row-major; init U(+-sqrt(3/hidden)) after torch.manual_seed(42) (it reseeds the GLOBAL RNG, kept)."""

    def __init__(self, input_dim, output_dim, hidden_dim, num_layers, activation="relu"):
        super().__init__()
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.hidden_dim = hidden_dim
        self.num_layers = num_layers
        self.activation = convert_activation(activation)
        self.output_activation = convert_activation("none")
        self.tensorcore_width = 16
        assert hidden_dim in [16, 32, 64, 128, 256], f"FFMLP only support hidden_dim in [16, 32, 64, 128, 256], but got {hidden_dim}"
        assert input_dim > 0 and input_dim % 16 == 0, f"FFMLP input_dim should be 16 * m (m  > 0), but got {input_dim}"
        assert output_dim <= 16, f"FFMLP current only supports output dim <= 16, but got {output_dim}"
        assert num_layers >= 2, f"FFMLP num_layers should be larger than 2 (3 matmuls), but got {num_layers}"
        self.padded_output_dim = int(math.ceil(output_dim / 16)) * 16
        self.num_parameters = hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + self.padded_output_dim)
        self.weights = nn.Parameter(torch.zeros(self.num_parameters))
        self._half = None
        self.reset_parameters()

    def cleanup(self):  # the reference frees its split-K streams here; we own none
        pass

    def __repr__(self):
        return (f"FFMLP: input_dim={self.input_dim} output_dim={self.output_dim} hidden_dim={self.hidden_dim} "
                f"num_layers={self.num_layers} activation={self.activation}")

    def reset_parameters(self):
        torch.manual_seed(42)
        std = math.sqrt(3 / self.hidden_dim)
        self.weights.data.uniform_(-std, std)

    def half_weights(self, fingerprint=None):
        """fp16 copy of `weights`, re-cast when the parameter was modified (see GridEncoder.half_table for the key)."""
        w = self.weights
        key = (w._version, w.data_ptr(), _cabi.fingerprint(w) if fingerprint is None else fingerprint)
        if self._half is None or self._half[0] != key:
            self._half = (key, w.detach().to(torch.half).contiguous())
        return self._half[1]

    def invalidate_caches(self):
        self._half = None

    def _load_from_state_dict(self, *args, **kwargs):
        self.invalidate_caches()
        return super()._load_from_state_dict(*args, **kwargs)

    def _apply(self, fn, *args, **kwargs):
        self.invalidate_caches()
        return super()._apply(fn, *args, **kwargs)

    def forward(self, inputs):
        B, C = inputs.shape
        outputs = ffmlp_forward(inputs, self.half_weights(), self.input_dim, self.padded_output_dim, self.hidden_dim,
                                self.num_layers, self.activation, self.output_activation, not self.training,
                                inputs.requires_grad)
        if self.padded_output_dim != self.output_dim:
            outputs = outputs[:B, :self.output_dim]
        return outputs
