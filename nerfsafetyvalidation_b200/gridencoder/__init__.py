"""Drop-in for the reference's `gridencoder` package (gridencoder/grid.py): `grid_encode`, `GridEncoder`.

Forward only (inference hot path); the backward / dy_dx side (gridencoder.cu:227-343) is SURVEY.md section 8f row 3.
Differences that are not visible to callers:
  * the fp16 copy of the table that the reference re-creates on EVERY call under autocast (grid.py:38-39,
    46.7 MB read + 23.3 MB written) is cached and refreshed only when the parameter changes;
  * the kernel writes [B, L*C] directly, so the reference's permute+reshape copy (grid.py:52) disappears.
"""
import numpy as np
import torch
import torch.nn as nn

from .. import _cabi

_gridtype_to_id = {"hash": 0, "tiled": 1}
_DTYPE_ID = {torch.float32: 0, torch.float16: 1}


def grid_encode(inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs=False, gridtype=0,
                align_corners=False, ref_rounding=True):
    """grid.py:19-59 (forward).  inputs [B,3] float in [0,1]; embeddings [rows,C] float32 or float16;
    offsets int32 [L+1].  Returns [B, L*C] in the embeddings' dtype.  `ref_rounding` (half tables only)
    reproduces the reference kernel's round-to-half after every corner bit for bit."""
    lib = _cabi.load()
    if calc_grad_inputs or (torch.is_grad_enabled() and (inputs.requires_grad or embeddings.requires_grad)):
        raise NotImplementedError("grid_encode: backward / dy_dx are outside the inference hot path built here; "
                                  "call under torch.no_grad()")
    _cabi.require_cuda(inputs, embeddings, offsets)
    inputs = inputs.float().contiguous()
    B, D = inputs.shape
    L = offsets.shape[0] - 1
    C = embeddings.shape[1]
    S = float(np.log2(per_level_scale))
    if torch.is_autocast_enabled() and C % 2 == 0 and embeddings.dtype != torch.float16:
        embeddings = embeddings.to(torch.half)  # callers that care use GridEncoder, which caches this cast
    if embeddings.dtype not in _DTYPE_ID:
        raise RuntimeError(f"grid_encode: embeddings must be float32 or float16, got {embeddings.dtype}")
    embeddings = embeddings.contiguous()
    offsets = offsets.int().contiguous()
    outputs = torch.empty(B, L * C, device=inputs.device, dtype=embeddings.dtype)
    _cabi.check(lib.ngp_grid_encode_forward(_cabi.ptr(inputs), _cabi.ptr(embeddings), _DTYPE_ID[embeddings.dtype],
                                            _cabi.ptr(offsets), _cabi.ptr(outputs), B, D, C, L, S, int(base_resolution),
                                            int(gridtype), int(bool(align_corners)), 1, int(bool(ref_rounding)), None,
                                            _cabi.stream()), "grid_encode_forward")
    return outputs


class GridEncoder(nn.Module):
    """gridencoder/grid.py:93-156, same constructor, parameters (`embeddings`, buffer `offsets`) and forward."""

    def __init__(self, input_dim=3, num_levels=16, level_dim=2, per_level_scale=2, base_resolution=16,
                 log2_hashmap_size=19, desired_resolution=None, gridtype="hash", align_corners=False):
        super().__init__()
        if desired_resolution is not None:
            per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
        self.input_dim = input_dim
        self.num_levels = num_levels
        self.level_dim = level_dim
        self.per_level_scale = per_level_scale
        self.log2_hashmap_size = log2_hashmap_size
        self.base_resolution = base_resolution
        self.output_dim = num_levels * level_dim
        self.gridtype = gridtype
        self.gridtype_id = _gridtype_to_id[gridtype]
        self.align_corners = align_corners

        offsets, offset = [], 0
        self.max_params = 2 ** log2_hashmap_size
        for i in range(num_levels):
            resolution = int(np.ceil(base_resolution * per_level_scale ** i))
            params_in_level = min(self.max_params, (resolution if align_corners else resolution + 1) ** input_dim)
            params_in_level = int(np.ceil(params_in_level / 8) * 8)
            offsets.append(offset)
            offset += params_in_level
        offsets.append(offset)
        self.register_buffer("offsets", torch.from_numpy(np.array(offsets, dtype=np.int32)))
        self.n_params = offsets[-1] * level_dim
        self.embeddings = nn.Parameter(torch.empty(offset, level_dim))
        self._half = None  # (version, data_ptr, tensor): cached fp16 table
        self.reset_parameters()

    def reset_parameters(self):
        std = 1e-4
        self.embeddings.data.uniform_(-std, std)

    def __repr__(self):
        return (f"GridEncoder: input_dim={self.input_dim} num_levels={self.num_levels} level_dim={self.level_dim} "
                f"resolution={self.base_resolution} -> {int(round(self.base_resolution * self.per_level_scale ** (self.num_levels - 1)))} "
                f"per_level_scale={self.per_level_scale:.4f} params={tuple(self.embeddings.shape)} gridtype={self.gridtype} "
                f"align_corners={self.align_corners}")

    def half_table(self, fingerprint=None):
        """fp16 copy of `embeddings` (the reference re-casts 46.7 MB per call, grid.py:38-39), re-cast only when the parameter
        was modified or moved.  Writes through `.data` do not bump `_version`, so the key also carries a content fingerprint
        (_cabi.fingerprint: 32 sampled elements; the caller may pass one it computed for several tensors in a single read-back)."""
        e = self.embeddings
        key = (e._version, e.data_ptr(), _cabi.fingerprint(e) if fingerprint is None else fingerprint)
        if self._half is None or self._half[0] != key:
            self._half = (key, e.detach().to(torch.half).contiguous())
        return self._half[1]

    def invalidate_caches(self):
        self._half = None

    def _load_from_state_dict(self, *args, **kwargs):
        self.invalidate_caches()
        return super()._load_from_state_dict(*args, **kwargs)

    def _apply(self, fn, *args, **kwargs):
        self.invalidate_caches()
        return super()._apply(fn, *args, **kwargs)

    def forward(self, inputs, bound=1):
        inputs = (inputs + bound) / (2 * bound)  # grid.py:144
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.view(-1, self.input_dim)
        use_half = torch.is_autocast_enabled() and self.level_dim % 2 == 0
        table = self.half_table() if use_half else self.embeddings
        outputs = grid_encode(inputs, table, self.offsets, self.per_level_scale, self.base_resolution,
                              inputs.requires_grad, self.gridtype_id, self.align_corners)
        return outputs.view(prefix_shape + [self.output_dim])
