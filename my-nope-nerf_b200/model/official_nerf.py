"""OfficialStaticNerf -- drop-in for the reference's model/official_nerf.py.

Same constructor, same `nn.Linear` submodules and state_dict keys (layers0.{0,2,4,6}, layers1.{0,2,4,6},
fc_density, fc_feature, rgb_layers.0, fc_rgb; weights (out,in) row-major fp32), same bias overrides.
The module only OWNS the parameters: the contraction itself runs inside the fused render kernels
(`Renderer` hands `param_list()` to the C ABI), or through `forward()` for point queries.
"""
import torch
import torch.nn as nn

from . import _cabi, _ops


class OfficialStaticNerf(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        D = cfg['model']['hidden_dim']
        pos_enc_levels = cfg['model']['pos_enc_levels']
        dir_enc_levels = cfg['model']['dir_enc_levels']
        pos_in_dims = (2 * pos_enc_levels + 1) * 3
        dir_in_dims = (2 * dir_enc_levels + 1) * 3
        if (D, pos_in_dims, dir_in_dims) != (256, 63, 27):
            # the reference hard-codes L=10 / L=4 in infer_occ/forward (model/official_nerf.py:61,87); the kernels
            # are specialised for the one architecture every shipped config uses.
            raise ValueError("nope_nerf_b200 kernels support hidden_dim=256, pos_enc_levels=10, dir_enc_levels=4")
        self.white_bkgd = cfg['rendering']['white_background']
        self.dist_alpha = cfg['rendering']['dist_alpha']
        self.occ_activation = cfg['model']['occ_activation']

        self.layers0 = nn.Sequential(
            nn.Linear(pos_in_dims, D), nn.ReLU(), nn.Linear(D, D), nn.ReLU(),
            nn.Linear(D, D), nn.ReLU(), nn.Linear(D, D), nn.ReLU())
        self.layers1 = nn.Sequential(
            nn.Linear(D + pos_in_dims, D), nn.ReLU(), nn.Linear(D, D), nn.ReLU(),
            nn.Linear(D, D), nn.ReLU(), nn.Linear(D, D), nn.ReLU())
        self.fc_density = nn.Linear(D, 1)
        self.fc_feature = nn.Linear(D, D)
        self.rgb_layers = nn.Sequential(nn.Linear(D + dir_in_dims, D // 2), nn.ReLU())
        self.fc_rgb = nn.Linear(D // 2, 3)

        self.fc_density.bias.data = torch.tensor([0.1]).float()
        self.sigmoid = nn.Sigmoid()
        if self.white_bkgd:
            self.fc_rgb.bias.data = torch.tensor([0.8, 0.8, 0.8]).float()
        else:
            self.fc_rgb.bias.data = torch.tensor([0.02, 0.02, 0.02]).float()

    def param_list(self):
        """The 24 tensors in the order the C ABI expects (reference state_dict order)."""
        sd = dict(self.named_parameters())
        return [sd[k] for k in _cabi.PARAM_ORDER]

    def forward(self, p, ray_d=None, only_occupancy=False, return_logits=False, return_addocc=False,
                noise=False, it=100000, **kwargs):
        """Point query (M,3) [+ view dirs (M,3)] -> rgb (M,3) [, density/alpha (M,1)].  Forward only: the training
        gradient path goes through the fused ray kernels, not through point queries."""
        if p.requires_grad or (ray_d is not None and ray_d.requires_grad):
            raise NotImplementedError("differentiable point queries are not part of the hot path (SURVEY.md 8f rank 3)")
        with torch.no_grad():
            rgb, dens = _ops.mlp_points(self.param_list(), p.reshape(-1, 3), None if ray_d is None else ray_d.reshape(-1, 3),
                                        self.occ_activation == 'softplus', self.dist_alpha)
        if only_occupancy:
            return dens
        if ray_d is None:
            return None
        return (rgb, dens) if return_addocc else rgb

    def infer_occ(self, p):
        """model/official_nerf.py:60-67 -> (x, density).  The 256-wide trunk feature `x` never leaves the kernels; only the raw
        density (before the activation) is returned, `x` is None.  Forward only."""
        with torch.no_grad():
            _, raw = _ops.mlp_points_density_grad(self.param_list(), p.reshape(-1, 3), want_raw=True)
        return None, raw.reshape(-1, 1)

    def gradient(self, p, it):
        """model/official_nerf.py:46-58: -d raw_density / d p, shape (M,1,3).  First order only (the reference builds the graph
        for `normal_loss`, which is not supported here)."""
        with torch.no_grad():
            g = _ops.mlp_points_density_grad(self.param_list(), p.reshape(-1, 3))
        return g.unsqueeze(1)


def encode_position(input, levels, inc_input):
    """model/official_nerf.py:99-119 as a stand-alone function: (...,C) -> (..., C * (2 * levels + inc_input)), column order
    [x, sin(2^0 x), cos(2^0 x), sin(2^1 x), ...].  The render kernels compute the same columns in their prologue
    (csrc/mlp_tc.cu write_encodings); this one is for callers that want the encoding itself."""
    freqs = 2.0 ** torch.arange(levels, dtype=input.dtype, device=input.device)
    ang = input.unsqueeze(-2) * freqs.view(-1, 1)                           # (..., levels, C)
    waves = torch.stack([ang.sin(), ang.cos()], dim=-2)                     # (..., levels, 2, C)
    waves = waves.reshape(input.shape[:-1] + (-1,))
    return torch.cat([input, waves], dim=-1) if inc_input else waves
