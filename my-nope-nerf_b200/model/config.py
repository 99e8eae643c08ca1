"""get_model -- drop-in for the reference's model/config.py.  The DPT monocular depth network is an offline
preprocessing tool and is not rebuilt; asking for it raises."""
from .network import nope_nerf


def get_model(renderer, cfg, device=None, **kwargs):
    if cfg['depth']['type'] == 'DPT':
        raise NotImplementedError("depth.type == 'DPT' builds the offline DPT network, which is out of scope; "
                                  "precompute depth maps with the reference's preprocess/dpt_depth.py")
    return nope_nerf(cfg, renderer, None, device)
