"""CheckpointIO -- same on-disk format as the reference's model/checkpoints.py: torch.save of
{module_name: state_dict, **scalars}; strict load_state_dict, so files are interchangeable with the reference."""
import os
import shutil
import time
import urllib.parse

import torch


class CheckpointIO(object):
    def __init__(self, checkpoint_dir='./chkpts', **kwargs):
        self.checkpoint_dir = checkpoint_dir
        self.module_dict = kwargs          # name -> module (or optimiser): everything with a state_dict
        os.makedirs(checkpoint_dir, exist_ok=True)

    def _path(self, filename):
        """Relative names live in the checkpoint directory (model/checkpoints.py:30-31,56-57)."""
        return filename if os.path.isabs(filename) else os.path.join(self.checkpoint_dir, filename)

    def register_modules(self, **kwargs):
        self.module_dict.update(kwargs)

    def save(self, filename, **kwargs):
        """model/checkpoints.py:29-41: the scalars (epoch_it, it, loss_val_best, ...) next to one state_dict per module."""
        payload = {name: module.state_dict() for name, module in self.module_dict.items()}
        torch.save({**kwargs, **payload}, self._path(filename))

    def backup_model_best(self, filename, **kwargs):
        """model/checkpoints.py:43-53: a COPY under a timestamp name; model_best.pt itself stays in place."""
        src = self._path(filename)
        if os.path.exists(src):
            dst_dir = os.path.join(self.checkpoint_dir, 'backup_model_best')
            os.makedirs(dst_dir, exist_ok=True)
            shutil.copy(src, os.path.join(dst_dir, '%s.pt' % time.time()))

    def load(self, filename, device=None, load_model_only=False):
        """model/checkpoints.py:55-64: an http(s) name goes through torch's model zoo, anything else is a file."""
        if is_url(filename):
            return self.load_url(filename)
        return self.load_file(filename, device, load_model_only)

    def load_url(self, url):
        from torch.utils import model_zoo
        return self.parse_state_dict(model_zoo.load_url(url, progress=True, check_hash=False))

    def load_file(self, filename, device=None, load_model_only=False):
        src = self._path(filename)
        if not os.path.exists(src):
            raise FileExistsError          # the exception type train.py:63-66 catches to start from scratch
        blob = torch.load(src, map_location='cpu' if device is None else device, weights_only=False)
        return self.parse_state_dict(blob, load_model_only)

    def parse_state_dict(self, state_dict, load_model_only=False):
        if load_model_only:
            # model/checkpoints.py:82-84: only the 'model' entry survives, so no scalars (epoch_it, it, ...) are returned and
            # the caller restarts its schedule
            state_dict = {'model': state_dict['model']}
        for name, module in self.module_dict.items():
            if name in state_dict:
                module.load_state_dict(state_dict[name])
            else:
                print('Warning: Could not find %s in checkpoint!' % name)
        return {k: v for k, v in state_dict.items() if k not in self.module_dict}


def is_url(name):
    return urllib.parse.urlparse(str(name)).scheme in ('http', 'https')
