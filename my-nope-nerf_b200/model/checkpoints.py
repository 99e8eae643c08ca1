"""CheckpointIO -- same on-disk format as the reference's model/checkpoints.py: torch.save of
{module_name: state_dict, **scalars}; strict load_state_dict, so files are interchangeable with the reference."""
import datetime
import os
import shutil

import torch


class CheckpointIO(object):
    def __init__(self, checkpoint_dir='./chkpts', **kwargs):
        self.module_dict = kwargs
        self.checkpoint_dir = checkpoint_dir
        if not os.path.exists(checkpoint_dir):
            os.makedirs(checkpoint_dir)

    def register_modules(self, **kwargs):
        self.module_dict.update(kwargs)

    def save(self, filename, **kwargs):
        if not os.path.isabs(filename):
            filename = os.path.join(self.checkpoint_dir, filename)
        outdict = dict(kwargs)
        for k, v in self.module_dict.items():
            outdict[k] = v.state_dict()
        torch.save(outdict, filename)

    def backup_model_best(self, filename, **kwargs):
        if not os.path.isabs(filename):
            filename = os.path.join(self.checkpoint_dir, filename)
        if os.path.exists(filename):
            backup_dir = os.path.join(self.checkpoint_dir, 'backup_model_best')
            if not os.path.exists(backup_dir):
                os.makedirs(backup_dir)
            # a COPY under a timestamp name (model/checkpoints.py:43-53): model_best.pt itself stays in place
            ts = datetime.datetime.now().timestamp()
            shutil.copy(filename, os.path.join(backup_dir, '%s.pt' % ts))

    def load(self, filename, device=None, load_model_only=False):
        if not os.path.isabs(filename):
            filename = os.path.join(self.checkpoint_dir, filename)
        if not os.path.exists(filename):
            raise FileExistsError
        state_dict = torch.load(filename, map_location=device if device is not None else 'cpu', weights_only=False)
        return self.parse_state_dict(state_dict, load_model_only)

    def parse_state_dict(self, state_dict, load_model_only=False):
        if load_model_only:
            # model/checkpoints.py:82-84: only the 'model' entry survives, so no scalars (epoch_it, it, ...) are returned and
            # the caller restarts its schedule
            state_dict = {'model': state_dict['model']}
        for k, v in self.module_dict.items():
            if k in state_dict:
                v.load_state_dict(state_dict[k])
            else:
                print('Warning: Could not find %s in checkpoint!' % k)
        return {k: v for k, v in state_dict.items() if k not in self.module_dict}
