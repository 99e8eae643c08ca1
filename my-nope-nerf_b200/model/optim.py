"""FlatAdam -- step glue: `torch.optim.Adam` semantics (amsgrad False, weight_decay 0, maximize False,
the settings of train.py:59,100,118) with all parameters of a group living in ONE flat fp32 buffer, so that an optimiser
step is one launch of `nnb_adam_step` per group instead of a multi-tensor sweep.

Drop-in for the callers' use of the optimiser: `param_groups[i]['lr']` (MultiStepLR in train.py:303-346), `zero_grad`,
`state_dict` / `load_state_dict` in torch's Adam format (`step`, `exp_avg`, `exp_avg_sq` per parameter; the two moment
tensors are views into the flat buffers)."""
import torch

from . import _ops


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self._flat = []
        for group in self.param_groups:
            ps = group['params']
            for p in ps:
                if not p.is_cuda or p.dtype != torch.float32:
                    raise ValueError("FlatAdam needs fp32 CUDA parameters (no CPU path)")
            flat = torch.cat([p.detach().reshape(-1) for p in ps]).contiguous()
            m, v = torch.zeros_like(flat), torch.zeros_like(flat)
            offs, o = [], 0
            for p in ps:
                n = p.numel()
                p.data = flat[o:o + n].view_as(p)      # the parameter now lives inside the flat buffer
                self.state[p] = {'step': torch.tensor(0.0), 'exp_avg': m[o:o + n].view_as(p), 'exp_avg_sq': v[o:o + n].view_as(p)}
                offs.append(o)
                o += n
            self._flat.append(dict(flat=flat, m=m, v=v, offs=offs, gbuf=None))

    def _flat_grad(self, ps, fb):
        """The gradients as one contiguous buffer: used in place when they already are consecutive views of one
        allocation (RenderRays.backward hands out the 24 MLP gradients that way), else gathered with one `cat`."""
        g0 = ps[0].grad
        base = g0.data_ptr()
        if all(p.grad.dtype == torch.float32 and p.grad.is_contiguous() and p.grad.data_ptr() == base + 4 * o
               for p, o in zip(ps, fb['offs'])):
            return g0, base
        if fb['gbuf'] is None:
            fb['gbuf'] = torch.empty_like(fb['flat'])
        torch.cat([p.grad.reshape(-1).float() for p in ps], out=fb['gbuf'])
        return fb['gbuf'], fb['gbuf'].data_ptr()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group, fb in zip(self.param_groups, self._flat):
            ps = group['params']
            have = [p.grad is not None for p in ps]
            if not any(have):
                continue                                 # torch skips parameters without gradient
            b1, b2 = group['betas']
            steps = [int(self.state[p]['step']) for p in ps]
            if all(have) and len(set(steps)) == 1:
                keep, gptr = self._flat_grad(ps, fb)
                _ops.adam_step_raw(fb['flat'].data_ptr(), gptr, fb['m'].data_ptr(), fb['v'].data_ptr(), fb['flat'].numel(),
                                   group['lr'], b1, b2, group['eps'], steps[0] + 1)
                del keep
            else:                                        # mixed: parameter by parameter, each with its own step count
                for p, o, h, st in zip(ps, fb['offs'], have, steps):
                    if h:
                        g = p.grad.contiguous().float()
                        _ops.adam_step_raw(fb['flat'].data_ptr() + 4 * o, g.data_ptr(), fb['m'].data_ptr() + 4 * o,
                                           fb['v'].data_ptr() + 4 * o, p.numel(), group['lr'], b1, b2, group['eps'], st + 1)
            for p, h in zip(ps, have):
                if h:
                    self.state[p]['step'] += 1
                    torch.autograd.graph.increment_version(p)   # written through a raw pointer: tell version-keyed caches
        return loss

    def load_state_dict(self, state_dict):
        """Accepts a torch.optim.Adam / FlatAdam state dict; the moments are copied into the flat buffers."""
        old = {p: self.state[p] for g in self.param_groups for p in g['params']}
        super().load_state_dict(state_dict)
        for p, mine in old.items():
            new = self.state.get(p)
            if new is not None and new is not mine:
                mine['exp_avg'].copy_(new['exp_avg'])
                mine['exp_avg_sq'].copy_(new['exp_avg_sq'])
                mine['step'] = torch.tensor(float(new['step']))
            self.state[p] = mine
