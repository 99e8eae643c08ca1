"""TEST INFRASTRUCTURE (not product code): a torch-CPU restatement of the MLP of model/official_nerf.py:60-96 that rounds
operands to bf16 at exactly the places the tensor-core kernels do (csrc/mlp_tc.cu), with fp32 accumulation:

  forward : every Linear sees bf16(activations) x bf16(weights); biases stay ~fp32 (hi/lo split in the kernel);
            activations are rounded to bf16 after the ReLU; the density and rgb heads read the fp32 accumulator;
  backward: data gradients use bf16(dY) x bf16(W); weight gradients use bf16(dY)^T x bf16(X); bias gradients sum bf16(dY).

It isolates kernel correctness from the bf16-vs-fp32 deviation: ReLU masks flip for pre-activations within rounding
distance of zero, which moves gradients by several percent relative to the fp32 reference without being a kernel
error.  tests/test_gpu_render_tc.py checks the CUDA path tightly against this emulation and loosely against the fp32
oracle (oracle/nope_oracle.py)."""
import torch
import torch.nn.functional as F

from .nope_oracle import posenc


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


class _RoundSTE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return bf(x)

    @staticmethod
    def backward(ctx, g):
        return g


class _LinearEmu(torch.autograd.Function):
    """y = x @ bf16(W)^T + b for x already bf16-representable."""

    @staticmethod
    def forward(ctx, x, W, b):
        ctx.save_for_backward(x, W)
        return x @ bf(W).t() + b

    @staticmethod
    def backward(ctx, dy):
        x, W = ctx.saved_tensors
        dyb = bf(dy)
        return dyb @ bf(W), dyb.t() @ x, dyb.sum(0)


class _HeadEmu(torch.autograd.Function):
    """Density / rgb heads: fp32 forward on the fp32 accumulator h; backward as the kernels do it:
    d_h exact fp32 (rgb head, `round_dh` False) or bf16(d_out) x bf16(W) (density head, folded into an MMA k-step);
    dW = bf16(d_out)^T bf16(h); db = sum d_out in fp32."""

    @staticmethod
    def forward(ctx, h, W, b, round_dh):
        ctx.save_for_backward(h, W)
        ctx.round_dh = round_dh
        return h @ W.t() + b

    @staticmethod
    def backward(ctx, d):
        h, W = ctx.saved_tensors
        dh = bf(d) @ bf(W) if ctx.round_dh else d @ W
        return dh, bf(d).t() @ bf(h), d.sum(0), None


def mlp_forward_bf16emu(P, pts, viewdir, occ_activation="softplus", dist_alpha=False):
    rnd, lin = _RoundSTE.apply, _LinearEmu.apply
    e = rnd(torch.cat([posenc(pts, 10)], dim=-1))
    h = e
    y = None
    for i in range(8):
        if i < 4:
            W, b = P["layers0.%d.weight" % (2 * i)], P["layers0.%d.bias" % (2 * i)]
        else:
            W, b = P["layers1.%d.weight" % (2 * (i - 4))], P["layers1.%d.bias" % (2 * (i - 4))]
        x = torch.cat([h, e], dim=-1) if i == 4 else h
        y = F.relu(lin(x, W, b))      # fp32 accumulator after ReLU
        h = rnd(y)                    # bf16 operand of the next layer
    raw = _HeadEmu.apply(y, P["fc_density.weight"], P["fc_density.bias"], True)
    dens = F.softplus(raw) if occ_activation == "softplus" else raw.relu()
    if not dist_alpha:
        dens = 1 - torch.exp(-1.0 * dens)
    feat = rnd(lin(h, P["fc_feature.weight"], P["fc_feature.bias"]))
    x = torch.cat([feat, rnd(posenc(viewdir, 4))], dim=-1)
    hc = F.relu(lin(x, P["rgb_layers.0.weight"], P["rgb_layers.0.bias"]))
    rgb = torch.sigmoid(_HeadEmu.apply(hc, P["fc_rgb.weight"], P["fc_rgb.bias"], False))
    return rgb, dens
