"""Ray choice (nnb_ray_choice, replaces torch.randperm(h*w)[:n] of model/training.py:277) and batch staging (nnb_copy_many).
Integer work: distinctness, range and byte-exact copies are exact checks; the scatter of the choice is a chi-square bound."""
import pytest
import torch

from util import dev

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_total,n_pick", [(465750, 4096), (465750, 465750), (1, 1), (2, 2), (5, 3), (4096, 4096), (1 << 20, 1024),
                                            ((1 << 20) + 1, 70000)])
def test_ray_choice_is_a_partial_permutation(n_total, n_pick):
    from model import _ops
    torch.manual_seed(3)
    a = _ops.ray_choice(n_total, n_pick, dev())
    assert a.dtype == torch.int64 and a.shape == (n_pick,)
    assert int(a.min()) >= 0 and int(a.max()) < n_total
    assert a.unique().numel() == n_pick, "indices must be distinct (sampling without replacement)"
    if n_pick == n_total:
        assert torch.equal(a.sort().values, torch.arange(n_total, device=dev()))
    # same generator state -> same choice; next draw -> another one
    torch.manual_seed(3)
    b = _ops.ray_choice(n_total, n_pick, dev())
    assert torch.equal(a, b)
    if n_total > 1000 and n_pick >= 1000:
        c = _ops.ray_choice(n_total, n_pick, dev())
        assert not torch.equal(a, c)


def test_ray_choice_scatter_is_uniform():
    """4,096 of 465,750 pixels, 400 draws: every one of 64 equal bins of the image receives its share (chi-square per draw
    around 63, never above 130 = p < 1e-6), the first pick is spread over the image, and no pixel is favoured across draws."""
    from model import _ops
    torch.manual_seed(11)
    n_total, n_pick, bins = 465750, 4096, 64
    counts = torch.zeros(n_total, device=dev())
    first = []
    worst = 0.0
    for _ in range(400):
        a = _ops.ray_choice(n_total, n_pick, dev())
        h = torch.bincount((a * bins) // n_total, minlength=bins).float()
        e = n_pick / bins
        worst = max(worst, float(((h - e) ** 2 / e).sum()))
        counts.index_add_(0, a, torch.ones(n_pick, device=dev()))
        first.append(int(a[0]))
    assert worst < 130.0, worst
    # per-pixel hit count over 400 draws ~ Binomial(400, 0.0088): mean 3.52, P(>= 20) ~ 1e-9 per pixel
    assert float(counts.max()) < 20 and abs(float(counts.mean()) - 400 * n_pick / n_total) < 1e-3
    f = torch.tensor(first, dtype=torch.float64) / n_total
    assert 0.42 < float(f.mean()) < 0.58 and float(f.min()) < 0.05 and float(f.max()) > 0.95


def test_copy_many_bytes_and_scalars():
    from model import _ops
    g = torch.Generator().manual_seed(2)
    shapes = [(1, 3, 375, 1242), (1, 375, 1242), (1, 4, 4), (7,), (0,), (1, 1, 33, 5)] * 4          # 24 > NNB_COPY_MAX: two launches
    src = [torch.randn(s, generator=g).to(dev()) for s in shapes]
    src[3] = torch.randint(0, 255, (7,), generator=g, dtype=torch.uint8).to(dev())                 # odd byte count
    src[5] = torch.randn(400, generator=g).to(dev())[1:166].view(1, 1, 33, 5)                        # 4-byte aligned only
    dst = [torch.zeros_like(s).contiguous() for s in src]
    a, b = torch.zeros(1, dtype=torch.int64, device=dev()), torch.zeros(1, dtype=torch.int64, device=dev())
    _ops.copy_many(list(zip(dst, src)), [(a, 17), (b, 1 << 40)])
    for d, s in zip(dst, src):
        assert torch.equal(d, s)
    assert int(a) == 17 and int(b) == 1 << 40
    _ops.copy_many([], [(a, 5)])
    assert int(a) == 5 and int(b) == 1 << 40


def test_mat4_mul_forward_and_backward_vs_torch():
    """nnb_mat4_mul_fwd / _bwd (relative pose, model/training.py:329,344) against torch's matmul and its autograd: fp32, 1e-6."""
    from model import _ops
    g = torch.Generator().manual_seed(9)
    a = torch.randn(4, 4, generator=g).to(dev()).requires_grad_()
    b = torch.randn(4, 4, generator=g).to(dev()).requires_grad_()
    w = torch.randn(4, 4, generator=g).to(dev())
    c = _ops.Mat4Mul.apply(a, b)
    (c * w).sum().backward()
    ga, gb = a.grad.clone(), b.grad.clone()
    a.grad = b.grad = None
    ref = a.double() @ b.double()
    (ref * w.double()).sum().backward()
    assert (c.double() - ref).abs().max() <= 1e-6 * float(ref.abs().max())
    assert (ga.double() - a.grad.double()).abs().max() <= 1e-6 * float(a.grad.abs().max())
    assert (gb.double() - b.grad.double()).abs().max() <= 1e-6 * float(b.grad.abs().max())
    # one-sided gradient (the reference frame's pose detached)
    a2 = a.detach().clone().requires_grad_()
    c2 = _ops.Mat4Mul.apply(a2, b.detach())
    (c2 * w).sum().backward()
    assert (a2.grad - ga).abs().max() == 0
