import torch
dev = torch.device("cuda")
n = 465750
out = torch.empty(4096, dtype=torch.int64, device=dev)
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(2):
        out.copy_(torch.randperm(n, device=dev)[:4096])
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        idx = torch.randperm(n, device=dev)[:4096]
        noise = torch.rand(1, 4096, 128, device=dev)
        out.copy_(idx)
    g.replay(); torch.cuda.synchronize(); a = out.clone(); na = noise.clone()
    g.replay(); torch.cuda.synchronize(); b = out.clone()
    print("randperm capturable; draws differ between replays:", not torch.equal(a, b), "unique:", a.unique().numel(), "noise differs:", not torch.equal(na, noise))
except Exception as e:
    print("randperm NOT capturable:", repr(e)[:300])
