"""Data parallelism over real NCCL: N ranks, one process per GPU, the all-reduced gradient of the
sharded step against the one-rank gradient of the same global batch.  Needs >= 2 GPUs (skipped otherwise; bench.py --gpus N
reports the same figure as `extras.dp_parity`, so the driver's scaling run carries it for 2 / 4 / 8 ranks)."""
import json
import os
import subprocess
import sys

import pytest
import torch

from util import ROOT

pytestmark = pytest.mark.gpu

WORKER = r'''
import json, os, sys
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "my-nope-nerf_b200"))
import torch
import synthetic_scene as syn
import bench_extras as X
from model import parallel
dp = parallel.init_from_env()
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
cfg = syn.default_cfg()
scene = syn.make_scene("cfg1")
out = X.dp_parity(dev, scene, cfg, dp, rays_global=1021)       # 1,021 rays: unequal shards
if dp.rank == 0:
    print("DP_PARITY " + json.dumps(out))
torch.distributed.barrier()
torch.distributed.destroy_process_group()
'''


def test_nccl_allreduced_gradient_equals_one_rank_gradient(tmp_path):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (bench.py --gpus N reports extras.dp_parity)")
    world = 2 if n < 4 else 4
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("DP_PARITY ")][-1]
    res = json.loads(line[len("DP_PARITY "):])
    print(json.dumps(res, indent=1))
    for key, tol in (("fp32_l1", 2e-5), ("fp32_invariant", 1e-3), ("bf16_l1", 1e-3), ("bf16_invariant", 2e-3)):
        assert res[key]["loss_rel_diff"] <= 1e-5, (key, res[key])
        assert res[key]["grad_rel_l2"] <= tol, (key, res[key])
