"""The drop-in boundary as the UNCHANGED callers see it: everything train.py, vis/render.py and
evaluation/eval.py import from `model` must resolve in the drop-in, and the host-side helpers of model.common must return what
the reference's functions of the same name return.  CPU only: nothing here launches a kernel."""
import ast
import json
import os
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

from util import ROOT, PKG

# `from model... import ...` / `import model ...` statements of the callers, as they stand in the reference:
# train.py:15,17   vis/render.py:11-14   evaluation/eval.py:12-15,21   evaluation/eval_poses.py:13   vis/vis_poses.py:7
CALLER_IMPORTS = {
    "train.py": ["import model as mdl", "from model.common import backup,  mse2psnr"],
    "vis/render.py": ["from model.checkpoints import CheckpointIO",
                      "from model.common import convert3x4_4x4,  interp_poses, interp_poses_bspline, generate_spiral_nerf",
                      "from model.extracting_images import Extract_Images", "import model as mdl"],
    "evaluation/eval.py": ["from model.checkpoints import CheckpointIO", "from model.common import compute_errors",
                           "from model.eval_images import Eval_Images", "import model as mdl", "from model.common import mse2psnr"],
    "evaluation/eval_poses.py": ["import model as mdl"],
    "vis/vis_poses.py": ["import model as mdl"],
}
# attributes of `mdl` those scripts touch
MDL_ATTRS = ["OfficialStaticNerf", "Renderer", "get_model", "LearnPose", "LearnFocal", "Learn_Distortion", "Trainer",
             "Trainer_pose", "CheckpointIO"]


def _model_imports_of(path):
    """The import statements of a script that name the `model` package, as source lines."""
    src = open(path).read()
    out = []
    for node in ast.walk(ast.parse(src)):
        if isinstance(node, ast.ImportFrom) and node.module and node.module.split(".")[0] == "model":
            out.append(ast.get_source_segment(src, node))
        elif isinstance(node, ast.Import) and any(a.name.split(".")[0] == "model" for a in node.names):
            out.append(ast.get_source_segment(src, node))
    return out


def test_caller_import_blocks_resolve_in_the_dropin():
    import model as mdl
    for script, lines in CALLER_IMPORTS.items():
        ns = {}
        for line in lines:
            exec(line, ns)      # ImportError here = the script would die at start-up with the drop-in on sys.path
    for a in MDL_ATTRS:
        assert hasattr(mdl, a), a


def test_recorded_import_blocks_are_the_callers_own():
    """Only where the reference is mounted: the recorded lines above still cover what the scripts import from `model`."""
    ref = os.environ.get("NOPE_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "model")):
        pytest.skip("reference tree not mounted")
    for script, lines in CALLER_IMPORTS.items():
        found = _model_imports_of(os.path.join(ref, script))
        ns = {}
        for l in found:
            exec(l, ns)
        norm = lambda s: " ".join(s.split())
        assert {norm(l) for l in found} <= {norm(l) for l in lines}, (script, found)


def test_generate_images_returns_the_keys_render_py_reads(tmp_path):
    """vis/render.py:104-109 reads out['img'], out['depth'], out['geo'], out['disp'] and np.stacks them; depth_out/<idx>.npy is
    written (model/extracting_images.py:113-114).  The render itself is stubbed: this checks the host side of the function."""
    import model as mdl
    ex = mdl.Extract_Images(types.SimpleNamespace(eval=lambda: None), {"extract_images": {"resolution": [6, 8]}}, device="cpu",
                            render_type="nope_nerf")
    g = torch.Generator().manual_seed(0)
    rgb, depth = torch.rand(6, 8, 3, generator=g), torch.rand(6, 8, generator=g) + 0.5
    ex.render_frame = lambda *a, **k: (rgb, depth)
    mdl.extracting_images._ops.Mat4Inverse = types.SimpleNamespace(apply=lambda m: torch.inverse(m))   # host stand-in for the 4x4 kernel
    try:
        data = {"img.camera_mat": torch.eye(4).unsqueeze(0), "img.scale_mat": torch.eye(4).unsqueeze(0), "img.idx": torch.tensor([3])}
        out = ex.generate_images(data, str(tmp_path), torch.eye(4).repeat(5, 1, 1), None, 0, False)
    finally:
        import importlib
        importlib.reload(mdl.extracting_images._ops)
    assert set(out) == {"img", "depth", "geo", "disp"} and out["geo"] is None
    assert out["img"].dtype == np.uint8 and out["img"].shape == (6, 8, 3)
    assert out["depth"].dtype == np.uint8 and out["depth"].shape == (6, 8)
    assert out["disp"].dtype == np.uint8 and out["disp"].shape == (6, 8, 3)
    np.testing.assert_array_equal(np.load(tmp_path / "depth_out" / "3.npy"), depth.numpy())
    for sub in ("img_out", "depth_out", "disp_out"):
        assert (tmp_path / sub / "0003.png").exists()
    # the reference's arithmetic on the same arrays (model/extracting_images.py:116-120)
    d = depth.numpy()
    np.testing.assert_array_equal(out["depth"], np.clip(255.0 / d.max() * (d - d.min()), 0, 255).astype(np.uint8))
    np.testing.assert_array_equal(out["img"], (rgb.numpy() * 255).astype(np.uint8))


_REF_HELPERS = r'''
import sys, json, io, contextlib
sys.path.insert(0, %r)
import numpy as np, torch
from oracle import ref_harness
C = ref_harness.load_reference().common
g = torch.Generator().manual_seed(3)
def poses(n):
    out = torch.eye(4).repeat(n, 1, 1)
    for i in range(n):
        a = torch.randn(3, generator=g) * 0.3
        out[i, :3, :3] = C.Exp(a)
        out[i, :3, 3] = torch.randn(3, generator=g)
    return out
P = poses(7)
times = np.array([0., 1., 2.5, 3., 4., 6., 7.])
hwf = np.tile(np.array([375., 1242., 725.]).reshape(1, 3, 1), (7, 1, 1))
res = {"P": P.tolist(),
       "interp": C.interp_poses(P, 11).tolist(),
       "bspline": C.interp_poses_bspline(P, 9, times, 3).tolist(),
       "at_times": C.get_poses_at_times(P, times, np.array([0.5, 2.0, 5.5])).tolist()}
with contextlib.redirect_stdout(io.StringIO()):
    res["spiral"] = C.generate_spiral_nerf(P, np.array([2., 4.]), 8, hwf).tolist()
res["spheric"] = C.create_spheric_poses(1.5, 0.2, 6).tolist()
gt = np.random.RandomState(0).rand(50) + 0.5
pr = np.random.RandomState(1).rand(50) + 0.5
res["errors"] = [float(v) for v in C.compute_errors(gt, pr)]
res["psnr"] = [float(C.mse2psnr(np.float64(m))) for m in (1e-3, 0.0, 0.25)]
pts = torch.randn(20, 3, generator=g) * 1.5
res["mip_in"] = pts.tolist(); res["mip"] = C.convert2mip(pts).tolist()
ro = torch.randn(16, 3, generator=g); ro[:, 2] -= 3.0
rd = torch.randn(16, 3, generator=g); rd[:, 2] = -rd[:, 2].abs() - 0.2
no, nd = C.get_ndc_rays_fxfy(torch.tensor([1.17, 3.87]), 1.0, ro, rd)
res["ndc"] = {"o": ro.tolist(), "d": rd.tolist(), "no": no.tolist(), "nd": nd.tolist()}
img = torch.rand(1, 2, 5, 7, generator=g); p = torch.rand(1, 9, 2, generator=g) * 5
res["gtv"] = {"img": img.tolist(), "p": p.tolist(), "v": C.get_tensor_values(img, p.clone(), mode="bilinear", align_corners=True).tolist()}
pw = torch.randn(1, 6, 3, generator=g)
cm = torch.diag(torch.tensor([1.2, -3.8, -1., 1.])).unsqueeze(0)
wm = torch.inverse(P[2]).unsqueeze(0)
res["cam"] = {"pw": pw.tolist(), "wm": wm.tolist(), "out": C.transform_to_camera_space(pw, cm, wm, torch.eye(4).unsqueeze(0)).tolist()}
pc = pw.clone(); pc[..., 2] = -pc[..., 2].abs() - 0.5
xy, valid = C.project_to_cam(pc, cm, torch.device("cpu"))
res["proj"] = {"pc": pc.tolist(), "xy": xy.tolist(), "valid": valid.tolist()}
v = torch.tensor([0.3, -1.2, 2.0])
res["skew"] = C.vec2skew(v).tolist()
# the loss terms the drop-in keeps as plain PyTorch (zero weight / False in every shipped config), model/losses.py:105-162,232-263
L = ref_harness.load_reference().losses
lo = L.Loss({"depth_loss_type": "l1", "with_ssim": True, "match_method": "dense"})
a = torch.rand(1, 9, 11, 3, generator=g); b = (a + 0.2 * torch.randn(1, 9, 11, 3, generator=g)).clamp(0, 1)
vm = torch.rand(1, 9, 11, 1, generator=g) > 0.3
tl = torch.randn(6, 3, generator=g)
d1, d2 = torch.rand(1, 40, generator=g), torch.rand(1, 40, generator=g)
rt, rtg = P[1:2], P[2:3]
w1, w2 = lo.get_weight_dist_loss(tl)
res["terms"] = {"a": a.tolist(), "b": b.tolist(), "vm": vm.tolist(), "tl": tl.tolist(), "d1": d1.tolist(), "d2": d2.tolist(),
                "ssim_map": L.compute_ssim_loss(a, b).tolist(), "rgb_s": float(lo.get_rgb_s_loss(a, b, vm)),
                "dist": [float(w1), float(w2)], "dc": [float(lo.get_depth_consistency_loss(d1, d2)), float(lo.get_depth_consistency_loss(d1, d2, d2 * 0.9, d1 * 1.1))],
                "tc": float(lo.get_t_cycle_loss(rt, rtg))}
# helpers the reference's modules call internally (sampling routines, positional encoding, SE(3) series, reprojection)
REF = ref_harness.load_reference()
rcfg = {"depth_range": [0.01, 10.0], "n_max_network_queries": 64000, "white_background": False}
rr = REF.rendering.Renderer(torch.nn.Identity(), rcfg, device=torch.device("cpu"))
zv = torch.linspace(0., 1., 16).view(1, 1, 16).repeat(1, 5, 1)
cw = torch.randn(1, 5, 3, generator=g) * 0.1; rv = torch.nn.functional.normalize(torch.randn(1, 5, 3, generator=g), dim=-1)
rv[..., 2] = -rv[..., 2].abs() - 0.2; cw[..., 2] -= 2.0
torch.manual_seed(11)
su = rr.sample_uniform(cw, rv, zv.clone(), True, [0.01, 10.0])
su0 = rr.sample_uniform(cw, rv, zv.clone(), False, [0.01, 10.0])
cmat = torch.diag(torch.tensor([1.17, -3.87, -1., 1.])).unsqueeze(0)
sn = rr.sample_ndc(cmat, cw, rv, zv.clone(), depth_range=[0., 1.])
x = torch.randn(6, 3, generator=g)
sm = torch.randn(7, generator=g) * 0.5
px = torch.rand(1, 12, 2, generator=g) * 1.6 - 0.8
dd = torch.rand(1, 12, 1, generator=g) + 1.0
rp, rpv = C.reprojection(px, dd, torch.inverse(P[3]).unsqueeze(0), torch.inverse(P[2]).unsqueeze(0), cm)
res["internal"] = {"zv": zv.tolist(), "cw": cw.tolist(), "rv": rv.tolist(),
    "su": [t.tolist() for t in su], "su0": [t.tolist() for t in su0], "sn": [t.tolist() for t in sn],
    "x": x.tolist(), "enc10": REF.official_nerf.encode_position(x, 10, True).tolist(), "enc4n": REF.official_nerf.encode_position(x, 4, False).tolist(),
    "homo": rr.transform_to_homogenous(pw).tolist(), "skew": C.skew_symmetric(x).tolist(), "sm": sm.tolist(),
    "tA": C.taylor_A(sm).tolist(), "tB": C.taylor_B(sm).tolist(), "tC": C.taylor_C(sm).tolist(),
    "px": px.tolist(), "dd": dd.tolist(), "rp": rp.tolist(), "rpv": rpv.tolist(),
    "mom": float(lo.mean_on_mask((a - b).abs(), vm))}
# LearnFocal parametrisations (model/intrinsics.py)
import importlib
FI = importlib.import_module("model.intrinsics")
FOCAL_CASES = [(True, False, 2, [1.3, 2.1]), (True, False, 1, [1.3, 2.1]), (True, True, 2, 1.7), (False, True, 1, 1.7), (True, False, 2, None), (True, True, 1, None), (True, False, 2, 0.9)]
res["focal"] = []
for rq, fxo, order, init in FOCAL_CASES:
    m = FI.LearnFocal(rq, fxo, order=order, init_focal=init)
    out = m(0)
    (out * torch.tensor([2.0, 3.0])).sum().backward() if rq else None
    res["focal"].append({"out": out.tolist(), "keys": sorted(m.state_dict().keys()), "req": [p.requires_grad for p in m.parameters()],
                         "grads": [None if p.grad is None else p.grad.tolist() for p in m.parameters()]})
# the SSIM evaluation/eval.py reports (third_party/pytorch_ssim, outside the model package)
sys.path.insert(0, ref_harness.REFERENCE_ROOT)
from third_party import pytorch_ssim
sys.path.remove(ref_harness.REFERENCE_ROOT)
i1 = torch.rand(1, 3, 24, 32, generator=g); i2 = (i1 + 0.1 * torch.randn(1, 3, 24, 32, generator=g)).clamp(0, 1)
res["ssim_eval"] = {"i1": i1.tolist(), "i2": i2.tolist(), "v": float(pytorch_ssim.ssim(i1, i2)), "vn": float(pytorch_ssim.ssim(i1, i2, use_padding=False)),
                    "per": pytorch_ssim.ssim(torch.cat([i1, i2]), torch.cat([i2, i2]), size_average=False).tolist()}
print(json.dumps(res))
'''


def test_host_helpers_match_the_reference(capsys):
    from oracle import ref_harness
    if not ref_harness.reference_available():
        pytest.skip("reference tree not mounted")
    r = subprocess.run([sys.executable, "-c", _REF_HELPERS % ROOT], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]
    ref = json.loads(r.stdout.strip().splitlines()[-1])
    from model import common as C
    T = lambda x: torch.tensor(x, dtype=torch.float32)
    P = T(ref["P"])
    times = np.array([0., 1., 2.5, 3., 4., 6., 7.])
    hwf = np.tile(np.array([375., 1242., 725.]).reshape(1, 3, 1), (7, 1, 1))
    close = lambda a, b: torch.testing.assert_close(a.float(), T(b), rtol=1e-6, atol=1e-6)
    close(C.interp_poses(P, 11), ref["interp"])
    close(C.interp_poses_bspline(P, 9, times, 3), ref["bspline"])
    close(C.get_poses_at_times(P, times, np.array([0.5, 2.0, 5.5])), ref["at_times"])
    close(C.generate_spiral_nerf(P, np.array([2., 4.]), 8, hwf), ref["spiral"])
    capsys.readouterr()
    np.testing.assert_allclose(C.create_spheric_poses(1.5, 0.2, 6), np.array(ref["spheric"]), rtol=1e-12, atol=1e-12)
    gt, pr = np.random.RandomState(0).rand(50) + 0.5, np.random.RandomState(1).rand(50) + 0.5
    np.testing.assert_allclose([float(v) for v in C.compute_errors(gt, pr)], ref["errors"], rtol=1e-12)
    assert [float(C.mse2psnr(np.float64(m))) for m in (1e-3, 0.0, 0.25)] == ref["psnr"]
    assert C.mse2psnr(np.float64(0.1)).dtype == np.float32
    close(C.convert2mip(T(ref["mip_in"])), ref["mip"])
    no, nd = C.get_ndc_rays_fxfy(torch.tensor([1.17, 3.87]), 1.0, T(ref["ndc"]["o"]), T(ref["ndc"]["d"]))
    close(no, ref["ndc"]["no"])
    close(nd, ref["ndc"]["nd"])
    close(C.get_tensor_values(T(ref["gtv"]["img"]), T(ref["gtv"]["p"]), mode="bilinear", align_corners=True), ref["gtv"]["v"])
    cm = torch.diag(torch.tensor([1.2, -3.8, -1., 1.])).unsqueeze(0)
    close(C.transform_to_camera_space(T(ref["cam"]["pw"]), cm, T(ref["cam"]["wm"]), torch.eye(4).unsqueeze(0)), ref["cam"]["out"])
    xy, valid = C.project_to_cam(T(ref["proj"]["pc"]), cm, torch.device("cpu"))
    close(xy, ref["proj"]["xy"])
    assert valid.tolist() == ref["proj"]["valid"]
    close(C.vec2skew(torch.tensor([0.3, -1.2, 2.0])), ref["skew"])
    # plain-PyTorch loss terms (with_ssim photometric term, pose smoothness, depth consistency, t-cycle)
    from model import losses as ML
    t = ref["terms"]
    lo = ML.Loss({"depth_loss_type": "l1", "with_ssim": True, "match_method": "dense"})
    a, b, vm = T(t["a"]), T(t["b"]), torch.tensor(t["vm"])
    close(ML.compute_ssim_loss(a, b), t["ssim_map"])
    assert abs(float(lo.get_rgb_s_loss(a, b, vm)) - t["rgb_s"]) < 1e-6
    assert float(lo.get_rgb_s_loss(a, b, torch.zeros_like(vm))) == 0.0
    w1, w2 = lo.get_weight_dist_loss(T(t["tl"]))
    assert abs(float(w1) - t["dist"][0]) < 1e-6 and abs(float(w2) - t["dist"][1]) < 1e-6
    d1, d2 = T(t["d1"]), T(t["d2"])
    assert abs(float(lo.get_depth_consistency_loss(d1, d2)) - t["dc"][0]) < 1e-5
    assert abs(float(lo.get_depth_consistency_loss(d1, d2, d2 * 0.9, d1 * 1.1)) - t["dc"][1]) < 1e-5
    assert abs(float(lo.get_t_cycle_loss(P[1:2], P[2:3])) - t["tc"]) < 1e-5
    # helpers the reference's modules call internally
    from model import rendering as MR, official_nerf as MN
    q = ref["internal"]
    rcfg = {"depth_range": [0.01, 10.0], "n_max_network_queries": 64000, "white_background": False, "num_points": 16, "outside_steps": 0}
    rr = MR.Renderer(torch.nn.Identity(), rcfg, device=torch.device("cpu"))
    zv, cw, rv = T(q["zv"]), T(q["cw"]), T(q["rv"])
    torch.manual_seed(11)
    for got, want in zip(rr.sample_uniform(cw, rv, zv.clone(), True, [0.01, 10.0]), q["su"]):
        close(got, want)
    for got, want in zip(rr.sample_uniform(cw, rv, zv.clone(), False, [0.01, 10.0]), q["su0"]):
        close(got, want)
    cmat = torch.diag(torch.tensor([1.17, -3.87, -1., 1.])).unsqueeze(0)
    for got, want in zip(rr.sample_ndc(cmat, cw, rv, zv.clone(), depth_range=[0., 1.]), q["sn"]):
        close(got, want)
    x = T(q["x"])
    torch.testing.assert_close(MN.encode_position(x, 10, True), T(q["enc10"]), rtol=0, atol=0)
    torch.testing.assert_close(MN.encode_position(x, 4, False), T(q["enc4n"]), rtol=0, atol=0)
    close(rr.transform_to_homogenous(T(ref["cam"]["pw"])), q["homo"])
    close(C.skew_symmetric(x), q["skew"])
    sm = T(q["sm"])
    close(C.taylor_A(sm), q["tA"]); close(C.taylor_B(sm), q["tB"]); close(C.taylor_C(sm), q["tC"])
    rp, rpv = C.reprojection(T(q["px"]), T(q["dd"]), torch.inverse(P[3]).unsqueeze(0), torch.inverse(P[2]).unsqueeze(0), cm)
    torch.testing.assert_close(rp, T(q["rp"]), rtol=1e-5, atol=1e-5)
    assert rpv.tolist() == q["rpv"]
    assert abs(float(lo.mean_on_mask((a - b).abs(), vm)) - q["mom"]) < 1e-6
    # the SSIM of the evaluation loop
    from model import eval_images as ME
    e = ref["ssim_eval"]
    i1, i2 = T(e["i1"]), T(e["i2"])
    assert abs(float(ME.ssim(i1, i2)) - e["v"]) < 1e-6 and abs(float(ME.ssim(i1, i2, use_padding=False)) - e["vn"]) < 1e-6
    close(ME.ssim(torch.cat([i1, i2]), torch.cat([i2, i2]), size_average=False), e["per"])
    # LearnFocal: every parametrisation (order 1 / 2, fx only, list / scalar / no initial focal)
    import model as mdl
    cases = [(True, False, 2, [1.3, 2.1]), (True, False, 1, [1.3, 2.1]), (True, True, 2, 1.7), (False, True, 1, 1.7), (True, False, 2, None), (True, True, 1, None), (True, False, 2, 0.9)]
    for (rq, fxo, order, init), want in zip(cases, ref["focal"]):
        m = mdl.LearnFocal(rq, fxo, order=order, init_focal=init)
        out = m(0)
        if rq:
            (out * torch.tensor([2.0, 3.0])).sum().backward()
        close(out.detach(), want["out"])
        assert sorted(m.state_dict().keys()) == want["keys"] and [p.requires_grad for p in m.parameters()] == want["req"]
        for p, gw in zip(m.parameters(), want["grads"]):
            assert (p.grad is None) == (gw is None)
            if gw is not None:
                close(p.grad, gw)


def test_backup_copies_config_and_sources(tmp_path, monkeypatch):
    """model/common.py:492-506."""
    from model import common as C
    work = tmp_path / "work"
    (work / "model").mkdir(parents=True)
    (work / "dataloading").mkdir()
    (work / "configs").mkdir()
    for f in ("train.py", "configs/default.yaml", "model/a.py", "dataloading/b.py", "my.yaml"):
        (work / f).write_text(f)
    monkeypatch.chdir(work)
    C.backup(str(tmp_path / "out"), "my.yaml")
    b = tmp_path / "out" / "backup"
    assert (b / "config.yaml").read_text() == "my.yaml"
    for f in ("train.py", "default.yaml", "model/a.py", "dataloading/b.py"):
        assert (b / f).exists(), f


def test_checkpoint_semantics(tmp_path):
    """load_model_only returns no scalars (model/checkpoints.py:82-84); backup_model_best copies and leaves the file in place
    (model/checkpoints.py:43-53)."""
    import model as mdl
    lin, other = torch.nn.Linear(2, 2), torch.nn.Linear(3, 1)
    io = mdl.CheckpointIO(str(tmp_path), model=lin, optimizer=other)
    io.save("model.pt", epoch_it=7, it=123, loss_val_best=0.5)
    assert io.load("model.pt") == {"epoch_it": 7, "it": 123, "loss_val_best": 0.5}
    assert io.load("model.pt", load_model_only=True) == {}
    io.save("model_best.pt", it=1)
    io.backup_model_best("model_best.pt")
    assert (tmp_path / "model_best.pt").exists()
    assert len(os.listdir(tmp_path / "backup_model_best")) == 1


def test_precision_default_and_override(monkeypatch):
    import model as mdl
    import synthetic_scene as syn
    cfg = syn.default_cfg()
    mk = lambda rc, **kw: mdl.Renderer(mdl.OfficialStaticNerf(cfg), rc, device="cpu", **kw)
    monkeypatch.delenv("NNB_PRECISION", raising=False)
    assert mk(cfg["rendering"]).precision == "bf16"                                   # 128 samples: the tensor-core tile fits
    assert mk(dict(cfg["rendering"], num_points=64)).precision == "fp32"
    assert mk(dict(cfg["rendering"], precision="fp32")).precision == "fp32"
    assert mk(cfg["rendering"], precision="fp32").precision == "fp32"
    monkeypatch.setenv("NNB_PRECISION", "fp32")
    assert mk(cfg["rendering"]).precision == "fp32"
    with pytest.raises(ValueError):
        mk(dict(cfg["rendering"], precision="fp16"))


def test_objects_build_from_the_reference_yaml_like_train_py():
    """train.py:50-158 builds the model stack from configs/default.yaml (+ a scene config).  Every key the drop-in reads must
    exist in the reference's YAML, and nothing the YAML lacks (precision, ray_choice) may be required.  CPU: construction only."""
    import yaml
    ref = os.environ.get("NOPE_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "configs")):
        pytest.skip("reference tree not mounted")
    import model as mdl
    cfg = yaml.safe_load(open(os.path.join(ref, "configs", "default.yaml")))
    scene_cfg = yaml.safe_load(open(os.path.join(ref, "configs", "V_KITTI", "intersection_d5.yaml")))
    for sec, kv in scene_cfg.items():          # dataloading.load_config: the scene file overrides the defaults section by section
        if isinstance(kv, dict):
            cfg.setdefault(sec, {}).update(kv)
    n_views = 8
    model = mdl.OfficialStaticNerf(cfg)
    renderer = mdl.Renderer(model, cfg['rendering'], device="cpu")
    nope_nerf = mdl.get_model(renderer, cfg, device="cpu")
    opt = torch.optim.Adam(nope_nerf.parameters(), lr=cfg['training']['learning_rate'], weight_decay=cfg['training']['weight_decay'])
    pose = mdl.LearnPose(n_views, cfg['pose']['learn_R'], cfg['pose']['learn_t'], cfg=cfg, init_c2w=None)
    opt_pose = torch.optim.Adam(pose.parameters(), lr=cfg['training']['pose_lr'])
    dist = mdl.Learn_Distortion(n_views, cfg['distortion']['learn_scale'], cfg['distortion']['learn_shift'], cfg)
    opt_dist = torch.optim.Adam(dist.parameters(), lr=cfg['training']['distortion_lr'])
    focal = mdl.LearnFocal(cfg['pose']['learn_focal'], cfg['pose']['fx_only'], order=cfg['pose']['focal_order'])
    trainer = mdl.Trainer(nope_nerf, opt, cfg['training'], device="cpu", optimizer_pose=opt_pose, pose_param_net=pose,
                          optimizer_focal=None, focal_net=None, optimizer_distortion=opt_dist, distortion_net=dist, cfg_all=cfg)
    assert renderer.precision in ("bf16", "fp32") and trainer.ray_choice == "feistel" and trainer.use_graph == "auto"
    assert trainer.n_training_points == cfg['training']['n_training_points']
    ex = mdl.Extract_Images(renderer, cfg, use_learnt_poses=cfg['pose']['learn_pose'], use_learnt_focal=cfg['pose']['learn_focal'],
                            device="cpu", render_type=cfg['rendering']['type'])
    assert list(ex.resolution) == list(cfg['extract_images']['resolution'])
    ck = mdl.CheckpointIO("/tmp/nnb_ckpt_probe", model=nope_nerf, optimizer=opt)
    assert focal is not None and ck is not None
    # the annealing schedule of train.py / model/training.py:204-212 on the YAML's own weights
    w = trainer.anneal(cfg['training']['pc_weight'][0], cfg['training']['pc_weight'][1], cfg['training']['scheduling_start'],
                       cfg['training']['annealing_epochs'], cfg['training']['scheduling_start'] + cfg['training']['annealing_epochs'] // 2)
    assert min(cfg['training']['pc_weight']) <= w <= max(cfg['training']['pc_weight'])
