"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference from /root/reference.

This module exists only in service of `oracle/make_golden.py` and the oracle-pinning
tests that can run in the build container.  /root/reference does not exist on the GPU
box, so nothing under `-m gpu`, `smoke()` or `bench.py` imports this file.

The reference needs three packages at import time that are absent here
(`matplotlib`, `imageio`, `timm`).  Empty stub modules are
registered in `sys.modules`; nothing on the hot path calls into them.  Two shims
let the loss stack run on CPU without editing the reference:
  * `torch.Tensor.cuda` -> identity (model/losses.py:168 calls `.cuda()` unconditionally)
  * `transform_to_world` default device -> cpu (model/common.py:113 defaults to cuda)
"""
import importlib
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("NOPE_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "model"))


def _stub(name, attrs=()):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for a in attrs:
        setattr(m, a, None)
    sys.modules[name] = m
    return m


_REF = None


def load_reference():
    """Return the reference's `model` package (imported under its own name `model`).

    The caller must not have our drop-in `model` package on sys.path before this.
    """
    global _REF
    if _REF is not None:
        return _REF
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    mpl = _stub("matplotlib")
    plt = _stub("matplotlib.pyplot")
    mpl.pyplot = plt
    _stub("imageio")
    timm = _stub("timm")
    timm.create_model = None
    # cv2 / PIL are present in the image; stub only if missing.
    for name in ("cv2",):
        try:
            importlib.import_module(name)
        except Exception:
            _stub(name)
    # make sure a previously imported drop-in `model` does not shadow the reference
    for k in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
        del sys.modules[k]
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        # shim 1: .cuda() is an identity on a CPU-only box
        if not torch.cuda.is_available():
            torch.Tensor.cuda = lambda self, *a, **k: self
        ref_model = importlib.import_module("model")
        ref_common = importlib.import_module("model.common")
        ref_training = importlib.import_module("model.training")
        # shim 2: default device of transform_to_world when world_mat is None
        _orig_ttw = ref_common.transform_to_world

        def _ttw(pixels, depth, camera_mat, world_mat=None, scale_mat=None, invert=True, device=None):
            if device is None:
                device = pixels.device
            return _orig_ttw(pixels, depth, camera_mat, world_mat, scale_mat, invert=invert, device=device)

        ref_training.transform_to_world = _ttw
    finally:
        sys.path.remove(REFERENCE_ROOT)
    _REF = types.SimpleNamespace(
        model=ref_model,
        common=ref_common,
        training=ref_training,
        losses=importlib.import_module("model.losses"),
        rendering=importlib.import_module("model.rendering"),
        official_nerf=importlib.import_module("model.official_nerf"),
        poses=importlib.import_module("model.poses"),
        distortions=importlib.import_module("model.distortions"),
        network=importlib.import_module("model.network"),
    )
    return _REF


def unload_reference():
    """Drop the reference's `model` modules from sys.modules (so the drop-in can be imported)."""
    global _REF
    for k in [k for k in sys.modules if k == "model" or k.startswith("model.")]:
        del sys.modules[k]
    _REF = None
