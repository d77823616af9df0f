"""Load the UNMODIFIED reference modules by file path (container only).

TEST INFRASTRUCTURE (see oracle/__init__.py).  /root/reference is read-only and
absent on the GPU box; anything that needs it must skip when ``available()`` is False.
Recipe: SURVEY.md Appendix C.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("F5_REFERENCE_ROOT", "/root/reference")
_SRC = os.path.join(REF_ROOT, "src", "f5_tts")


def available() -> bool:
    return os.path.isfile(os.path.join(_SRC, "model", "cfm.py"))


def _load(dotted, relpath):
    if dotted in sys.modules and getattr(sys.modules[dotted], "__file__", None):
        return sys.modules[dotted]
    spec = importlib.util.spec_from_file_location(dotted, os.path.join(_SRC, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[dotted] = mod
    spec.loader.exec_module(mod)
    return mod


def load():
    """Returns a namespace with the reference's CFM, DiT, UNetT, modules and utils."""
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    from . import shims

    shims.install()
    for pkg in ("f5_tts", "f5_tts.model", "f5_tts.model.backbones"):
        if pkg not in sys.modules:
            m = types.ModuleType(pkg)
            m.__path__ = []
            sys.modules[pkg] = m
    utils = _load("f5_tts.model.utils", "model/utils.py")
    modules = _load("f5_tts.model.modules", "model/modules.py")
    dit = _load("f5_tts.model.backbones.dit", "model/backbones/dit.py")
    unett = _load("f5_tts.model.backbones.unett", "model/backbones/unett.py")
    cfm = _load("f5_tts.model.cfm", "model/cfm.py")
    return types.SimpleNamespace(
        utils=utils, modules=modules, dit=dit, unett=unett, cfm=cfm, CFM=cfm.CFM, DiT=dit.DiT, UNetT=unett.UNetT
    )
