"""oracle/ref_loader.py -- TEST INFRASTRUCTURE ONLY.

Imports the UNMODIFIED reference modules from /root/reference (build container only; that path
does not exist on the GPU box) on top of the oracle's open3d shim, with import stubs for the
packages that are absent here (matplotlib).  Used to validate ref_port.py / o3d_core.py and to
generate tests/golden/ vectors.  Nothing is copied: modules are loaded from where they lie.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REF = os.environ.get("IBL_REFERENCE_ROOT", "/root/reference")
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "o3d_shim")
_cache = {}


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "utils"))


def _stub(name, **attrs):
    if name in sys.modules:
        return
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m


def _oracle_open3d():
    """Load the shim under a private name so it never shadows a product-side ``open3d``."""
    if "oracle_open3d" not in sys.modules:
        spec = importlib.util.spec_from_file_location("oracle_open3d", os.path.join(_SHIM, "open3d", "__init__.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["oracle_open3d"] = mod
        spec.loader.exec_module(mod)
    return sys.modules["oracle_open3d"]


def load(relpath: str):
    """Load e.g. 'utils/depth_utils.py' from the reference tree as a private module."""
    if relpath in _cache:
        return _cache[relpath]
    if not available():
        raise RuntimeError("reference tree not present")
    _stub("matplotlib")
    _stub("matplotlib.pyplot")
    o3d = _oracle_open3d()
    saved = sys.modules.get("open3d")
    sys.modules["open3d"] = o3d
    try:
        name = "ref_" + relpath.replace("/", "_").replace(".py", "")
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if saved is not None:
            sys.modules["open3d"] = saved
        else:
            sys.modules.pop("open3d", None)
    _cache[relpath] = mod
    return mod


def open3d():
    return _oracle_open3d()
