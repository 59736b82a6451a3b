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


def load_object_memory():
    """Import the UNMODIFIED reference object_memory/object_memory.py (and object_info.py) as the package
    ``object_memory`` with import stubs for what is absent here (detector, LoRA module, imageio, Objectron IoU).
    The reference's own utils/{depth_utils,similarity_volume,fpfh_register}.py are loaded from the tree and run on the
    oracle's open3d shim.  -> (module object_memory.object_memory, module object_memory.object_info)."""
    if "object_memory.object_memory" in _cache:
        return _cache["object_memory.object_memory"], _cache["object_memory.object_info"]
    if not available():
        raise RuntimeError("reference tree not present")
    _stub("matplotlib")
    _stub("matplotlib.pyplot")
    _stub("imageio", imread=lambda p: None)
    o3d = _oracle_open3d()
    saved = {k: sys.modules.get(k) for k in ("open3d", "utils", "object_memory")}
    sys.modules["open3d"] = o3d
    try:
        def from_tree(name, rel):
            spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[name] = mod
            spec.loader.exec_module(mod)
            return mod
        utils_pkg = types.ModuleType("utils")
        utils_pkg.__path__ = []
        sys.modules["utils"] = utils_pkg
        _stub_force("utils.logging", conditional_log=lambda m, s: None, get_mem_stats=lambda: ("0", "0"))
        _stub_force("utils.IoU_ops", calculate_obj_aligned_3d_IoU=lambda *a, **k: 0.0)
        for m in ("depth_utils", "similarity_volume", "fpfh_register"):
            setattr(utils_pkg, m, from_tree(f"utils.{m}", f"utils/{m}.py"))
        pkg = types.ModuleType("object_memory")
        pkg.__path__ = []
        sys.modules["object_memory"] = pkg

        class _NoDetector:
            @staticmethod
            def setup(**kw):
                pass

            @staticmethod
            def find(*a, **k):
                raise RuntimeError("detector is out of scope")
        _stub_force("object_memory.object_finder", ObjectFinder=_NoDetector)
        _stub_force("object_memory.object_finder_phrases", check_if_floor=lambda t: False)
        _stub_force("object_memory.lora_module", LoraRevolver=object, LoraConfig=object)
        info = from_tree("object_memory.object_info", "object_memory/object_info.py")
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            om = from_tree("object_memory.object_memory", "object_memory/object_memory.py")
    finally:
        for k, v in saved.items():
            if v is not None:
                sys.modules[k] = v
            else:
                sys.modules.pop(k, None)
        for k in [k for k in sys.modules if k.startswith(("utils.", "object_memory."))]:
            sys.modules.pop(k, None)
    _cache["object_memory.object_memory"], _cache["object_memory.object_info"] = om, info
    return om, info


def _stub_force(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
