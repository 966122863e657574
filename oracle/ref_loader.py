"""Loads the compiled reference core (oracle/_ref/mettagrid_c*.so). TEST INFRASTRUCTURE.

``load_mettagrid_c()`` works anywhere the .so travelled to (the GPU box included).
``load_reference_python()`` additionally imports the reference's Python package from
/root/reference -- only possible in the build container; used by tests/golden/make_golden.py.
"""

from __future__ import annotations

import glob
import importlib.machinery
import importlib.util
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def ref_so_path() -> str | None:
    # MG_REF_DIR=_ref_ordered selects the build with id-ordered clipper sets (oracle/build_ref_ordered.sh)
    hits = sorted(glob.glob(os.path.join(_HERE, os.environ.get("MG_REF_DIR", "_ref"), "mettagrid_c*.so")))
    return hits[0] if hits else None


def available() -> bool:
    return ref_so_path() is not None


def load_mettagrid_c():
    global _mod
    if _mod is not None:
        return _mod
    so = ref_so_path()
    if so is None:
        raise ImportError("oracle/_ref/mettagrid_c*.so not built (run oracle/build_ref.sh where /root/reference exists)")
    name = "mettagrid.mettagrid_c" if "mettagrid" in sys.modules else "mettagrid_c"
    loader = importlib.machinery.ExtensionFileLoader(name, so)
    spec = importlib.util.spec_from_file_location(name, so, loader=loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    _mod = mod
    return mod


def load_reference_python(root: str = "/root/reference"):
    """Import the reference's own Python package around the compiled core (container only)."""
    src = os.path.join(root, "packages", "mettagrid", "python", "src")
    if not os.path.isdir(src):
        raise ImportError(f"{src} not present")
    if src not in sys.path:
        sys.path.insert(0, src)
    cog = os.path.join(root, "packages", "cogames", "src")
    if os.path.isdir(cog) and cog not in sys.path:
        sys.path.insert(0, cog)
    for name in ("boto3", "botocore", "botocore.exceptions"):  # pulled in by mettagrid/util/file.py:12 only
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    exc = sys.modules["botocore.exceptions"]
    for n in ("NoCredentialsError", "ClientError"):
        if not hasattr(exc, n):
            setattr(exc, n, type(n, (Exception,), {}))
    import mettagrid  # noqa: F401

    mod = load_mettagrid_c() if _mod is not None else None
    if mod is None:
        so = ref_so_path()
        loader = importlib.machinery.ExtensionFileLoader("mettagrid.mettagrid_c", so)
        spec = importlib.util.spec_from_file_location("mettagrid.mettagrid_c", so, loader=loader)
        mod = importlib.util.module_from_spec(spec)
        sys.modules["mettagrid.mettagrid_c"] = mod
        loader.exec_module(mod)
        globals()["_mod"] = mod
    sys.modules["mettagrid.mettagrid_c"] = mod
    mettagrid.mettagrid_c = mod
    return mod
