"""TEST INFRASTRUCTURE ONLY — loads the UNMODIFIED reference from /root/reference.

Only `oracle/gen_golden.py` and container-side tests (`-m "not gpu"`, skipped when
/root/reference is absent) may import this file.  Nothing on the product path does.

The reference (`/root/reference/models/step1.py:6`) star-imports `models/utils.py`, whose
`from timm.models.layers import DropPath` (`models/utils.py:11`) is dead code but breaks
the import in this image (timm is not installed).  We inject an empty `timm.models.layers`
stub; no reference file is modified.  (SURVEY.md Appendix D.)

The reference's package is called `models`, the same import path our drop-in uses, so the
reference modules are imported under that name and then re-homed in `sys.modules` as
`ref_models.*`, restoring whatever `models` package was loaded before.
"""
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("NCONV_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "step1.py"))


_CACHE = {}


def load_reference():
    """Return (ref_step1_module, ref_step2_module) of the unmodified reference."""
    if "mods" in _CACHE:
        return _CACHE["mods"]
    if not reference_available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    import torch.nn as nn

    for name in ("timm", "timm.models", "timm.models.layers"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["timm.models.layers"].DropPath = nn.Identity

    # models/step2.py:8-9 imports torchvision; its import-time torch.library registrations
    # inspect every loaded module, which chokes on a namespace package called `models`, so
    # torchvision has to be fully imported before the reference's `models` appears.
    try:
        import torchvision.models  # noqa: F401
        import torchvision.transforms  # noqa: F401
    except ImportError:
        pass

    saved = {k: sys.modules.pop(k) for k in list(sys.modules)
             if k == "models" or k.startswith("models.")}
    # A regular package named `models` anywhere on sys.path (ours has an __init__.py) would
    # shadow the reference's namespace package, so hide such entries during the import.
    old_path = list(sys.path)
    sys.path[:] = [REFERENCE_ROOT] + [p for p in old_path
                                      if not os.path.isdir(os.path.join(p or ".", "models"))]
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            import models.step1 as r1  # noqa
            import models.step2 as r2  # noqa
    finally:
        sys.path[:] = old_path
        for k in list(sys.modules):
            if k == "models" or k.startswith("models."):
                sys.modules["ref_" + k] = sys.modules.pop(k)
        sys.modules.update(saved)
    _CACHE["mods"] = (r1, r2)
    return r1, r2


def load_reference_utils():
    """The unmodified /root/reference/utils.py (calculate_loss*, get_optimizer) as a module."""
    if "utils" in _CACHE:
        return _CACHE["utils"]
    if not os.path.isfile(os.path.join(REFERENCE_ROOT, "utils.py")):
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    import importlib.util
    load_reference()                      # installs the timm stub, imports torchvision first
    saved = {k: sys.modules.pop(k) for k in list(sys.modules)
             if k in ("models", "dataset", "utils") or k.startswith(("models.", "dataset."))}
    old_path = list(sys.path)
    sys.path[:] = [REFERENCE_ROOT] + [p for p in old_path
                                      if not os.path.isdir(os.path.join(p or ".", "models"))]
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REFERENCE_ROOT, "utils.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
    finally:
        sys.path[:] = old_path
        for k in list(sys.modules):
            if k in ("models", "dataset") or k.startswith(("models.", "dataset.")):
                sys.modules["ref_" + k] = sys.modules.pop(k)
        sys.modules.update(saved)
    _CACHE["utils"] = mod
    return mod


class two_arg_step1:
    """Context manager: the shipped `SETP1_NCONV.forward(self, x0_d)` takes one tensor but
    `SETP2_BP_*` call it with two (`models/step2.py:62,107`).  The intended
    semantics is "stack both depth frames on the batch axis" (matches `torch.cat((rgb0,
    rgb1), dim=0)` at `models/step2.py:63` and the `[0:1]/[1:2]` slicing at `:77,:126`).
    This patches the *class attribute in memory only* for the duration of the block."""

    def __enter__(self):
        import torch
        r1, _ = load_reference()
        self._cls = r1.SETP1_NCONV
        self._orig = self._cls.forward
        orig = self._orig

        def fwd(mod, a, b=None):
            return orig(mod, a if b is None else torch.cat((a, b), 0))

        self._cls.forward = fwd
        return self

    def __exit__(self, *exc):
        self._cls.forward = self._orig
        return False
