"""Import the UNMODIFIED reference (`/root/reference`, read-only) in this container.

Only used by `tests/golden/make_golden.py` (fixture generation) and by the optional
CPU-side cross-checks that skip when `/root/reference` is absent (it does not exist on
the GPU box).  Four third-party modules the reference imports at package-import time are
not installed here (omegaconf, hydra, kaldiio, librosa); none is touched on the offline
Paraformer path when the config is passed inline, so empty stubs are injected
(SURVEY.md section 8c, Appendix A).
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("FUNASR_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "funasr"))


def import_reference():
    """Returns the reference `funasr` package (tables filled)."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    for name in ("kaldiio", "librosa", "hydra"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    if "omegaconf" not in sys.modules:
        om = types.ModuleType("omegaconf")
        om.DictConfig = type("DictConfig", (dict,), {})
        om.ListConfig = type("ListConfig", (list,), {})
        om.OmegaConf = type("OmegaConf", (), {})
        sys.modules["omegaconf"] = om
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import funasr  # noqa: F401  (fills funasr.register.tables)

    return funasr
