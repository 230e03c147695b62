"""Import the UNMODIFIED reference (`/root/reference`, read-only) in this container.

Used by `tests/golden/make_golden*.py` (fixture generation), by the tests that drive the reference's own
`AutoModel` (CPU cross-checks here; on the GPU box the drop-in test and the large-sample parity gate) and by
`bench.py --impl reference`.  `/root/reference` does not exist on the GPU box: there the unmodified copy under
`oracle/_ref/pyref` (written by `make -C oracle`, git-ignored) is imported instead.  Four third-party modules the reference imports at package-import time are
not installed here (omegaconf, hydra, kaldiio, librosa); none is touched on the offline
Paraformer path when the config is passed inline, so empty stubs are injected
(SURVEY.md section 8c, Appendix A).
"""
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
# /root/reference in the build container; on the GPU box the unmodified copy that `make -C oracle` put under
# oracle/_ref/pyref (git-ignored, travels with the snapshot)
_CANDIDATES = [os.environ.get("FUNASR_REFERENCE_ROOT", ""), "/root/reference",
               os.path.join(os.path.dirname(os.path.dirname(_HERE)), "oracle", "_ref", "pyref")]
REFERENCE_ROOT = next((c for c in _CANDIDATES if c and os.path.isdir(os.path.join(c, "funasr"))), "/root/reference")


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
