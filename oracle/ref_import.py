"""TEST INFRASTRUCTURE ONLY - never imported by the product package.

Imports the *unmodified* reference modules from /root/reference when that tree
exists (it does in the build container, it does NOT on the GPU box).  The
reference does `import gensim.downloader as api` at module top
(simple_two_tower.py:5, enhanced_two_tower.py:5) and gensim is not installed,
so two empty stub modules are registered first; nothing in the model / loss
classes touches them.
"""
import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("TT_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "enhanced_two_tower.py"))


def load_reference():
    """Return (simple_two_tower, margin_two_tower, enhanced_two_tower) modules."""
    if not reference_available():
        raise RuntimeError(f"reference tree not present at {REF_ROOT}")
    for name in ("gensim", "gensim.downloader"):
        sys.modules.setdefault(name, types.ModuleType(name))
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    return tuple(importlib.import_module(m) for m in
                 ("simple_two_tower", "margin_two_tower", "enhanced_two_tower"))
