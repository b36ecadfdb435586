"""Import the upstream reference (pepborrell/cool-chic) from /root/reference.

Only the golden-vector GENERATOR (`make_golden.py`) and ad-hoc validation scripts
run in the build container use this.  Nothing under `-m gpu` tests, `smoke()` or
`bench.py` imports it: /root/reference does not exist on the GPU box.

The reference imports `fvcore.nn` (absent in this image) purely to print MAC
counts (coolchic/enc/component/coolchic.py:15,611-650); an in-memory stub keeps
the numerics untouched.
"""

import collections
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("COOLCHIC_REFERENCE", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "coolchic"))


def import_reference():
    """Returns the `coolchic` top-level package of the reference."""
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    os.environ.setdefault("WANDB_MODE", "disabled")
    if "fvcore" not in sys.modules:
        fv, fvnn = types.ModuleType("fvcore"), types.ModuleType("fvcore.nn")

        class _FlopStub:
            def __init__(self, *a, **k):
                pass

            def unsupported_ops_warnings(self, *a):
                pass

            def uncalled_modules_warnings(self, *a):
                pass

            def total(self):
                return 0

            def by_module(self):
                return collections.defaultdict(int)

        fvnn.FlopCountAnalysis = _FlopStub
        fvnn.flop_count_table = lambda f: "fvcore-stub"
        fv.nn = fvnn
        sys.modules["fvcore"] = fv
        sys.modules["fvcore.nn"] = fvnn
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import coolchic  # noqa: F401

    return coolchic
