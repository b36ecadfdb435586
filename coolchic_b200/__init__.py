"""coolchic_b200 -- B200-native (sm_100a) overfitting step for Cool-chic.

Drop-in for the hot path of pepborrell/cool-chic: the module tree mirrors the
reference (`coolchic.enc.component.{coolchic,frame,nlcoolchic}`, `coolchic.enc.training.{loss,train}`)
and every forward / backward / optimiser step runs in hand-written CUDA kernels behind the
C ABI of include/ccb.h.  There is no CPU fallback.
"""

from . import _native  # noqa: F401

__all__ = ["_native"]
