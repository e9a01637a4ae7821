"""torchbooster_b200 -- B200-native batch-transform + metrics hot path of TorchBooster.

Drop-in modules (same call signatures as the reference): ``lmdb``, ``dataset``, ``metrics``,
``distributed``.  Additions: ``transforms`` (``BatchTransform``), ``staging``.  All device work goes
through the C ABI in ``include/torchbooster_b200.h``; there is no CPU fallback.
"""
__version__ = "0.1.0"
