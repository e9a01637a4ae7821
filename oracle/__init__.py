"""oracle/ -- CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in ``torchbooster_b200/`` imports this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and the CPU legs of ``bench.py`` (``cpu_baseline`` and
``--impl reference``) may import or execute it, and only as the checker / the timed CPU
baseline -- never as the product path.

What is restated, and what pins it
----------------------------------
The reference (yliess86/TorchBooster) is pure Python and holds no transform arithmetic of its
own: its examples build ``torchvision.transforms`` pipelines over PIL images
(``examples/img_cls/resnet/resnet.py:94-107``, ``examples/img_stt/adain/adain.py:167-171``) and
its metrics are three torch one-liners (``torchbooster/metrics.py:11-75``).  The arithmetic lives
in un-vendored, *unpinned* third-party code (``setup.py:5-16`` lists no requirements).  The
versions installed in this image are therefore declared the oracle versions:

    torch 2.11.0 (CPU ops) / torchvision 0.26.0 (PIL backend) / Pillow 12.2.0 / numpy 2.3.5

* ``transform_oracle``  -- literally calls torchvision's functional PIL ops per sample
  (``F.pad -> F.crop -> F.hflip -> F.resize -> F.center_crop -> F.to_tensor -> F.normalize``).
  That *is* the reference path, so it needs no pin.
* ``resize_oracle``     -- integer restatement of Pillow's 8 bpc antialiased bilinear resample
  (the file the CUDA kernel is written against).  Pinned: bit-exact against Pillow itself in
  ``tests/test_oracle.py`` and against ``tests/golden/resize_*.npz``.
* ``params_oracle``     -- torchvision's own ``RandomCrop.get_params`` / ``torch.rand(1) < p``
  draws under ``torch.manual_seed``.
* ``augment_oracle``    -- ``T.RandomRotation`` / ``T.RandAugment`` (``resnet.py:99-100``): torchvision's
  own ``F.rotate`` / ``autoaugment._apply_op`` and its own draws, plus a numpy restatement of
  Pillow's arithmetic for each operation, pinned bit-exact against torchvision live and against
  ``tests/golden/augment_torchvision.npz`` (the reference's ``T.Compose`` run as written).
* ``metrics_oracle``    -- restates ``torchbooster/metrics.py:27`` (accuracy) and ``:73-75``
  (RunningAverage).  Pinned: ``tests/golden/metrics_reference.npz`` was produced by importing
  the reference's ``torchbooster.metrics`` from ``/root/reference``
  (``tests/golden/make_golden.py``).  Top-k and the confusion matrix have no reference
  counterpart; they are defined by the rank rule and cross-checked against
  ``torch.topk`` / ``torch.bincount`` on tie-free data.
* LMDB (``torchbooster/lmdb.py``) -- py-lmdb is absent from this image and the reference holds
  no ``.mdb`` fixture: **parity unpinned** for the byte source (DESIGN.md says the same).

The reference's own test-suite (``test/test_config.py``) holds no golden vector for this path.
"""
