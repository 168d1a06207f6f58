"""CPU oracle for the loss / metric hot path — TEST INFRASTRUCTURE ONLY.

This package restates, with plain CPU torch ops issued in the reference's own
order, every function of the per-pixel loss + metric path of
RUA1027/Lowlight_Image_Enhancement.  It exists so
that the CUDA kernels can be checked against something; it is NOT a product
path and NOT a fallback:

* only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
  ``--impl reference`` legs of ``bench.py`` may import it;
* nothing under ``lowlight_image_enhancement_b200/`` imports it (a test
  enforces that), and the product raises when its CUDA library is missing.

Pinning status (DESIGN.md §Oracle has the table):

* ``oracle.psf_ops``  — PINNED: bit-compared with the reference modules
  ``metrics/phys_consistency.py``, ``NewBP_model/newbp_layer.py`` and
  ``NewBP_model/losses.py`` (imported from /root/reference by
  ``tests/golden/make_golden.py``; vectors committed under ``tests/golden``).
* ``oracle.quality``  — PSNR family and ``ssim_linear`` PINNED the same way
  (``metrics/linear.py``, ``metrics/channelwise.py``, ``metrics/psnr.py``).
  The kornia-0.6.12 SSIM loss and the torchmetrics-1.2.0 SSIM are restated
  from the published algorithms of those pinned releases; neither package is
  installed here, so those two policies are **parity unpinned** (cross-checked
  only against ``ssim_linear`` away from the borders).
* ``oracle.colour``   — ``_deltaE00_lab_map``, ``_ciede2000``, ``_sobel_magnitude``
  and ``torch.quantile`` semantics PINNED against the reference functions on
  the 18 Lab pairs of ``standard_tests/data/ciede2000_pairs.json`` and seeded
  tensors; kornia ``rgb_to_lab`` is restated → sRGB-input entry points are
  **parity unpinned** at the Lab conversion (loose cross-check: cv2 Lab).
* ``oracle.pixel_losses`` — PINNED against the reference's BasicSR loss classes
  (``tests/golden/make_golden_basicsr.py``).
* ``oracle.ingest``   — PINNED: the reference dataset's ``__getitem__`` arithmetic
  (bit-compared) and the ``metrics/evaluation.py`` tensor helpers
  (``tests/golden/make_golden_ingest.py``).
"""
