"""B200-native (sm_100a) loss / metric hot path of RUA1027/Lowlight_Image_Enhancement.

Drop-in modules (same names, signatures and error behaviour as the reference):

    lowlight_image_enhancement_b200.metrics.phys_consistency   <- metrics/phys_consistency.py
    lowlight_image_enhancement_b200.metrics.linear             <- metrics/linear.py
    lowlight_image_enhancement_b200.metrics.channelwise        <- metrics/channelwise.py
    lowlight_image_enhancement_b200.metrics.psnr               <- metrics/psnr.py
    lowlight_image_enhancement_b200.metrics.ssim               <- metrics/ssim.py
    lowlight_image_enhancement_b200.metrics.color_error        <- metrics/color_error.py
    lowlight_image_enhancement_b200.NewBP_model.newbp_layer    <- NewBP_model/newbp_layer.py
    lowlight_image_enhancement_b200.NewBP_model.losses         <- NewBP_model/losses.py
    lowlight_image_enhancement_b200.lowlight_metrics           <- NAFNet_base/basicsr/metrics/lowlight_metrics.py

All arithmetic runs in hand-written CUDA kernels behind the C ABI of include/llie_b200.h.
There is no CPU fallback: without the built library or without a CUDA device, calls raise.
"""
__version__ = "0.1.0"
