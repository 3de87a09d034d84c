"""unet_b200 -- B200-native (sm_100a) per-frame cable-inspection hot path.

Host-side mirror of the reference's Python contracts over libunetb200.so:
  NestedUNet                      (src/models/unetpp.py)
  preprocess_image, map_roi_to_original, detect_burrs_on_cable, visualize_two_stage   (infer_two_stage_burr.py)
  detect_burrs_enhanced, enhance_image                           (infer_enhanced_burr.py)
  InspectionPipeline              batched body of the inference loop + multi-GPU sharding
"""
__all__ = ["NestedUNet", "InspectionPipeline", "preprocess_image", "map_roi_to_original",
           "detect_burrs_on_cable", "detect_burrs_enhanced", "enhance_image", "bgr2gray",
           "normalize_frame", "visualize_two_stage", "overlay_blend", "get_burr_mask_rulebased", "compute_diameter_metrics", "analyze_defects", "compute_thickness_profile", "filter_cable_by_shape", "constrain_tape_to_ring", "synth"]


def __getattr__(name):          # lazy: importing the package must not need torch.cuda
    if name == "NestedUNet":
        from .model import NestedUNet
        return NestedUNet
    if name == "InspectionPipeline":
        from .pipeline import InspectionPipeline
        return InspectionPipeline
    if name in ("preprocess_image", "map_roi_to_original", "detect_burrs_on_cable", "detect_burrs_enhanced",
                "enhance_image", "bgr2gray", "normalize_frame", "visualize_two_stage", "overlay_blend",
                "get_burr_mask_rulebased", "compute_diameter_metrics", "analyze_defects", "compute_thickness_profile",
                "filter_cable_by_shape", "constrain_tape_to_ring"):
        from . import ops
        return getattr(ops, name)
    raise AttributeError(name)
