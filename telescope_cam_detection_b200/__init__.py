"""B200-native per-frame detection hot path of telescope_cam_detection.

Same class names and call surface as the reference's src/ modules;
the work underneath runs in libtcd.so (hand-written sm_100a CUDA behind a C ABI).
"""
__version__ = "0.1.0"

_LAZY = {
    "YOLOXDetector": ".yolox_detector",
    "InferenceEngine": ".inference_engine_yolox",
    "SharedInferenceCoordinator": ".shared_inference_coordinator",
    "PendingInference": ".shared_inference_coordinator",
    "MotionFilter": ".motion_filter",
    "AdaptiveMotionFilter": ".motion_filter",
    "EmptyFrameFilter": ".empty_frame_filter",
    "TwoStageDetectionPipeline": ".two_stage_pipeline_yolox",
    "SpeciesClassifier": ".species_classifier",
    "RTDETRDetector": ".rtdetr_detector",
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        return getattr(importlib.import_module(_LAZY[name], __name__), name)
    raise AttributeError(name)
