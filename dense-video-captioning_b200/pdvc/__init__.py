"""PDVC (callers of the hot path + the glue needed to time a training step) on top of the CUDA MSDeformAttn modules.
Not imported by the package root: `from dense_video_captioning_b200.pdvc import build, PDVCConfig, Trainer`."""
from .model import PDVC, PDVCConfig, build
from .train import CONFIGS, Trainer, synthetic_video, to_device

__all__ = ["PDVC", "PDVCConfig", "build", "CONFIGS", "Trainer", "synthetic_video", "to_device"]
