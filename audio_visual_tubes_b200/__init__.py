"""avtubes-b200: sm_100a CUDA kernels for the audio-visual similarity + hard-way
localisation head and the cIoU/AUC evaluation of tonymisic/audio-visual-tubes, behind the
reference's Python call signatures.  See DESIGN.md."""
from . import _lib                                       # noqa: F401
from .head import (HeadConfig, hardway_head, hardway_loss, audio_normalize,    # noqa: F401
                   hardway_value_and_grad, hardway_value_and_grad_host)
from . import dist                                       # noqa: F401
from .step import HeadStep, HostHeadPipeline                               # noqa: F401
from .losses import NPRatio, PropagationLoss, hardway16_objective   # noqa: F401
from .model import AVENet, FullModel, HardWayAttention   # noqa: F401
from .audio import log_spectrogram                      # noqa: F401
from .evaluation import (Evaluator, evaluate_maps, evaluate_videos, mTC, normalize_img, rasterise_gt,     # noqa: F401
                         testset_gt, testset_gt_frame)

__all__ = ["AVENet", "FullModel", "HardWayAttention", "HeadConfig", "hardway_head", "hardway_loss", "hardway_value_and_grad", "hardway_value_and_grad_host", "dist", "HeadStep", "HostHeadPipeline", "NPRatio", "PropagationLoss", "hardway16_objective",
           "log_spectrogram", "Evaluator", "evaluate_maps", "evaluate_videos", "mTC", "normalize_img", "rasterise_gt", "testset_gt",
           "testset_gt_frame"]
