"""B200-native adversarial-SDF training step (drop-in for the reference's ``src.model``)."""
from .functional import upload_panel  # noqa: F401
from .model import AssetPricingGAN, GraphedStep, MacroLSTM, MomentNetwork, SDFNetwork  # noqa: F401
from .training import FusedClipAdam, train_epoch  # noqa: F401

__all__ = ["AssetPricingGAN", "SDFNetwork", "MomentNetwork", "MacroLSTM", "GraphedStep", "upload_panel",
           "FusedClipAdam", "train_epoch"]
