"""B200-native adversarial-SDF training step (drop-in for the reference's ``src.model``)."""
from .data import AssetPricingDataset, create_data_loaders, create_small_sample  # noqa: F401
from .functional import upload_panel  # noqa: F401
from .model import AssetPricingGAN, GraphedStep, MacroLSTM, MomentNetwork, SDFNetwork  # noqa: F401
from .training import FusedClipAdam, train_epoch  # noqa: F401

__all__ = ["AssetPricingGAN", "SDFNetwork", "MomentNetwork", "MacroLSTM", "GraphedStep", "upload_panel",
           "FusedClipAdam", "train_epoch", "AssetPricingDataset", "create_data_loaders", "create_small_sample"]
