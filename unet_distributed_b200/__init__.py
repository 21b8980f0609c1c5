"""unet_distributed_b200 -- the UNet training hot path of dmsuehir/unet_distributed on B200 (sm_100a).

Host side mirrors the reference's Python surface (model.define_model, dice_coef/dice_coef_loss, data.get_epoch,
settings_dist, the test_dist flags); all arithmetic runs in libunetb200.so (hand-written CUDA) behind a C ABI
declared in include/unetb200.h.  Importing this package never builds or falls back: a missing library raises."""

__all__ = ["model", "data", "preprocess", "settings_dist", "plan", "trainer", "parallel", "optim", "synthetic"]
__version__ = "0.1.0"
