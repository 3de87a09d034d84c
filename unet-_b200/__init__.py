"""unet_b200: B200-native per-frame cable inspection hot path."""
