"""b200cgan: B200-native cGAN patch-calibrator train step (drop-in for the reference cgan/ modules)."""
