"""b200cgan: B200-native cGAN patch-calibrator train step (drop-in for the reference's cgan/ modules).

    from b200cgan.models import GeneratorUNet, Discriminator, weights_init_normal
    from b200cgan.losses import HybridLoss, apply_delta_to_bbox, iou_metric, compute_gradient_penalty
    from b200cgan.train_step import CGANTrainer

All device arithmetic runs in libb2c.so (hand-written sm_100a CUDA behind the C-ABI in include/b2c.h); there is no
CPU, cuDNN/cuBLAS or Triton fallback."""
