"""s2lc_b200 — B200-native EfficientNet-UNet training step (hot path of
MaxWolf-01/sentinel2-landcover-classification), hand-written sm_100a CUDA behind a C ABI.

Public surface mirrors the reference for this path:
  efficientnet_unet.EfficientNetConfig / EfficientnetUnet   (src/modules/efficientnet_unet.py)
  losses.get_loss / FocalLoss / LossType                    (src/losses.py)
  train_segmentation.SegmentationStep / DataParallelTrainer (src/train_segmentation.py:109-147)
"""
__version__ = "0.1.0"
