"""picd_b200 -- a Blackwell (sm_100a) native data-parallel image-classification trainer.

Capabilities follow RanjanBalappa/pytorch-imageclassification-distributed
(reference: train.py, nn/classifier.py, dp/loader.py, ddp_utils.py, utils.py);
the architecture does not: NHWC bf16 activations, hand-written tcgen05/TMA
implicit-GEMM convolutions, fused BN/ReLU/residual kernels, and a gradient
all-reduce over NVLink peer memory fused with the optimizer update.

Sub-packages
    ops       kernels (ctypes -> libpicd_b200.so) with pure-torch oracles
    models    ResNet / Inception-v3 / EfficientNet-B3 / ViT-B/16 + Classifier
    parallel  process-group helpers, sharded sampler, P2P arena, fused DP engine
    data      folder dataset, augmentations, on-device synthetic source
    engine    train/val epoch loops, CLI
    utils     meters, checkpoint, tracing, clocks
"""
__version__ = "0.1.0"
