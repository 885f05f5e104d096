"""Model zoo: the reference's backbones (nn/classifier.py:11-23) plus the driver's
ResNet-18 and ViT-B/16, all built from the fused ops in :mod:`picd_b200.ops`."""
from .classifier import BACKBONES, Classifier
from .efficientnet import EfficientNet, efficientnet
from .inception import Inception3, InceptionOutputs, inception_v3
from .layers import (BatchNorm2d, Conv2d, Dropout, LayerNorm, Linear, MLPHead, ReLU, SyncBatchNorm,
                     conv_bn_act, convert_sync_batchnorm, prepare_input)
from .resnet import ResNet, resnet18, resnet34, resnet50, resnet101
from .vit import VisionTransformer, vit_b_16, vit_tiny_test

__all__ = [n for n in dir() if not n.startswith("_")]
