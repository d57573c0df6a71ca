"""Model zoo: the architectures the reference scripts instantiate (SwinIR, the
sub-pixel SR CNN, the perceptual-loss VGG) and the ones BASELINE.json benchmarks
(ResNet-18/50, ViT-B/16)."""
from .perceptual import PerceptualLoss, VGGFeatures, feat_loss
from .resnet import ResNet, resnet18, resnet34, resnet50, resnet101
from .sr_net import Net
from .swinir import SwinIR
from .vit import VisionTransformer, vit_b_16, vit_s_16, vit_tiny

__all__ = ["ResNet", "resnet18", "resnet34", "resnet50", "resnet101", "VisionTransformer", "vit_b_16", "vit_s_16",
           "vit_tiny", "SwinIR", "Net", "PerceptualLoss", "VGGFeatures", "feat_loss"]
