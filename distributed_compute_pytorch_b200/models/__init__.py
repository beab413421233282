"""Model families: the reference ConvNet plus the north-star ResNet-18/50 and GPT-2 small."""
from .convnet import ConvNet  # noqa: F401
from .resnet import resnet18, resnet50, ResNet  # noqa: F401
from .gpt2 import GPT2, GPT2Config  # noqa: F401
