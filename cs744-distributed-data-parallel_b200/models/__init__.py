"""Model definitions (VGG family for CIFAR-sized inputs)."""
from .vgg import (VGG, VGG11, VGG13, VGG16, VGG19, build_model, conv_blocks,  # noqa: F401
                  param_specs, grad_ready_order, state_dict_keys)
