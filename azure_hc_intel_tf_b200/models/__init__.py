"""Model registry — the ``model_config.get_model_config(name, 'imagenet')`` role of tf_cnn_benchmarks
(invoked through ``--model`` at benchmark-scripts/run-tf-sing-ucx-openmpi.sh:66)."""
from __future__ import annotations

from .builder import BuildConfig, Model
from . import resnet, zoo

_REGISTRY = {
    "resnet50": resnet.resnet50,
    "resnet50_v1.5": resnet.resnet50_v15,
    "resnet101": resnet.resnet101,
    "resnet152": resnet.resnet152,
    "vgg11": lambda cfg: zoo.VGG(cfg, (1, 1, 2, 2, 2), "vgg11"),
    "vgg16": lambda cfg: zoo.VGG(cfg, (2, 2, 3, 3, 3), "vgg16"),
    "vgg19": lambda cfg: zoo.VGG(cfg, (2, 2, 4, 4, 4), "vgg19"),
    "inception3": zoo.InceptionV3,
    "googlenet": zoo.GoogLeNet,
    "alexnet": zoo.AlexNet,
    "trivial": zoo.Trivial,
    "lenet": zoo.LeNet,
    "overfeat": zoo.Overfeat,
}


def available_models():
    return sorted(_REGISTRY)


def create_model(name: str, kernels: str = "cudnn", num_classes: int = 1001) -> Model:
    if name not in _REGISTRY:
        raise ValueError(f"unknown --model={name}; available: {', '.join(available_models())}")
    cfg = BuildConfig(kernels=kernels, num_classes=num_classes)
    return _REGISTRY[name](cfg)
