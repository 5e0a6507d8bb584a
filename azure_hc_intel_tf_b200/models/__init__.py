"""Model registry — the ``model_config.get_model_config(name, 'imagenet')`` role of tf_cnn_benchmarks
(invoked through ``--model`` at benchmark-scripts/run-tf-sing-ucx-openmpi.sh:66)."""
from __future__ import annotations

from .builder import BuildConfig, Model
from . import mobilenet, nasnet, resnet, zoo, zoo2

_REGISTRY = {
    "resnet50": resnet.resnet50,
    "resnet50_v1.5": resnet.resnet50_v15,
    "resnet101": resnet.resnet101,
    "resnet152": resnet.resnet152,
    # the TF official-models ResNets tf_cnn_benchmarks exposes as official_resnet*: 18 / 34 are basic-block networks,
    # 50 / 101 / 152 are the v1.5 form (stride in the 3x3), the _v2 names the pre-activation form
    "official_resnet18": resnet.official_resnet18,
    "official_resnet34": resnet.official_resnet34,
    "official_resnet50": lambda cfg: resnet.ResNet(cfg, (3, 4, 6, 3), True, "official_resnet50"),
    "official_resnet101": lambda cfg: resnet.ResNet(cfg, (3, 4, 23, 3), True, "official_resnet101"),
    "official_resnet152": lambda cfg: resnet.ResNet(cfg, (3, 8, 36, 3), True, "official_resnet152"),
    "official_resnet50_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 4, 6, 3), "official_resnet50_v2"),
    "official_resnet101_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 4, 23, 3), "official_resnet101_v2"),
    "official_resnet152_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 8, 36, 3), "official_resnet152_v2"),
    "vgg11": lambda cfg: zoo.VGG(cfg, (1, 1, 2, 2, 2), "vgg11"),
    "vgg16": lambda cfg: zoo.VGG(cfg, (2, 2, 3, 3, 3), "vgg16"),
    "vgg19": lambda cfg: zoo.VGG(cfg, (2, 2, 4, 4, 4), "vgg19"),
    "inception3": zoo.InceptionV3,
    "googlenet": zoo.GoogLeNet,
    "alexnet": zoo.AlexNet,
    "trivial": zoo.Trivial,
    "lenet": zoo.LeNet,
    "overfeat": zoo.Overfeat,
    "inception4": zoo2.InceptionV4,
    "mobilenet": mobilenet.MobileNetV2,
    "nasnet": nasnet.nasnet_mobile,
    "nasnetlarge": nasnet.nasnet_large,
    "mobilenet_v2": lambda cfg: mobilenet.MobileNetV2(cfg, "mobilenet_v2"),
    "resnet50_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 4, 6, 3), "resnet50_v2"),
    "resnet101_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 4, 23, 3), "resnet101_v2"),
    "resnet152_v2": lambda cfg: zoo2.ResNetV2(cfg, (3, 8, 36, 3), "resnet152_v2"),
    # CIFAR-10 models (32x32 input, 10 classes + the unused background class)
    **{f"resnet{d}": (lambda cfg, d=d: zoo2.ResNetCifar(cfg, d, False, f"resnet{d}")) for d in (20, 32, 44, 56, 110)},
    **{f"resnet{d}_v2": (lambda cfg, d=d: zoo2.ResNetCifar(cfg, d, True, f"resnet{d}_v2")) for d in (20, 32, 44, 56, 110)},
    "densenet40_k12": lambda cfg: zoo2.DenseNetCifar(cfg, (12, 12, 12), 12, "densenet40_k12"),
    "densenet100_k12": lambda cfg: zoo2.DenseNetCifar(cfg, (32, 32, 32), 12, "densenet100_k12"),
    "densenet100_k24": lambda cfg: zoo2.DenseNetCifar(cfg, (32, 32, 32), 24, "densenet100_k24"),
}

_CIFAR = {n for n in _REGISTRY if n.startswith("densenet") or (n.startswith("resnet") and n.split("_")[0][6:] in
                                                                ("20", "32", "44", "56", "110"))}


def dataset_of(name: str) -> str:
    """'cifar10' or 'imagenet': which synthetic input shape / class count a model is defined for."""
    return "cifar10" if name in _CIFAR else "imagenet"


def available_models():
    return sorted(_REGISTRY)


def create_model(name: str, kernels: str = "cudnn", num_classes: int = None) -> Model:
    if name not in _REGISTRY:
        raise ValueError(f"unknown --model={name}; available: {', '.join(available_models())}")
    if num_classes is None:
        num_classes = 11 if dataset_of(name) == "cifar10" else 1001
    cfg = BuildConfig(kernels=kernels, num_classes=num_classes)
    return _REGISTRY[name](cfg)
