"""Fused sm_100a executors (hand-written kernels only; see csrc/)."""
from __future__ import annotations


def make_fused_executor(net):
    name = net.base.name
    kinds = {l.spec.kind for l in net.layers}
    if name.startswith("densenet"):
        from .densenet import DenseNetExecutor
        return DenseNetExecutor(net)
    if name.startswith("mobilenetv2"):
        from .mobilenet import MobileNetV2Executor
        return MobileNetV2Executor(net)
    if name == "vgg16" or kinds <= {"input", "conv", "maxpool", "gap", "dense"}:
        from .vgg import ChainExecutor
        return ChainExecutor(net)
    if kinds <= {"input", "conv", "maxpool", "dropout", "flatten", "dense"}:
        from .small_cnn import SmallCnnExecutor
        return SmallCnnExecutor(net)
    raise NotImplementedError(
        f"no fused sm_100a pipeline for model {name!r}; pass executor='reference' to run the PyTorch "
        "reference executor on this device explicitly")
