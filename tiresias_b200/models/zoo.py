"""Model registry: every name in the reference's checkpoint-size table
(model/model_factory.py:19-55) and per-tensor table (core/models.py:8-20) resolves to an
executable random-init model plus the synthetic batch shape it trains on.

Names with their own architecture here: resnet50/101/152, vgg11/16/19, alexnet, googlenet, inception3,
mobilenet_v1_025 (models/vision.py, models/inception.py), transformer, BERT_Chinese and the three OpenNMT
LSTM checkpoints (models/transformer.py, models/rnn.py) — 15 of the table's 17 names, each within a few per cent
of its checkpoint size (tests/test_models_cpu.py), so that the communication the reference prices as
"model size / bandwidth" moves the right number of bytes. inception2 and inception4 remain *size-class
stand-ins* (``STAND_IN``): the architecture here whose gradient volume is closest.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Tuple

import torch
from torch import nn

from . import inception, rnn, transformer, vision

# Checkpoint sizes in MB, verbatim data from model/model_factory.py:19-55.
MODEL_SIZES_MB: Dict[str, float] = {
    "4_layers_brnn": 1300, "transformer": 1100, "1_layer_bilstm_opennmt": 900, "BERT_Chinese": 350,
    "2_layers_lstm_gigaword": 330, "mobilenet_v1_025": 15, "googlenet": 26, "inception2": 43,
    "inception3": 104, "inception4": 176, "alexnet": 233, "vgg11": 519, "vgg19": 549, "vgg16": 528,
    "resnet50": 97, "resnet101": 555, "resnet152": 737,
}
CNN_MODELS = ("mobilenet_v1_025", "googlenet", "inception2", "inception3", "inception4", "alexnet",
              "vgg11", "vgg16", "vgg19", "resnet50", "resnet101", "resnet152")
CUDA_INIT_SIZE_MB = 525           # model/model_factory.py:66


@dataclass(frozen=True)
class ModelSpec:
    name: str
    build: Callable[[], nn.Module]
    kind: str                      # "image" | "tokens"
    sample_shape: Tuple[int, ...]  # per-sample input shape (H, W, C) or (seq,)
    num_classes: int


def _img(name: str, fn: Callable[[], nn.Module]) -> ModelSpec:
    return ModelSpec(name, fn, "image", (224, 224, 3), 1000)


def _tok(name: str, fn: Callable[[], nn.Module], seq: int = 128, vocab: int = 32000) -> ModelSpec:
    return ModelSpec(name, fn, "tokens", (seq,), vocab)


REGISTRY: Dict[str, ModelSpec] = {
    "resnet50": _img("resnet50", vision.resnet50),
    "resnet101": _img("resnet101", vision.resnet101),
    "resnet152": _img("resnet152", vision.resnet152),
    "vgg11": _img("vgg11", vision.vgg11),
    "vgg16": _img("vgg16", vision.vgg16),
    "vgg19": _img("vgg19", vision.vgg19),
    "alexnet": _img("alexnet", vision.alexnet),
    "googlenet": _img("googlenet", inception.googlenet),
    "inception3": _img("inception3", inception.inception3),
    "mobilenet_v1_025": _img("mobilenet_v1_025", inception.mobilenet_v1_025),
    "2_layers_lstm_gigaword": _tok("2_layers_lstm_gigaword", rnn.lstm_gigaword, seq=64, vocab=50000),
    "1_layer_bilstm_opennmt": _tok("1_layer_bilstm_opennmt", rnn.bilstm_opennmt, seq=64, vocab=70000),
    "4_layers_brnn": _tok("4_layers_brnn", rnn.brnn_4_layers, seq=64, vocab=90000),
    "BERT_Chinese": _tok("BERT_Chinese", rnn.bert_chinese, seq=128, vocab=21128),
    "transformer": _tok("transformer", transformer.transformer_base),
    "transformer_base": _tok("transformer_base", transformer.transformer_base),
    "transformer_big": _tok("transformer_big", transformer.transformer_big),
}

STAND_IN: Dict[str, str] = {"inception2": "googlenet", "inception4": "inception3"}


def build(spec: ModelSpec, fused: bool = True) -> nn.Module:
    """Instantiate ``spec``; ``fused=False`` asks for the stock-PyTorch variant (baseline arm) when the
    builder has one."""
    import inspect
    if "fused" in inspect.signature(spec.build).parameters:
        return spec.build(fused=fused)
    return spec.build()


def resolve(name: str) -> ModelSpec:
    """Model spec for a trace ``model_name``; unknown names fall back to inception3, mirroring
    ``get_model``'s default pick (core/models.py:49-52)."""
    if name in REGISTRY:
        return REGISTRY[name]
    if name in STAND_IN:
        return REGISTRY[STAND_IN[name]]
    return REGISTRY["inception3"]


def synthetic_batch(spec: ModelSpec, batch: int, pinned: bool = True, seed: int = 0):
    """Host-side synthetic batch of the named shape (uint8 NHWC images or int64 tokens)."""
    g = torch.Generator().manual_seed(seed)
    if spec.kind == "image":
        x = torch.randint(0, 256, (batch, *spec.sample_shape), dtype=torch.uint8, generator=g)
        y = torch.randint(0, spec.num_classes, (batch,), dtype=torch.int64, generator=g)
    else:
        x = torch.randint(0, spec.num_classes, (batch, spec.sample_shape[0]), dtype=torch.int64, generator=g)
        y = torch.roll(x, -1, dims=1)
    if pinned and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    return x, y


def param_bytes(model: nn.Module, bytes_per_el: int = 4) -> int:
    return sum(p.numel() for p in model.parameters()) * bytes_per_el
