"""Model cost tables used by the simulator (reference orphans O1-O3: model/model_factory.py:4-106,
core/models.py:8-72).

* ``model_sizes``       checkpoint size in MB per model name = gradient bytes per iteration
                        (README.md:5 "use checkpoint size as the gradient size").
* ``estimate_gpu_utilization``  the memory-utilisation heuristic (CNN: 525 MB CUDA context + 3x
                        weights + activations; non-CNN: 0.52).
* ``get_model`` / ``get_model_with_scale``  per-parameter-tensor sizes in MB for nine CNNs — the
                        message sizes of the PS push/pull buckets.
* ``model_for_memory``  maps a live-schema trace row (which has no model name) to a model by its
                        peak GPU memory.

The tensor tables are built from the layer structure instead of being listed element by element:
VGG = conv stacks + three FC tensors (392 / 64 / 15.6 MB), ResNet = stem + N bottleneck tensors +
the stage-4/FC tail, etc.; ``tests/test_sim_models.py`` pins lengths and totals.
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

from . import util

model_sizes: Dict[str, float] = {
    "4_layers_brnn": 1300, "transformer": 1100, "1_layer_bilstm_opennmt": 900, "BERT_Chinese": 350,
    "2_layers_lstm_gigaword": 330, "mobilenet_v1_025": 15, "googlenet": 26, "inception2": 43,
    "inception3": 104, "inception4": 176, "alexnet": 233, "vgg11": 519, "vgg19": 549, "vgg16": 528,
    "resnet50": 97, "resnet101": 555, "resnet152": 737,
}
cnn_models: List[str] = ["mobilenet_v1_025", "googlenet", "inception2", "inception3", "inception4",
                         "alexnet", "vgg11", "vgg19", "vgg16", "resnet50", "resnet101", "resnet152"]
_CUDA_INIT_SIZE = 525

worker_mem = 5            # GB, core/models.py:25-27
ps_mem = 8
per_worker_mem = 0.2


def get_model_size(name: str) -> float:
    """Checkpoint MB for ``name``; unknown names price as inception3 (``get_model``'s fallback)."""
    return model_sizes.get(name, model_sizes["inception3"])


def is_cnn(name: str) -> bool:
    return name in cnn_models


def estimate_gpu_utilization(size, is_cnn, gpu_mem_cap, batch_size=64, per_data_size=0.012,
                             add_noise=True, rng=None):
    """Fraction of GPU memory a training job of ``size`` MB occupies (gpu_mem_cap in GB)."""
    if is_cnn:
        weights_and_backprop = _CUDA_INIT_SIZE + (size * 3)
        forward = 3.5 * (per_data_size * batch_size)          # AlexNet activation heuristic
        backward = forward * 2
        at_least = (weights_and_backprop + forward + backward) / ((gpu_mem_cap - 0.07) * 1024)
    else:
        at_least = 0.52
    if add_noise:
        mean, std = 0.52, 0.26
        draw = (rng if rng is not None else np.random).normal(mean, std)
        return max(round(draw, 2), round(at_least + std, 2))
    return round(float(at_least), 2)


# ---- per-tensor tables --------------------------------------------------------------------------------
_FC_VGG = [392.0, 64.0, 15.6]
_RES_TAIL = [8.0, 2.0, 9.0, 4.0, 4.0, 9.0, 4.0, 4.0, 9.0, 4.0, 7.8]
_INC4_A = [1.5, 1.1, 1.5, 1.1, 1.3, 1.5]                 # repeated inception-A style group
_INC4_C = [2.3, 1.1, 1.1, 2.3, 2.0, 2.6, 1.5, 1.5, 1.5, 1.5]


def _vgg(n64_128: int, n256: int) -> List[float]:
    return [1.1] + [2.3] * n64_128 + [4.5] + [9.0] * n256 + _FC_VGG


def _resnet(n_small: int) -> List[float]:
    return [2.0] + [2.3] * n_small + _RES_TAIL


def _inception4() -> List[float]:
    t = [1.3, 5.1, 1.5, 2.0] + _INC4_A * 7
    t = t[:-1] + [1.5, 1.3, 1.8, 2.2, 3.5, 1.5, 1.5]
    t += _INC4_C * 3
    return t[:-2] + [5.9]


m_names = ["vgg19", "vgg16", "vgg11", "alexnet", "resnet152", "resnet101", "resnet50", "inception4",
           "inception3"]
m_tensors: List[List[float]] = [
    _vgg(3, 7), _vgg(2, 5), _vgg(1, 3),
    [1.2, 2.5, 5.1, 3.4, 144.0, 64.0, 15.6],
    _resnet(36), _resnet(23), _resnet(6),
    _inception4(),
    [3.8, 2.1, 1.3, 1.6, 1.9, 1.7, 1.7, 2.2, 5.9, 1.7, 1.7, 2.5, 3.0, 1.7, 1.7, 3.5, 5.9, 1.7, 1.7, 1.5, 7.8],
]
m_mem = [1] * len(m_names)
# The reference keeps a measured per-model GPU share next to the live table, commented out (core/models.py:21);
# with the live all-ones table a job takes whole GPUs and `dlas-gpu-pack` degenerates to `dlas-gpu`.
# ``--pack_mem_util table`` (set_mem_table) switches the packing schedulers to it.
m_mem_measured = [0.60, 0.55, 0.45, 0.13, 0.85, 0.70, 0.50, 0.85, 0.80]
_use_measured_mem = False


def set_mem_table(measured: bool) -> None:
    global _use_measured_mem
    _use_measured_mem = bool(measured)


def get_model(model_name: str) -> dict:
    """{name, ind, tensors, mem_util}; unknown names resolve to inception3."""
    if model_name in m_names:
        idx = m_names.index(model_name)
    else:
        idx = m_names.index("inception3")
        util.print_fn("No model match, pick %s" % m_names[idx])
    return {"name": m_names[idx], "ind": idx, "tensors": m_tensors[idx],
            "mem_util": (m_mem_measured if _use_measured_mem else m_mem)[idx]}


def get_model_with_scale(model_name: str, model_scale) -> dict:
    ret = get_model(model_name)
    ret["tensors"] = ret["tensors"] * int(model_scale)
    total = 0.0
    for t in ret["tensors"]:
        total += t
    ret["total_size"] = round(total, 1)
    return ret


# ---- live-schema rows -> model --------------------------------------------------------------------------
_SIZE_CLASSES = [                        # (peak GPU MiB upper bound, model)
    (2 * 1024, "resnet50"), (4 * 1024, "inception3"), (6 * 1024, "alexnet"), (8 * 1024, "vgg16"),
    (12 * 1024, "resnet152"),
]


def model_for_memory(peak_mib: float) -> str:
    """Pick an executable model for a trace row from its peak GPU memory."""
    for bound, name in _SIZE_CLASSES:
        if peak_mib <= bound:
            return name
    return "transformer"
