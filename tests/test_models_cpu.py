"""Model registry and trainer planning logic that need no GPU."""
import pytest
import torch
from torch import nn

from tiresias_b200.models import zoo
from tiresias_b200.ops.batchnorm import FusedBatchNormAct
from tiresias_b200.parallel.trainer import plan_buckets
from tiresias_b200.sim import models as sim_models


def test_every_reference_model_name_resolves_to_an_executable_model():
    assert set(zoo.MODEL_SIZES_MB) == set(sim_models.model_sizes)          # 17 names, same table
    for name in zoo.MODEL_SIZES_MB:
        spec = zoo.resolve(name)
        assert spec.kind in ("image", "tokens") and callable(spec.build)
    assert zoo.resolve("no_such_model").name == zoo.resolve("inception3").name   # get_model's default pick


def test_parameter_counts_match_the_architectures():
    counts = {n: sum(p.numel() for p in zoo.build(zoo.resolve(n)).parameters())
              for n in ("resnet50", "vgg16", "alexnet")}
    assert counts == {"resnet50": 25557032, "vgg16": 138357544, "alexnet": 61100840}
    # checkpoint MB of the reference table ~ 4 bytes per parameter
    assert abs(counts["resnet50"] * 4 / 2 ** 20 - zoo.MODEL_SIZES_MB["resnet50"]) < 2
    assert abs(counts["vgg16"] * 4 / 2 ** 20 - zoo.MODEL_SIZES_MB["vgg16"]) < 2


def test_fused_and_stock_variants_agree_on_cpu():
    torch.manual_seed(0)
    fused, stock = zoo.build(zoo.resolve("resnet50"), True), zoo.build(zoo.resolve("resnet50"), False)
    remap = {}
    for k, v in fused.state_dict().items():
        parts = k.split(".")
        if parts[-2].startswith("bn") or parts[-2].endswith("_bn"):
            parts.insert(-1, "bn")
        remap[".".join(parts)] = v
    stock.load_state_dict(remap, strict=False)
    x = torch.randn(2, 3, 64, 64)
    assert torch.allclose(fused(x), stock(x), atol=1e-4)


def test_fused_bn_module_cpu_fallback_matches_batchnorm2d():
    torch.manual_seed(1)
    mod, ref = FusedBatchNormAct(16, relu=True), nn.BatchNorm2d(16)
    x, res = torch.randn(4, 16, 5, 5), torch.randn(4, 16, 5, 5)
    assert torch.allclose(mod(x, res), torch.relu(ref(x) + res), atol=1e-5)
    assert torch.allclose(mod.running_var, ref.running_var, atol=1e-6)
    mod.to(torch.bfloat16)
    assert mod.running_mean.dtype == torch.float32 and mod.weight.dtype == torch.bfloat16


def test_bucket_plan_covers_every_parameter_once_and_shards_evenly():
    model = zoo.build(zoo.resolve("resnet50"))
    for world in (1, 2, 8):
        buckets, total = plan_buckets(model, world, bucket_elems=4 << 20, pull_min_elems=1 << 21, allow_pull=True)
        seen = set()
        for b in buckets:
            assert b.length % (8 * world) == 0 and b.offset % 64 == 0
            for p, off in b.params:
                assert id(p) not in seen and b.offset <= off and off + p.numel() <= b.offset + b.length
                seen.add(id(p))
            lo, n = b.shard(world - 1, world)
            assert lo + n == b.offset + b.length
        assert len(seen) == len(list(model.parameters())) and total >= sum(p.numel() for p in model.parameters())
        assert buckets[-1].offset + buckets[-1].length <= total


def test_conv_pass_routing_spec():
    from tiresias_b200.ops.conv import _parse_ops
    assert _parse_ops("") == {"p": set(), "k": set()}
    assert _parse_ops("1") == {"p": set("fdw"), "k": set("fdw")}
    assert _parse_ops("p:fw,k:f") == {"p": {"f", "w"}, "k": {"f"}}
    with pytest.raises(ValueError):
        _parse_ops("q:f")
    with pytest.raises(ValueError):
        _parse_ops("p:x")


def test_tc_conv_modules_fall_back_to_torch_on_cpu():
    """Off the GPU the tcgen05 convolution modules are plain convolutions with Conv2d's parameter shape."""
    from tiresias_b200.ops.conv import TcConv1x1, TcConv2d
    torch.manual_seed(0)
    x = torch.randn(2, 64, 9, 9)
    c1, c3 = TcConv1x1(64, 16), TcConv2d(64, 24, 3, 2, 1)
    assert c1.weight.shape == (16, 64, 1, 1) and c3.weight.shape == (24, 64, 3, 3)
    assert torch.allclose(c1(x), torch.nn.functional.conv2d(x, c1.weight))
    assert torch.allclose(c3(x), torch.nn.functional.conv2d(x, c3.weight, stride=2, padding=1))
    with pytest.raises(ValueError):
        TcConv2d(48, 64, 3)


def test_samplers_degrade_without_nvml():
    """On a box without NVML (this CPU container) the clock / utilisation samplers report nothing instead of
    failing, so the loopback executor and CPU runs keep working."""
    from tiresias_b200.utils.clocks import ClockSampler, UtilSampler, rejected
    u = UtilSampler(0).start().stop()
    c = ClockSampler(0).start().stop()
    if u["samples"] == 0:                       # no NVML here
        assert u["gpu_util_avg"] is None and u["gpu_util_max"] is None
    else:                                       # a GPU box: values are percentages
        assert 0 <= u["gpu_util_avg"] <= 100
    assert "reasons" in c and not rejected({"reasons": ["sw_power_cap"]}) and rejected({"reasons": ["hw_slowdown"]})


def test_zoo_architectures_carry_the_table_s_gradient_volume():
    """googlenet / inception3 / the three OpenNMT LSTMs / BERT_Chinese have their own architectures whose fp32
    parameter bytes sit within ~20 % of the reference's checkpoint-size table (model/model_factory.py:19-55):
    the quantity the PS/worker cost model is calibrated on. (mobilenet_v1_025 is the real 0.47 M-parameter
    network; the table's 15 MB is a TF checkpoint with optimizer slots.)"""
    for name, tol in (("googlenet", 0.05), ("inception3", 0.08), ("2_layers_lstm_gigaword", 0.05),
                      ("1_layer_bilstm_opennmt", 0.05), ("4_layers_brnn", 0.05), ("BERT_Chinese", 0.2)):
        spec = zoo.resolve(name)
        assert spec.name == name, "stand-in where an architecture is expected"
        with torch.device("meta"):
            model = zoo.build(spec, fused=False)
        mb = sum(p.numel() for p in model.parameters()) * 4 / 1e6
        assert abs(mb - zoo.MODEL_SIZES_MB[name]) <= tol * zoo.MODEL_SIZES_MB[name], (name, mb)
    with torch.device("meta"):
        mobile = zoo.build(zoo.resolve("mobilenet_v1_025"), fused=False)
    assert 0.4e6 < sum(p.numel() for p in mobile.parameters()) < 0.6e6
    assert zoo.resolve("inception2").name == "googlenet" and zoo.resolve("inception4").name == "inception3"


@pytest.mark.parametrize("name,hw", [("googlenet", 64), ("inception3", 96), ("mobilenet_v1_025", 64)])
def test_new_cnn_architectures_run_fused_and_stock_on_cpu(name, hw):
    torch.manual_seed(0)
    spec = zoo.resolve(name)
    for fused in (True, False):
        model = zoo.build(spec, fused=fused)
        y = model(torch.randn(2, 3, hw, hw))
        assert y.shape == (2, 1000) and torch.isfinite(y).all()


def test_seq2seq_lstm_forward_backward_on_cpu():
    from tiresias_b200.models.rnn import BertEncoder, Seq2SeqLSTM
    torch.manual_seed(0)
    for bidir in (False, True):
        model = Seq2SeqLSTM(vocab=97, emb=16, hidden=32, layers=2, bidirectional=bidir, fused=False)
        tokens = torch.randint(0, 97, (3, 9))
        logits = model(tokens)
        assert logits.shape == (3, 9, 97)
        torch.nn.functional.cross_entropy(logits.reshape(-1, 97), torch.roll(tokens, -1, 1).reshape(-1)).backward()
        assert all(p.grad is not None for p in model.parameters())
    bert = BertEncoder(vocab=101, d_model=32, n_layers=2, n_heads=4, d_ff=64, max_len=16, fused=False)
    assert bert.head.weight is bert.tok.weight                                   # tied LM head
    assert bert(torch.randint(0, 101, (2, 7))).shape == (2, 7, 101)


def test_skip_gradient_is_parked_before_the_main_branch_consumes_it():
    """ops/conv.py GradBox: the skip branch, built last in the forward, runs first in the backward, so the main
    branch's first op finds the skip gradient parked and autograd receives one (already summed) gradient."""
    from tiresias_b200.ops.conv import GradBox, park_skip_grad
    seen = []

    class Main(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, box):
            ctx.box = box
            return x * 2.0

        @staticmethod
        def backward(ctx, g):
            box = ctx.box
            skip, box.grad, box.main_done = box.grad, None, True
            seen.append(skip is not None)
            return g * 2.0 + (skip if skip is not None else 0.0), None

    leaf = torch.arange(6.0).reshape(2, 3).requires_grad_(True)
    x = leaf * 1.0
    box = GradBox()
    main = torch.tanh(Main.apply(x, box)) * 5.0            # a few more nodes on the main branch
    y = main + 3.0 * park_skip_grad(x, box)
    y.sum().backward()
    want = 5.0 * (1 - torch.tanh(2.0 * leaf.detach()) ** 2) * 2.0 + 3.0
    assert seen == [True]
    assert torch.allclose(leaf.grad, want)
    # and when the main branch runs first (skip built first), nothing is lost: autograd adds as usual
    seen.clear()
    leaf.grad = None
    x = leaf * 1.0
    box = GradBox()
    skip = 3.0 * park_skip_grad(x, box)
    y = torch.tanh(Main.apply(x, box)) * 5.0 + skip
    y.sum().backward()
    assert torch.allclose(leaf.grad, want)


def test_new_op_modules_fall_back_to_torch_on_cpu():
    """FusedLayerNorm, MaxPool, TcConv3x3Bias and fused_attention on CPU tensors are their torch counterparts (the
    native kernels serve bf16 CUDA tensors only), so model wiring can be tested without a GPU."""
    from tiresias_b200.ops.attention import attention_reference, fused_attention
    from tiresias_b200.ops.conv import TcConv3x3Bias
    from tiresias_b200.ops.layernorm import FusedLayerNorm
    from tiresias_b200.ops.pooling import MaxPool
    torch.manual_seed(0)
    x = torch.randn(3, 10, 512)
    ln, ref = FusedLayerNorm(512), torch.nn.LayerNorm(512)
    ref.load_state_dict(ln.state_dict())
    assert torch.equal(ln(x), ref(x))
    img = torch.randn(2, 8, 13, 17)
    for k, s, p in ((3, 2, 1), (2, 2, 0), (3, 2, 0)):
        assert torch.equal(MaxPool(k, s, p)(img), torch.nn.functional.max_pool2d(img, k, s, p))
    conv = TcConv3x3Bias(64, 128, relu=True)
    xin = torch.randn(1, 64, 6, 6)
    want = torch.relu(torch.nn.functional.conv2d(xin, conv.weight, conv.bias, padding=1))
    assert torch.allclose(conv(xin), want)
    assert sorted(n for n, _ in conv.named_parameters()) == ["bias", "weight"]
    qkv = torch.randn(2 * 16, 3 * 2 * 64)
    assert torch.equal(fused_attention(qkv, 2, 16, 2, True), attention_reference(qkv, 2, 16, 2, True))


def test_release_stream_caches_is_keyed_by_device_and_stream():
    """Trainer.close() drops the scratch of ITS streams only (ops/native.py::release_stream_caches)."""
    from tiresias_b200.ops import batchnorm, layernorm, native
    saved = [dict(c) for c in (batchnorm._scratch, layernorm._scratch, native._wgrad_ws)]
    try:
        batchnorm._scratch[(0, 111)] = torch.zeros(1)
        batchnorm._scratch[(0, 222)] = torch.zeros(1)
        layernorm._scratch[(0, 111, 512)] = torch.zeros(1)
        native._wgrad_ws[(1, 111)] = torch.zeros(1)

        assert native.drop_python_stream_caches(0, 111) == 2
        assert (0, 111) not in batchnorm._scratch and (0, 222) in batchnorm._scratch
        assert (0, 111, 512) not in layernorm._scratch and (1, 111) in native._wgrad_ws
    finally:
        for c, s in zip((batchnorm._scratch, layernorm._scratch, native._wgrad_ws), saved):
            c.clear()
            c.update(s)
