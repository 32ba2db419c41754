"""End-to-end trainer checks on one GPU: the fused PS path (world = 1) against the plain-PyTorch
update path on the same model, eager against CUDA-graph replay, ring mode, checkpoint/resume."""
import pytest
import torch
from torch import nn

from tiresias_b200.models.zoo import ModelSpec, resolve, synthetic_batch
from tiresias_b200.ops import native
from tiresias_b200.ops.linear import TcLinear

pytestmark = pytest.mark.gpu


class TinyNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(3, 16, 3, 2, 1, bias=False)
        self.bn = nn.BatchNorm2d(16)
        self.fc1 = TcLinear(16 * 8 * 8, 256, act=native.ACT_RELU)
        self.fc2 = TcLinear(256, 16)

    def forward(self, x):
        x = torch.relu(self.bn(self.conv(x)))
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc2(self.fc1(x))


TINY = ModelSpec("tiny", TinyNet, "image", (16, 16, 3), 16)


def _run(mode, steps, use_graph, cuda_device, seed=0):
    from tiresias_b200.parallel.trainer import Trainer
    tr = Trainer(TINY, batch_size=32, mode=mode, use_graph=use_graph, device=cuda_device,
                 bucket_mb=0.05, lr=0.05, seed=seed)
    losses = []
    for i in range(steps):
        x, y = synthetic_batch(TINY, 32, seed=i % 2)
        tr.step(x, y)
        losses.append(tr.loss())
    assert tr.check_status() == 0
    params = tr.param_flat.float().clone()
    launches = tr.native_launches
    nb = len(tr.buckets)
    tr.close()
    return losses, params, launches, nb


def test_linear_backward_matches_torch(cuda_device):
    torch.manual_seed(0)
    lin = TcLinear(256, 136).to(cuda_device).to(torch.bfloat16)
    x = torch.randn(40, 256, device=cuda_device).to(torch.bfloat16).requires_grad_(True)
    y = lin(x)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().float().requires_grad_(True)
    wr = lin.weight.detach().float().requires_grad_(True)
    br = lin.bias.detach().float().requires_grad_(True)
    yr = torch.nn.functional.linear(xr, wr, br)
    yr.backward(g.float())
    assert torch.allclose(y.float(), yr, rtol=3e-2, atol=3e-2)
    assert torch.allclose(x.grad.float(), xr.grad, rtol=3e-2, atol=5e-2)
    assert torch.allclose(lin.weight.grad.float(), wr.grad, rtol=3e-2, atol=1e-1)
    assert torch.allclose(lin.bias.grad.float(), br.grad, rtol=3e-2, atol=1e-1)


def test_ps_matches_torch_update_path(cuda_device):
    l_ps, p_ps, launches, nb = _run("ps", 4, False, cuda_device)
    l_ref, p_ref, _, _ = _run("nccl", 4, False, cuda_device)
    assert nb >= 2
    assert launches > 0
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_ps, l_ref)), (l_ps, l_ref)
    assert torch.allclose(p_ps, p_ref, rtol=2e-2, atol=2e-2)
    assert l_ps[-1] < l_ps[0]


def test_graph_replay_matches_eager(cuda_device):
    l_e, p_e, _, _ = _run("ps", 7, False, cuda_device)
    l_g, p_g, launches, _ = _run("ps", 7, True, cuda_device)
    assert launches > 0
    assert all(abs(a - b) < 5e-2 for a, b in zip(l_e, l_g)), (l_e, l_g)
    assert torch.allclose(p_e, p_g, rtol=2e-2, atol=2e-2)


def test_ring_mode_world1(cuda_device):
    l_r, p_r, _, _ = _run("ring", 4, False, cuda_device)
    l_ps, p_ps, _, _ = _run("ps", 4, False, cuda_device)
    assert torch.allclose(p_r, p_ps, rtol=2e-2, atol=2e-2)


def test_checkpoint_resume(cuda_device):
    from tiresias_b200.parallel.trainer import Trainer
    tr = Trainer(TINY, batch_size=32, mode="ps", use_graph=False, device=cuda_device, bucket_mb=0.05)
    x, y = synthetic_batch(TINY, 32, seed=0)
    for _ in range(2):
        tr.step(x, y)
    state = tr.state_dict()
    tr.step(x, y)
    want = tr.loss()
    p_want = tr.param_flat.float().clone()
    tr.close()
    tr2 = Trainer(TINY, batch_size=32, mode="ps", use_graph=False, device=cuda_device, bucket_mb=0.05, seed=123)
    tr2.load_state_dict(state)
    tr2.step(x, y)
    got = tr2.loss()
    assert abs(got - want) < 5e-2
    assert torch.allclose(tr2.param_flat.float(), p_want, rtol=2e-2, atol=2e-2)
    tr2.close()


def test_resnet50_step(cuda_device):
    from tiresias_b200.parallel.trainer import Trainer
    spec = resolve("resnet50")
    tr = Trainer(spec, batch_size=8, mode="ps", use_graph=False, device=cuda_device)
    x, y = synthetic_batch(spec, 8)
    tr.step(x, y)
    l0 = tr.loss()
    assert l0 == l0 and 0 < l0 < 20
    assert tr.check_status() == 0
    tr.close()


@pytest.mark.parametrize("name", ["googlenet", "mobilenet_v1_025", "vgg11", "transformer"])
def test_other_model_families_step(cuda_device, name):
    """Channel counts that are not multiples of 64 / 128 (GoogLeNet's 96- and 208-wide 3x3 layers, MobileNet's
    depthwise stack) must route to the kernels that serve them, not crash in an eligibility query."""
    from tiresias_b200.parallel.trainer import Trainer
    spec = resolve(name)
    tr = Trainer(spec, batch_size=4, mode="ps", use_graph=False, device=cuda_device)
    x, y = synthetic_batch(spec, 4)
    for _ in range(2):
        tr.step(x, y)
    loss = tr.loss()
    assert loss == loss and 0 < loss < 30
    assert tr.check_status() == 0
    tr.close()


def test_profile_phases_reports_every_phase(cuda_device):
    from tiresias_b200.parallel.trainer import Trainer
    tr = Trainer(TINY, batch_size=32, mode="ps", use_graph=False, device=cuda_device, bucket_mb=0.05)
    x, y = synthetic_batch(TINY, 32, seed=0)
    tr.step(x, y)
    ph = tr.profile_phases(x, y, steps=3)
    assert set(ph) == {"h2d", "preprocess", "forward", "backward", "comm_tail", "total"}
    assert all(v >= 0 for v in ph.values()) and ph["total"] >= ph["forward"] + ph["backward"] - 1e-3
    assert tr.check_status() == 0
    tr.close()


def test_conv1x1_on_tcgen05_matches_torch(cuda_device):
    from tiresias_b200.ops.conv import TcConv1x1
    torch.manual_seed(0)
    conv = TcConv1x1(64, 256).to(cuda_device).to(torch.bfloat16).to(memory_format=torch.channels_last)
    x = torch.randn(4, 64, 14, 14, device=cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x.requires_grad_(True)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    wr = conv.weight.detach().float().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr)
    yr.backward(dy.float())
    assert y.shape == yr.shape and torch.allclose(y.float(), yr, rtol=3e-2, atol=3e-2)
    assert torch.allclose(x.grad.float(), xr.grad, rtol=3e-2, atol=5e-2)
    assert conv.weight.grad.shape == conv.weight.shape
    assert torch.allclose(conv.weight.grad.float(), wr.grad, rtol=3e-2, atol=2e-1)


def test_fit_with_prefetch_equals_step_by_step(cuda_device):
    """Trainer.fit (next batch copied on its own stream while the current step runs) must train exactly
    like step(x, y) + loss() on the same batches."""
    from tiresias_b200.parallel.trainer import Trainer
    batches = [synthetic_batch(TINY, 32, pinned=True, seed=i % 3) for i in range(7)]
    tr = Trainer(TINY, batch_size=32, mode="ps", use_graph=True, device=cuda_device, bucket_mb=0.05, lr=0.05)
    step_losses = []
    for x, y in batches:
        tr.step(x, y)
        step_losses.append(tr.loss())
    p_step = tr.param_flat.float().clone()
    tr.close()
    tr = Trainer(TINY, batch_size=32, mode="ps", use_graph=True, device=cuda_device, bucket_mb=0.05, lr=0.05)
    fit_losses = list(tr.fit(batches))
    assert tr.check_status() == 0
    p_fit = tr.param_flat.float().clone()
    with pytest.raises(RuntimeError):
        tr.step()                                   # nothing staged
    tr.close()
    assert len(fit_losses) == len(batches)
    assert fit_losses == step_losses
    assert torch.equal(p_fit, p_step)
