"""Client-batched explicit forward/backward == per-client autograd of the reference-named nn.Module (CPU, fp32)."""
import copy

import pytest
import torch
import torch.nn.functional as F

from fedml_b200.models import LogisticRegression
from fedml_b200.models.batched import TorchBackend, make_batched
from fedml_b200.models.cv import BasicBlock, Bottleneck, ResNet
from fedml_b200.parallel.arena import ArenaLayout


def _stack_clients(make_model, C, seed=0):
    torch.manual_seed(seed)
    models = [make_model() for _ in range(C)]
    for m in models:   # make BN affine params non-trivial
        for n, p in m.named_parameters():
            if "bn" in n or "downsample.1" in n:
                p.data.add_(0.1 * torch.randn_like(p))
    layout = ArenaLayout(models[0])
    stack = torch.stack([layout.flatten(m.state_dict()) for m in models])
    return models, layout, stack


@pytest.mark.parametrize("block,layers", [(Bottleneck, [1, 1, 1]), (BasicBlock, [1, 1, 1])])
def test_batched_resnet_matches_autograd(block, layers):
    C, B = 3, 4
    mk = lambda: ResNet(block, layers, num_classes=10, stem="cifar")
    models, layout, stack = _stack_clients(mk, C)
    gstack = torch.zeros_like(stack)
    W, G = layout.stacked_views(stack), layout.stacked_views(gstack)
    bm = make_batched(models[0], TorchBackend(torch.float32))
    torch.manual_seed(1)
    x_u8 = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8)
    y = torch.randint(0, 10, (C, B))
    active = torch.ones(C)
    loss = bm.train_step(x_u8, y, W, G, active)
    mean = torch.tensor((0.4914, 0.4822, 0.4465)).view(1, 3, 1, 1)
    std = torch.tensor((0.2470, 0.2435, 0.2616)).view(1, 3, 1, 1)
    for c, m in enumerate(models):
        m.train()
        xin = (x_u8[c].permute(0, 3, 1, 2).float() / 255.0 - mean) / std
        out = m(xin)
        l = F.cross_entropy(out, y[c])
        l.backward()
        assert abs(float(l) - float(loss[c])) < 1e-4
        for name, p in m.named_parameters():
            torch.testing.assert_close(G[name][c], p.grad, rtol=2e-3, atol=2e-5, msg=lambda s, n=name: f"{n}: {s}")
        sd = m.state_dict()
        for name in sd:
            if "running" in name or "num_batches" in name:
                torch.testing.assert_close(W[name][c], sd[name].float(), rtol=1e-4, atol=1e-5)
    # eval-mode logits
    lg = bm.logits(x_u8, W)
    for c, m in enumerate(models):
        m.eval()
        xin = (x_u8[c].permute(0, 3, 1, 2).float() / 255.0 - mean) / std
        torch.testing.assert_close(lg[c], m(xin), rtol=1e-3, atol=1e-4)


def test_batched_resnet_inactive_client_keeps_bn_stats():
    C, B = 2, 4
    mk = lambda: ResNet(Bottleneck, [1, 1, 1], num_classes=10, stem="cifar")
    models, layout, stack = _stack_clients(mk, C)
    before = stack.clone()
    W, G = layout.stacked_views(stack), layout.stacked_views(torch.zeros_like(stack))
    bm = make_batched(models[0], TorchBackend(torch.float32))
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8)
    bm.train_step(x, torch.randint(0, 10, (C, B)), W, G, torch.tensor([1.0, 0.0]))
    assert torch.equal(stack[1], before[1])           # frozen client: buffers untouched (params are never touched here)
    assert not torch.equal(stack[0], before[0])


def test_batched_lr_matches_autograd():
    C, B = 4, 6
    models, layout, stack = _stack_clients(lambda: LogisticRegression(784, 10), C)
    W, G = layout.stacked_views(stack), layout.stacked_views(torch.zeros_like(stack))
    bm = make_batched(models[0], TorchBackend(torch.float32))
    x = torch.rand(C, B, 784); y = torch.randint(0, 10, (C, B))
    loss = bm.train_step(x, y, W, G, torch.ones(C))
    for c, m in enumerate(models):
        l = F.cross_entropy(m(x[c]), y[c]); l.backward()
        assert abs(float(l) - float(loss[c])) < 1e-5
        torch.testing.assert_close(G["linear.weight"][c], m.linear.weight.grad, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(G["linear.bias"][c], m.linear.bias.grad, rtol=1e-4, atol=1e-6)
