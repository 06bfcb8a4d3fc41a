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


def _tiny_bert(dropout=0.0):
    from fedml_b200.models.nlp import DistilBertForSequenceClassification
    return DistilBertForSequenceClassification(num_labels=5, seq_classif_dropout=dropout, vocab_size=97, dim=32, n_layers=2,
                                               n_heads=4, hidden_dim=64, max_pos=24, dropout=dropout,
                                               attention_dropout=dropout)


def test_batched_distilbert_matches_autograd():
    """Explicit client-batched DistilBERT forward/backward == autograd of the reference-named module (padding tokens,
    key mask, ignored rows included)."""
    C, B, S = 3, 4, 16
    torch.manual_seed(0)
    models = [_tiny_bert() for _ in range(C)]
    for m in models:      # non-trivial LayerNorm affine parameters and biases
        for n, p in m.named_parameters():
            if "LayerNorm" in n or "layer_norm" in n or n.endswith(".bias"):
                p.data.add_(0.1 * torch.randn_like(p))
    layout = ArenaLayout(models[0])
    stack = torch.stack([layout.flatten(m.state_dict()) for m in models])
    W, G = layout.stacked_views(stack), layout.stacked_views(torch.zeros_like(stack))
    bm = make_batched(models[0], TorchBackend(torch.float32))
    assert bm is not None and bm.input_kind == "token_ids"
    ids = torch.randint(1, 97, (C, B, S))
    ids[:, :, 12:] = 0                     # padding tail -> key mask
    ids[0, 1, 5:] = 0
    y = torch.randint(0, 5, (C, B))
    y[1, 3] = -1                            # padded row of a ragged last batch
    scale = 1.0 / (y >= 0).sum(dim=1).float()
    loss = bm.train_step(ids, y, W, G, torch.ones(C), row_scale=scale)
    for c, m in enumerate(models):
        m.train()
        out = m(ids[c])
        l = F.cross_entropy(out, y[c], ignore_index=-1)
        l.backward()
        assert abs(float(l) - float(loss[c])) < 1e-4
        for name, p in m.named_parameters():
            torch.testing.assert_close(G[name][c], p.grad, rtol=2e-3, atol=2e-5, msg=lambda s, n=name: f"{n}: {s}")
    lg = bm.logits(ids, W)
    for c, m in enumerate(models):
        m.eval()
        torch.testing.assert_close(lg[c], m(ids[c]), rtol=1e-3, atol=1e-4)


def test_batched_distilbert_dropout_masks_are_consistent():
    """With dropout on, the explicit backward uses the same masks as the forward: a finite-difference probe of the
    loss along the analytic gradient direction matches (the step counter is frozen for the probe)."""
    C, B, S = 2, 3, 8
    torch.manual_seed(3)
    model = _tiny_bert(dropout=0.25)
    layout = ArenaLayout(model)
    stack = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)])
    gstack = torch.zeros_like(stack)
    bm = make_batched(model, TorchBackend(torch.float32))
    ids = torch.randint(1, 97, (C, B, S)); y = torch.randint(0, 5, (C, B))
    l0 = bm.train_step(ids, y, layout.stacked_views(stack), layout.stacked_views(gstack), torch.ones(C))
    step = int(bm._step_dev.item())
    direction = gstack / gstack.norm()
    eps = 2e-2

    def probe(sign):
        bm._step_dev.fill_(step - 1)          # same masks as the analytic step
        return bm.train_step(ids, y, layout.stacked_views(stack + sign * eps * direction),
                             layout.stacked_views(torch.zeros_like(gstack)), torch.ones(C)).double().sum()

    fd = float(probe(+1) - probe(-1)) / (2 * eps)
    an = float((gstack.double() * direction.double()).sum())
    assert an > 0 and abs(fd - an) < 0.03 * an, (fd, an)
    assert l0.shape == (C,)


def test_ragged_batches_with_batchnorm_train_sequentially():
    """a BatchNorm model whose clients have n % batch_size != 0 (no drop_last) must not take the client-batched
    path (padding rows would enter the batch statistics): the round trains client by client and equals the per-client
    PyTorch loop."""
    import copy
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.simulation import SimulatorNCCL
    from fedml_b200.simulation.trainers import create_model_trainer
    a = fedml.init(Arguments.from_dict(dict(
        training_type="simulation", backend="NCCL", dataset="cifar10", model="resnet56", client_num_in_total=4,
        client_num_per_round=2, comm_round=1, epochs=1, batch_size=8, learning_rate=0.05, partition_method="homo",
        frequency_of_the_test=0, synthetic_train_size=4 * 13, log_level="ERROR", using_gpu=False, random_seed=0,
        shuffle_clients=False, drop_last=False)))             # no drop_last: 13 samples = one full batch + 5 ragged
    ds, ncls = fedml.data.load(a)
    torch.manual_seed(0)
    model = fedml.models.create(a, ncls)
    sim = SimulatorNCCL(a, torch.device("cpu"), ds, copy.deepcopy(model)).simulator
    assert sim.batched is not None and sim.batched.uses_batch_statistics
    init = copy.deepcopy(model.state_dict())
    sim.train_one_round(0)
    assert sim._ragged_round and sim._warned_ragged
    # oracle: the same two clients trained one after another with the stock trainer, sample-weighted average
    from fedml_b200.simulation.sp.fedavg_api import client_sampling
    picked = client_sampling(0, 4, 2)
    acc, tot = None, 0
    for cid in picked:
        m = copy.deepcopy(model); m.load_state_dict(init)
        tr = create_model_trainer(m, a); tr.set_id(cid)
        tr.train(ds[5][cid], torch.device("cpu"), a)
        n = ds[4][cid]; tot += n
        sd = {k: v.double() * n for k, v in m.state_dict().items()}
        acc = sd if acc is None else {k: acc[k] + sd[k] for k in acc}
    got = sim.global_state_dict()
    for k, v in acc.items():
        torch.testing.assert_close(got[k].double(), v / tot, rtol=1e-4, atol=1e-5, msg=k)


def test_batched_groupnorm_resnet_matches_autograd():
    """ResNet with GroupNorm (two channels per group, the reference's resnet18_gn normalisation: model/cv/resnet_gn.py,
    group_normalization.py) trains client-batched: explicit forward / backward == per-client autograd."""
    C, B = 3, 5
    mk = lambda: ResNet(BasicBlock, [1, 1, 1], num_classes=10, stem="cifar", group_norm=2)   # noqa: E731
    models, layout, stack = _stack_clients(mk, C)
    assert not any("running" in k for k in models[0].state_dict())
    gstack = torch.zeros_like(stack)
    W, G = layout.stacked_views(stack), layout.stacked_views(gstack)
    bm = make_batched(models[0], TorchBackend(torch.float32))
    assert bm is not None and not bm.uses_batch_statistics
    torch.manual_seed(1)
    x_u8 = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8)
    y = torch.randint(0, 10, (C, B))
    loss = bm.train_step(x_u8, y, W, G, torch.ones(C))
    mean = torch.tensor((0.4914, 0.4822, 0.4465)).view(1, 3, 1, 1)
    std = torch.tensor((0.2470, 0.2435, 0.2616)).view(1, 3, 1, 1)
    for c, m in enumerate(models):
        m.train()
        xin = (x_u8[c].permute(0, 3, 1, 2).float() / 255.0 - mean) / std
        l = F.cross_entropy(m(xin), y[c])
        l.backward()
        assert abs(float(l) - float(loss[c])) < 1e-4
        for name, p in m.named_parameters():
            torch.testing.assert_close(G[name][c], p.grad, rtol=2e-3, atol=2e-5, msg=lambda s, n=name: f"{n}: {s}")
    lg = bm.logits(x_u8, W)
    for c, m in enumerate(models):
        m.eval()
        xin = (x_u8[c].permute(0, 3, 1, 2).float() / 255.0 - mean) / std
        torch.testing.assert_close(lg[c], m(xin), rtol=1e-3, atol=1e-4)


def _cnn(p1=0.0, p2=0.0, only_digits=False):
    from fedml_b200.models.cv import CNN_DropOut
    m = CNN_DropOut(only_digits=only_digits)
    m.dropout_1.p, m.dropout_2.p = p1, p2
    return m


def test_batched_cnn_dropout_matches_autograd():
    """FEMNIST CNN (reference model/cv/cnn.py:74-142): valid convolutions as centre crops of same-padded ones, pooling with
    a window offset, (h, w, c)-ordered fc staging - gradients equal per-client autograd when dropout is off."""
    C, B = 3, 5
    models, layout, stack = _stack_clients(lambda: _cnn(), C, seed=5)
    gstack = torch.zeros_like(stack)
    bm = make_batched(models[0], TorchBackend(torch.float32))
    assert type(bm).__name__ == "BatchedCNNDropOut"
    x = torch.randn(C, B, 784); y = torch.randint(0, 62, (C, B)); y[1, -1] = -1          # one padded row
    W, G = layout.stacked_views(stack), layout.stacked_views(gstack)
    loss = bm.train_step(x, y, W, G, torch.ones(C))
    for c, m in enumerate(models):
        m.train()
        keep = y[c] >= 0
        l = F.cross_entropy(m(x[c][keep]), y[c][keep], reduction="sum") / B
        m.zero_grad(); l.backward()
        assert abs(float(loss[c]) - float(l)) < 1e-5
        for k, p in m.named_parameters():
            torch.testing.assert_close(G[k][c], p.grad, rtol=2e-4, atol=2e-6, msg=k)
    with torch.no_grad():
        torch.testing.assert_close(bm.logits(x, W)[0], models[0].eval()(x[0]), rtol=1e-4, atol=1e-5)


def test_batched_cnn_dropout_masks_are_consistent():
    """Dropout on: the backward pass re-derives the forward's masks (finite-difference probe along the gradient)."""
    C, B = 2, 4
    models, layout, stack = _stack_clients(lambda: _cnn(0.25, 0.5, True), C, seed=6)
    gstack = torch.zeros_like(stack)
    bm = make_batched(models[0], TorchBackend(torch.float32))
    x = torch.randn(C, B, 28, 28); y = torch.randint(0, 10, (C, B))
    l0 = bm.train_step(x, y, layout.stacked_views(stack), layout.stacked_views(gstack), torch.ones(C))
    step = int(bm._step_dev.item())
    direction = gstack / gstack.norm()
    eps = 1e-2

    def probe(sign):
        bm._step_dev.fill_(step - 1)
        return bm.train_step(x, y, layout.stacked_views(stack + sign * eps * direction),
                             layout.stacked_views(torch.zeros_like(gstack)), torch.ones(C)).double().sum()

    fd = float(probe(+1) - probe(-1)) / (2 * eps)
    an = float((gstack.double() * direction.double()).sum())
    assert an > 0 and abs(fd - an) < 0.05 * an, (fd, an)
    l1 = bm.train_step(x, y, layout.stacked_views(stack), layout.stacked_views(gstack), torch.ones(C))
    assert not torch.equal(l0, l1)                     # next step, new masks
