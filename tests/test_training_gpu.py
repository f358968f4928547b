"""GPU: the training-step API -- the CUDA-graph step replays exactly the eager step, the host feeder pipeline
delivers the right batches in order, flat-bucket optimizers keep parameters and gradients as views."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(seed=0):
    from replicann_b200.arch.resnet import ResNet
    torch.manual_seed(seed)
    return ResNet((1, 1), bottleneck=True, n_classes=16).cuda().to(torch.bfloat16)


def _batches(n, seed=3):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(8, 3, 64, 64, generator=g), torch.randint(0, 16, (8,), generator=g)) for _ in range(n)]


def test_cuda_graph_step_matches_eager_step():
    from replicann_b200.optim import SGD
    from replicann_b200.training import Trainer
    data = _batches(8)
    losses = {}
    for graph in (False, True):
        m = _model()
        tr = Trainer(m, SGD(m.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4), cuda_graph=graph)
        losses[graph] = [tr.step(x.cuda(), y.cuda()).item() for x, y in data]
        if graph:
            assert tr.graph_captured
        losses[(graph, "w")] = torch.cat([p.detach().float().flatten() for p in m.parameters()])
    # identical kernels with identical arguments: the replayed step is bitwise the eager step
    assert losses[True] == losses[False]
    assert torch.equal(losses[(True, "w")], losses[(False, "w")])
    assert all(l == l and abs(l) < 1e4 for l in losses[False])   # finite


def test_cuda_graph_accepts_adamw_and_dropout():
    """AdamW keeps its step count on the device and dropout draws from a device-side epoch, so neither rules a
    captured step out any more: the AdamW graph step must replay the eager AdamW step bit for bit."""
    from replicann_b200.arch.alexnet import AlexNet
    from replicann_b200.optim import SGD, AdamW
    from replicann_b200.training import Trainer
    data = _batches(7)
    res = {}
    for graph in (False, True):
        m = _model()
        opt = AdamW(m.parameters(), lr=1e-3, weight_decay=1e-2)
        tr = Trainer(m, opt, cuda_graph=graph)
        res[graph] = [tr.step(x.cuda(), y.cuda()).item() for x, y in data]
        res[(graph, "w")] = torch.cat([p.detach().float().flatten() for p in m.parameters()])
        assert opt.state_dict()["steps"] == len(data)
        if graph:
            assert tr.graph_captured
    assert res[True] == res[False]
    assert torch.equal(res[(True, "w")], res[(False, "w")])
    a = AlexNet().cuda().to(torch.bfloat16)  # p_dropout = 0.5, training mode: accepted (device-side RNG epoch)
    Trainer(a, SGD(a.parameters(), lr=0.1), cuda_graph=True)


@pytest.mark.parametrize("graph", [False, True])
def test_host_feeder_delivers_batches_in_order(graph):
    from replicann_b200.optim import SGD
    from replicann_b200.training import HostFeeder, Trainer
    data = [(x.pin_memory(), y.pin_memory()) for x, y in _batches(7)]
    m1, m2 = _model(), _model()
    t1 = Trainer(m1, SGD(m1.parameters(), lr=0.05, momentum=0.9))
    ref = [t1.step(x.cuda(), y.cuda()).item() for x, y in data]
    t2 = Trainer(m2, SGD(m2.parameters(), lr=0.05, momentum=0.9), cuda_graph=graph)
    feeder = HostFeeder(t2)
    got = []
    feeder.submit(*data[0])
    for i in range(len(data)):
        if i + 1 < len(data):
            feeder.submit(*data[i + 1])   # prefetch the next batch while this one trains
        got.append(feeder.step().item())
    assert got == ref
    with pytest.raises(RuntimeError):
        feeder.step()


def test_flat_buckets_are_views_and_adamw_steps():
    from replicann_b200.optim import AdamW
    m = _model()
    opt = AdamW(m.parameters(), lr=1e-3)
    for g in opt.groups:
        for p in g.params:
            assert p.data.untyped_storage().data_ptr() == g.work.untyped_storage().data_ptr()
            assert p.grad.untyped_storage().data_ptr() == g.grad.untyped_storage().data_ptr()
    x, y = _batches(1)[0]
    from replicann_b200.nn.layers import cross_entropy
    before = torch.cat([p.detach().float().flatten() for p in m.parameters()])
    for _ in range(2):
        opt.zero_grad()
        cross_entropy(m(x.cuda()), y.cuda()).backward()
        opt.step()
    after = torch.cat([p.detach().float().flatten() for p in m.parameters()])
    assert torch.isfinite(after).all() and not torch.equal(before, after)
    sd = opt.state_dict()
    opt.load_state_dict(sd)


def test_vgg_alexnet_dropout_train_mode_runs():
    """Dropout active (p = 0.5): the step runs, loss is finite, and two steps with different counters differ."""
    from replicann_b200.arch.alexnet import AlexNet
    from replicann_b200.optim import SGD
    from replicann_b200.training import Trainer
    torch.manual_seed(0)
    m = AlexNet(n_classes=16).cuda().to(torch.bfloat16)
    tr = Trainer(m, SGD(m.parameters(), lr=0.01, momentum=0.9))
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(4, 3, 224, 224, generator=g).cuda(), torch.randint(0, 16, (4,), generator=g).cuda()
    l1, l2 = tr.step(x, y).item(), tr.step(x, y).item()
    assert l1 == l1 and l2 == l2


def test_graph_step_with_dropout_draws_new_masks_every_replay():
    """A captured step starts with rnn_rng_advance(): the Philox offsets frozen in the graph are shifted by a
    device-side epoch, so consecutive replays of the SAME graph on the SAME batch must give different losses
    (different dropout masks), while an eval-mode forward stays deterministic."""
    import torch
    from replicann_b200.arch.alexnet import AlexNet
    from replicann_b200.optim import SGD
    from replicann_b200.training import Trainer
    torch.manual_seed(0)
    m = AlexNet(n_classes=16, p_dropout=0.5).cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(8, 3, 224, 224, generator=g).cuda()
    y = torch.randint(0, 16, (8,), generator=g).cuda()
    tr = Trainer(m, SGD(m.parameters(), lr=0.0), cuda_graph=True, graph_warmup=1)   # lr 0: weights stay put
    losses = [float(tr.step(x, y).item()) for _ in range(6)]
    assert tr.graph_captured
    replayed = losses[1:]
    assert len({round(v, 6) for v in replayed}) >= 4, replayed       # fresh masks on (nearly) every replay
    m.eval()
    with torch.no_grad():
        a, b = m(x), m(x)
    assert torch.equal(a, b)
