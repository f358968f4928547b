"""CPU (gloo, world_size 2): host-side logic of the data-parallel path -- bucket planning, ready-hooks that
launch a bucket's all-reduce as soon as its last gradient is produced, and the summed result."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_plan_buckets():
    from replicann_b200.parallel import plan_buckets
    assert plan_buckets([10, 10, 10, 10], 2, 40) == [(0, 2), (2, 4)]
    assert plan_buckets([100], 2, 40) == [(0, 1)]
    assert plan_buckets([1, 1, 1], 4, 1 << 20) == [(0, 3)]
    assert plan_buckets([], 2, 8) == []
    # the last bucket is split so that only a small tail is exchanged after backward has ended
    assert plan_buckets([10, 10, 10, 10, 3, 2], 2, 1 << 20, tail_bytes=12) == [(0, 4), (4, 6)]
    assert plan_buckets([10, 10, 10], 2, 40, tail_bytes=4) == [(0, 2), (2, 3)]      # nothing small enough to split off
    assert plan_buckets([5], 2, 40, tail_bytes=4) == [(0, 1)]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from replicann_b200.nn import _functions as Fn
        from replicann_b200.optim import SGD
        from replicann_b200.parallel import DataParallel
        torch.manual_seed(rank)  # different initial weights per rank: broadcast must make them equal
        model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Linear(32, 8), torch.nn.Linear(8, 4))
        opt = SGD(model.parameters(), lr=0.1, momentum=0.9)
        dp = DataParallel(model, opt, bucket_bytes=1024)  # small buckets -> several of them
        assert dp.n_buckets >= 2
        w0 = torch.cat([g.work.clone() for g in opt.groups])
        gathered = [torch.zeros_like(w0) for _ in range(world)]
        dist.all_gather(gathered, w0)
        assert torch.equal(gathered[0], gathered[1]), "parameters were not broadcast from rank 0"
        assert opt.grad_scale == 1.0 / world
        params = list(model.parameters())[::-1]
        shared = params[0]   # reports two gradient kernels per step (a module applied twice in one forward)
        # first step: the number of gradient kernels per parameter is learned, finish() launches everything
        dp.begin_step()
        for p in params:
            Fn.grad_target(p)[0].zero_()
            Fn._notify(p)
        Fn._notify(shared)
        assert not any(b.launched for b in dp._buckets)
        dp.finish()
        assert all(b.launched for b in dp._buckets)
        opt.zero_grad()
        # simulate backward: gradients appear in reverse registration order, each reported through the hook
        dp.begin_step()
        launched_before_finish = 0
        for i, p in enumerate(params):
            tgt, acc = Fn.grad_target(p)
            assert acc is False
            tgt.copy_(torch.full_like(p, float(rank + 1) * (i + 1)))
            Fn._notify(p)
            if p is shared:
                # the first of its two kernels must not release the bucket: the second accumulation is to come
                assert not dp._param_bucket[id(p)].launched
                Fn._notify(p)
            launched_before_finish = sum(b.launched for b in dp._buckets)
        assert launched_before_finish == dp.n_buckets  # every bucket went out during "backward"
        dp.finish()
        for i, p in enumerate(params):
            expect = float(sum(r + 1 for r in range(world)) * (i + 1))
            assert torch.allclose(p.grad, torch.full_like(p, expect)), (i, p.grad.flatten()[:3], expect)
        # a step in which one parameter never gets a gradient: finish() still reduces its bucket
        opt.zero_grad()
        dp.begin_step()
        for i, p in enumerate(params[:-1]):
            tgt, _ = Fn.grad_target(p)
            tgt.fill_(1.0)
            Fn._notify(p)
            if p is shared:
                Fn._notify(p)
        assert sum(b.launched for b in dp._buckets) < dp.n_buckets
        dp.finish()
        assert all(b.launched for b in dp._buckets)
        dp.close()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_data_parallel_buckets_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=100) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}:\n{msg}"
