"""Data-parallel gradient bucketing over gloo (world_size 2, CPU): the host logic of unet_zoo_b200.parallel."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from unet_zoo_b200 import parallel as P


def test_plan_buckets_reverse_order_and_sizes():
    b = P.plan_buckets([10, 20, 30, 40], bucket_bytes=200)
    assert b == [[3, 2], [1, 0]]
    assert P.plan_buckets([5], 1 << 20) == [[0]]
    assert sorted(i for g in P.plan_buckets([7] * 9, 60) for i in g) == list(range(9))


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(100 + rank)                       # different weights per rank before the wrap
    net = nn.Sequential(nn.Linear(8, 16), nn.BatchNorm1d(16), nn.Linear(16, 4))
    dp = P.DataParallel(net, bucket_mb=0.0002)          # ~200-byte buckets -> several buckets
    ref = [torch.zeros_like(p) for p in net.parameters()]
    for p, r in zip(net.parameters(), ref):
        r.copy_(p.detach())
    gathered = [torch.zeros_like(ref[0]) for _ in range(world)]
    dist.all_gather(gathered, ref[0])
    same_weights = all(torch.equal(g, gathered[0]) for g in gathered)          # rank 0's parameters everywhere
    params = list(net.parameters())
    pg = net._uz_make_param_grads(params, torch.device("cpu"))
    nb = len(pg.buckets)
    for i, p in enumerate(params):
        pg.view(p).fill_(float(rank + 1) * (i + 1))      # rank-dependent "gradients"
    for p in reversed(params):                           # backward completes parameters in reverse order
        pg.ready(p)
    pg.finish()
    ok = all(torch.allclose(pg.view(p), torch.full_like(p, 1.5 * (i + 1))) for i, p in enumerate(params))
    sd_keys = list(dp.state_dict())
    # defer_sync mode (CUDA-graph step at N > 1): backward fills a plain flat buffer, one all-reduce afterwards
    refused = False
    net._last_param_grads = pg
    try:
        dp.sync_gradients()                              # the bucketed backward above already averaged
    except RuntimeError:
        refused = True
    dp.defer_sync = True
    pg2 = net._uz_make_param_grads(params, torch.device("cpu"))
    for i, p in enumerate(params):
        pg2.view(p).fill_(float(rank + 1) * (i + 2))
        pg2.ready(p)                                     # no hook: nothing is launched
    net._last_param_grads = pg2                          # what _HotPathFn.backward records
    dp.sync_gradients()
    ok2 = (type(pg2) is P.ParamGrads and refused and
           all(torch.allclose(pg2.view(p), torch.full_like(p, 1.5 * (i + 2))) for i, p in enumerate(params)))
    out[rank] = (same_weights, ok and ok2, nb, all(pg.launched), sd_keys == list(net.state_dict()))
    dist.destroy_process_group()


def test_bucketed_allreduce_world2():
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    for r in range(2):
        same_weights, averaged, nb, launched, keys = out[r]
        assert same_weights and averaged and launched and keys
        assert nb >= 2


def _padded_worker(rank, world, port, out):
    """multiresunet's zero-padded stand-ins under data parallelism: the padded gradients are gathered into the flat
    buffer at the end of backward, and only then do the buckets see their parameters complete and all-reduce."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import unet_zoo_b200
    from unet_zoo_b200.models._padded import Shadow
    torch.manual_seed(7)
    model = unet_zoo_b200.create_model("multiresunet", filters=8)
    dp = P.DataParallel(model, bucket_mb=0.05)
    params = list(model.parameters())
    sh = Shadow(model, model._specs(), torch.device("cpu"))
    sh.refresh(True)
    pg = model._uz_make_param_grads(params, torch.device("cpu"))
    sh.begin_backward(pg)
    early = 0
    for (t, _off) in sh.tensors:
        pg.view(t).copy_(t.detach() * float(rank + 1))   # rank-dependent "gradient" of every padded operand
        pg.ready(t)                                      # reported by the tape: must not complete any bucket
        early += sum(pg.launched)
    sh.end_backward(pg, params)                          # gather + report every parameter: buckets launch now
    launched = all(pg.launched)
    pg.finish()
    ok = all(torch.allclose(pg.view(p), p.detach() * 1.5) for p in params)
    out[rank] = (early == 0, launched, ok, len(pg.buckets), type(dp.module).__name__)
    dist.destroy_process_group()


def test_padded_operands_with_bucketed_allreduce_world2():
    mgr = mp.Manager()
    out = mgr.dict()
    port = 31500 + os.getpid() % 2000
    mp.spawn(_padded_worker, args=(2, port, out), nprocs=2, join=True)
    for r in range(2):
        nothing_early, launched, averaged, nb, name = out[r]
        assert nothing_early and launched and averaged and nb >= 2 and name == "MultiResUnet"
