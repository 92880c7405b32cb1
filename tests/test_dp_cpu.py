"""Data-parallel logic on CPU (gloo, world_size 2): R replicas with sync-BatchNorm and averaged gradients must take
exactly the same optimisation step as one process on the concatenated batch (SURVEY §4 test plan item 2)."""
import os
import socket

import torch
import torch.distributed as dist
from dcnn_b200.utils.spawn import spawn_bounded

from dcnn_b200 import nn as dnn
from dcnn_b200.train import TrainStep


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model():
    m = (dnn.SequentialBuilder("dp_test").input([3, 8, 8])
         .conv2d(8, 3, 3, 1, 1, 1, 1, True, "c1").batchnorm(1e-5, 0.1, True, "bn1").activation("relu", "r1")
         .basic_residual_block(8, 16, 2, "rb")
         .avgpool2d(4, 4, 1, 1, 0, 0, "ap").flatten("f").dense(5, True, "fc").build())
    m.set_seed(3)
    m.initialize()
    return m


def _data():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(8, 3, 8, 8, generator=g)
    t = torch.nn.functional.one_hot(torch.randint(0, 5, (8,), generator=g), 5).float().reshape(8, 5, 1, 1)
    return x, t


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dcnn_b200.parallel.dp import DataParallel
    torch.set_num_threads(1)
    m = _model()
    opt = dnn.SGD(0.05, 0.9)  # SGD: linear in the gradients (Adam would amplify the rounding noise of the analytically-zero conv-bias-before-BN gradients)
    dp = DataParallel(m, opt, dist.group.WORLD, sync_bn=True)
    step = TrainStep(m, opt, dnn.SoftmaxCrossEntropyLoss(), data_parallel=dp)
    x, t = _data()
    lo, hi = rank * 4, rank * 4 + 4
    for _ in range(2):
        step(x[lo:hi], t[lo:hi])
    if rank == 0:
        torch.save([p.clone() for p in m.parameters()] + list(m.buffers().values()), out)
    dist.destroy_process_group()


def test_dp2_syncbn_equals_single_process(tmp_path):
    out = str(tmp_path / "dp.pt")
    spawn_bounded(_worker, args=(2, _free_port(), out), nprocs=2, timeout_s=300)
    got = torch.load(out)
    m = _model()
    opt = dnn.SGD(0.05, 0.9)  # SGD: linear in the gradients (Adam would amplify the rounding noise of the analytically-zero conv-bias-before-BN gradients)
    opt.attach(m.parameters(), m.gradients())
    step = TrainStep(m, opt, dnn.SoftmaxCrossEntropyLoss())
    x, t = _data()
    for _ in range(2):
        step(x, t)
    ref = [p for p in m.parameters()] + list(m.buffers().values())
    assert len(got) == len(ref)
    for a, b in zip(got, ref):
        assert torch.allclose(a, b, atol=2e-5, rtol=1e-4), (a - b).abs().max()


def test_bucket_plan_covers_the_buffer_and_shards_evenly():
    """plan_buckets (the DCNN_DP_BUCKETS planner): buckets are contiguous, start at unit boundaries, cover [0, n), and the
    per-rank shards of every bucket partition it in multiples of 4 with consistent optimizer-state offsets."""
    from dcnn_b200.models import create_resnet9_cifar10
    from dcnn_b200.parallel.dp import plan_buckets
    m = create_resnet9_cifar10()
    m.set_seed(0)
    m.initialize()
    numels = [p.numel() for p in m.parameters()]
    offsets, total = [], 0
    for nn_ in numels:
        offsets.append(total)
        total += (nn_ + 3) // 4 * 4
    units = m._get_units()
    unit_params, i = [], 0
    for u in units:
        c = sum(len(l.parameters()) for l in u.layers)
        unit_params.append((i, i + c))
        i += c
    assert i == len(numels)
    for world in (2, 3, 8):
        n = (total + 4 * world - 1) // (4 * world) * (4 * world)
        for k in (1, 2, 4, 7, 100):
            per_rank = [plan_buckets(unit_params, numels, offsets, n, k, r, world) for r in range(world)]
            b0 = per_rank[0]
            assert 1 <= len(b0) <= max(1, k)
            assert b0[0]["a"] == 0 and b0[-1]["a"] + b0[-1]["n"] == n
            assert all(x["a"] + x["n"] == y["a"] for x, y in zip(b0, b0[1:]))
            assert [b["first_unit"] for b in b0] == sorted({b["first_unit"] for b in b0})
            for bi, bk in enumerate(b0):
                assert bk["n"] % 4 == 0
                if bi > 0:       # a bucket starts where the first parameter of its first unit starts
                    lo = next(unit_params[u][0] for u in range(bk["first_unit"], len(units))
                              if unit_params[u][1] > unit_params[u][0])
                    assert bk["a"] == offsets[lo]
                cuts = [per_rank[r][bi]["sb"] for r in range(world)] + [per_rank[-1][bi]["se"]]
                assert cuts[0] == 0 and cuts[-1] == bk["n"] and cuts == sorted(cuts) and all(c % 4 == 0 for c in cuts)
                assert all(per_rank[r][bi]["se"] == per_rank[r + 1][bi]["sb"] for r in range(world - 1))
            for r in range(world):
                off = 0
                for bk in per_rank[r]:
                    assert bk["state_off"] == off
                    off += bk["se"] - bk["sb"]


def test_state_layout_partitions_the_flat_parameter_space():
    """The sharded optimizer state (one shard per rank, or one per bucket and rank) covers every flat index exactly once
    and its offsets inside a rank's state arrays are dense -- what checkpoint assembly relies on."""
    from dcnn_b200.parallel.dp import state_layout
    for world in (2, 3, 8):
        n = 4 * world * 37
        buckets = [{"a": 0, "n": 4 * 11}, {"a": 44, "n": 4 * 50}, {"a": 244, "n": n - 244}]
        for bk in (None, buckets):
            seen = [0] * n
            for r in range(world):
                off = 0
                for b, e, o in state_layout(n, world, r, bk):
                    assert o == off and b % 4 == 0 and e % 4 == 0 and b <= e
                    off += e - b
                    for i in range(b, e):
                        seen[i] += 1
            assert seen == [1] * n


def _fd_receiver(rank, tag, out):
    from dcnn_b200.parallel.dp import recv_fd
    fd = recv_fd(tag, rank + 1, timeout_s=30)
    data = os.pread(fd, 1 << 16, 0)       # the duplicates share one file offset: read positionally
    os.close(fd)
    with open(out + f".{rank + 1}", "wb") as f:
        f.write(data)


def test_fd_passing_between_processes(tmp_path):
    """The descriptor hand-off NvlsBuffer uses for the multicast object (SCM_RIGHTS over abstract UNIX sockets): two
    receiver processes each get a working duplicate of a descriptor opened by the sender."""
    import time
    import torch.multiprocessing as mp
    from dcnn_b200.parallel.dp import send_fd_to_ranks
    src = tmp_path / "payload.bin"
    src.write_bytes(b"multicast-handle-stand-in" * 10)
    tag = f"dcnn-test-{os.getpid()}-{time.monotonic_ns()}"
    out = str(tmp_path / "got")
    ctx = mp.spawn(_fd_receiver, args=(tag, out), nprocs=2, join=False)
    fd = os.open(str(src), os.O_RDONLY)
    try:
        send_fd_to_ranks(tag, [1, 2], fd, timeout_s=30)
    finally:
        os.close(fd)
    deadline = time.monotonic() + 60
    while not ctx.join(timeout=1.0):
        assert time.monotonic() < deadline
    for r in (1, 2):
        assert open(out + f".{r}", "rb").read() == src.read_bytes()
