"""SPMD PipelineEngine on CPU (gloo): stages x microbatches (x data-parallel replicas) == single device."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
from dcnn_b200.utils.spawn import spawn_bounded

from dcnn_b200 import nn as dnn
from dcnn_b200.parallel.pipeline import build_schedule
from dcnn_b200.partitioner import NaivePartitioner
from tests.test_pipeline_inprocess import _data, _model, _reference_run


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_schedules():
    assert build_schedule("sync", 0, 2, 3) == [("F", 0), ("F", 1), ("F", 2), ("B", 0), ("B", 1), ("B", 2)]
    # last stage alternates immediately; with the textbook warm-up the first stage runs S-1 forwards ahead (1F1B)
    assert build_schedule("semi_async", 2, 3, 4, 1) == [("F", 0), ("B", 0), ("F", 1), ("B", 1), ("F", 2), ("B", 2),
                                                        ("F", 3), ("B", 3)]
    assert build_schedule("semi_async", 0, 3, 4, 1) == [("F", 0), ("F", 1), ("F", 2), ("B", 0), ("F", 3), ("B", 1),
                                                        ("B", 2), ("B", 3)]
    # default: two microbatches of lead per stage of distance to the end, so transfers hide behind compute
    assert build_schedule("semi_async", 1, 3, 6) == [("F", 0), ("F", 1), ("F", 2), ("B", 0), ("F", 3), ("B", 1),
                                                     ("F", 4), ("B", 2), ("F", 5), ("B", 3), ("B", 4), ("B", 5)]
    assert build_schedule("semi_async", 0, 3, 6)[:5] == [("F", 0), ("F", 1), ("F", 2), ("F", 3), ("F", 4)]
    for factor in (1, 2, 3):
        for st in range(4):
            for m in (2, 5):
                ops = build_schedule("semi_async", st, 4, m, factor)
                assert sorted(ops) == sorted([("F", i) for i in range(m)] + [("B", i) for i in range(m)])
                assert all(ops.index(("F", i)) < ops.index(("B", i)) for i in range(m))
                assert [o for o in ops if o[0] == "B"] == [("B", i) for i in range(m)]


def _proc(rank, world, port, out, stages, reps, micro, schedule):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dcnn_b200.parallel.pipeline import PipelineEngine, make_pipeline_groups
    stage, rep, links, dpg = make_pipeline_groups(stages, reps)
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), stages)
    eng = PipelineEngine(full, parts, stage, dnn.SGD(0.05, 0.9), dnn.SoftmaxCrossEntropyLoss(), micro, links, "cpu",
                         schedule=schedule, dp_group=dpg, seed=77)
    x, t = _data(8)
    per = 8 // reps
    xs, ts = x[rep * per:(rep + 1) * per], t[rep * per:(rep + 1) * per]
    losses = []
    for _ in range(3):
        loss, _ = eng.train_step(xs, ts)
        if loss is not None:
            losses.append(float(loss))
    gathered = [None] * world
    dist.all_gather_object(gathered, (stage, rep, [p.clone() for p in eng.model.parameters()], losses))
    if rank == 0:
        torch.save(gathered, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("stages,reps,micro,schedule", [(2, 1, 2, "semi_async"), (3, 1, 4, "sync"),
                                                        (2, 2, 2, "semi_async")])
def test_engine_equals_single_device(tmp_path, stages, reps, micro, schedule):
    out = str(tmp_path / "eng.pt")
    world = stages * reps
    spawn_bounded(_proc, args=(world, _free_port(), out, stages, reps, micro, schedule), nprocs=world, timeout_s=300)
    got = torch.load(out)
    if reps == 1:
        ref, ref_losses = _reference_run(stages, micro, 3)
    else:
        # replicas with sync-BN and averaged gradients == one device, where every pipeline microbatch is the
        # concatenation of the replicas' microbatches.  Build that reference explicitly.
        ref, ref_losses = _reference_dp(stages, micro, reps, 3)
    params = []
    for s in range(stages):
        entry = [g for g in got if g[0] == s and g[1] == 0][0]
        params.extend(entry[2])
    for p, q in zip(params, ref.parameters()):
        assert torch.allclose(p, q, atol=2e-5, rtol=1e-4), (p - q).abs().max()
    if reps == 1:
        losses = [g for g in got if g[0] == stages - 1][0][3]
        for a, b in zip(losses, ref_losses):
            assert abs(a - b) < 1e-4


def _reference_dp(stages, micro, reps, steps):
    import dcnn_b200 as dc
    from dcnn_b200.train import split_microbatches
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), stages)
    subs = full.split(parts)
    for i, s in enumerate(subs):
        s.set_seed(77 + 1000003 * parts[i].start_layer)
        s.initialize()
    seq = dc.Sequential("ref")
    for s in subs:
        for l in s.layers:
            seq.add(l)
    opt = dnn.SGD(0.05, 0.9)
    opt.attach(seq.parameters(), seq.gradients())
    loss_fn = dnn.SoftmaxCrossEntropyLoss()
    x, t = _data(8)
    per = 8 // reps
    losses = []
    for _ in range(steps):
        xr = [split_microbatches(x[r * per:(r + 1) * per], micro) for r in range(reps)]
        tr = [split_microbatches(t[r * per:(r + 1) * per], micro) for r in range(reps)]
        gs = []
        for m in range(micro):
            xm = torch.cat([xr[r][m] for r in range(reps)])
            tm = torch.cat([tr[r][m] for r in range(reps)])
            y = seq.forward(xm, m)
            loss, g, _ = loss_fn.forward_backward(y, tm)
            gs.append(g)
        for m in reversed(range(micro)):
            seq.backward(gs[m], m)
        opt.update(zero_grads=True)
    return seq, losses


def _proc_eval_gather(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dcnn_b200.parallel.pipeline import PipelineEngine, make_pipeline_groups
    stage, rep, links, dpg = make_pipeline_groups(world, 1)
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), world)
    eng = PipelineEngine(full, parts, stage, dnn.SGD(0.05, 0.9), dnn.SoftmaxCrossEntropyLoss(), 2, links, "cpu", seed=77)
    x, t = _data(8)
    for _ in range(2):
        eng.train_step(x, t)
    ev = eng.eval_step(x, t)
    again = eng.eval_step(x, t)                          # evaluation must not change any state
    gathered = eng.gather_full_model(0)
    if rank == 0:
        gathered.save_to_file(out + "_model", save_state=True)     # + BatchNorm running statistics sidecar
    if eng.is_last:
        torch.save({"loss": float(ev[0]), "correct": float(ev[1]), "loss2": float(again[0])}, out + "_eval.pt")
    dist.barrier()
    dist.destroy_process_group()


def test_eval_step_and_gathered_checkpoint_match_single_device(tmp_path):
    import dcnn_b200 as dc
    out = str(tmp_path / "eg")
    spawn_bounded(_proc_eval_gather, args=(2, _free_port(), out), nprocs=2, timeout_s=300)
    ev = torch.load(out + "_eval.pt")
    assert abs(ev["loss"] - ev["loss2"]) < 1e-7
    # the checkpoint written from the pipelined run is a plain single-device model that evaluates identically
    m = dc.Sequential.from_file(out + "_model")
    m.set_training(False)
    x, t = _data(8)
    y = m.forward(x, 0)
    loss, _, correct = dnn.SoftmaxCrossEntropyLoss().forward_backward(y, t, want_grad=False)
    # eval_step averages the per-microbatch mean losses (2 equal microbatches here == the batch mean)
    assert abs(float(loss) - ev["loss"]) < 1e-5 and float(correct) == ev["correct"]
    ref, _ = _reference_run(2, 2, 2)
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=2e-5, rtol=1e-4)
