"""Coordinator (ref include/pipeline/coordinator.hpp:30-599, in_process_coordinator.hpp:23-60,
distributed_coordinator.hpp:26-50).

Same surface: ``initialize()`` (partition + per-stage ``StageConfig``), ``deploy_stages()`` (CONFIG_TRANSFER ->
CONFIG_RECEIVED), ``start()/stop()``, ``forward/backward(tensor, mb)``, ``join(cmd, n, timeout)``,
``async_process_batch`` (semi-async), ``sync_process_batch`` (GPipe order of
examples/sync_pipeline_coordinator.cpp:122-180), ``update_parameters()``, profiling broadcast, plus the operations the
reference leaves as stubs: ``collect_current_parameters`` (gather stage shards in layer order -> one reference-format
checkpoint), ``send_params``, ``checkpoint``, ``health_check``, ``barrier``, ``report_load``.

B200-first changes: labels ride with the forward job so the LAST stage computes loss+gradient and turns the microbatch
around without a round-trip through the coordinator; the first stage does not ship dL/dinput back (the reference
transfers and discards it); timeouts raise ``PipelineTimeout`` (fail-fast) instead of printing a warning and carrying
on with a hung pipeline (coordinator.hpp:180-183)."""
from __future__ import annotations

import time
from typing import List, Optional, Sequence

import torch

from ...nn.loss import Loss
from ...nn.optimizers import Optimizer
from ...nn.sequential import Partition, Sequential
from ...partitioner import NaivePartitioner, Partitioner
from .communicator import Communicator, InProcessCommunicator
from .protocol import CommandType, Endpoint, Job, Message, StageConfig
from .stage import PipelineStage


class PipelineTimeout(RuntimeError):
    pass


class Coordinator:
    def __init__(self, model: Sequential, optimizer: Optimizer):
        self.model, self.optimizer = model, optimizer
        self.num_stages = 0
        self.num_microbatches = 1
        self.stage_names: List[str] = []
        self.partitions: List[Partition] = []
        self.stage_configs: List[StageConfig] = []
        self.partitioner: Partitioner = NaivePartitioner()
        self.loss_function: Optional[Loss] = None
        self.comm: Optional[Communicator] = None
        self.coordinator_endpoint: Optional[Endpoint] = None
        self.remote_endpoints: List[Endpoint] = []
        self.is_deployed = False
        self.loss_on_last_stage = True
        self.seed: Optional[int] = None
        self.errors: List[str] = []

    # -- configuration ---------------------------------------------------------------------------------
    def set_partitioner(self, p: Partitioner):
        self.partitioner = p

    def set_loss_function(self, loss: Loss):
        if loss is None:
            raise ValueError("Loss function cannot be null")
        self.loss_function = loss

    def add_message_callback(self, callback=None):
        """Call ``callback()`` whenever a message reaches the coordinator (the reference wires its condition variable
        this way, coordinator.hpp:95-101; waits here already block on the communicator's own condition)."""
        self.comm.set_message_notification_callback(callback)

    def set_num_microbatches(self, n: int) -> bool:
        if n <= 0:
            print("Number of microbatches must be positive")
            return False
        self.num_microbatches = int(n)
        return True

    def initialize(self):
        self.stage_names = [f"stage_{i}" for i in range(self.num_stages)]
        self.partitions = self.partitioner.get_partitions(self.model.get_layers(), self.num_stages)
        self._initialize_topology()

    def _initialize_topology(self):
        if len(self.remote_endpoints) != self.num_stages:
            raise RuntimeError("number of stage endpoints must equal number of stages")
        stages = self.model.split(self.partitions)
        opt_cfg = self.optimizer.get_config()
        loss_cfg = self.loss_function.get_config() if (self.loss_function and self.loss_on_last_stage) else None
        self.stage_configs = []
        for i, st in enumerate(stages):
            mc = st.get_config()
            if self.seed is not None:
                mc["seed"] = self.seed + 1000003 * self.partitions[i].start_layer
            prev_ep = self.coordinator_endpoint if i == 0 else self.remote_endpoints[i - 1]
            next_ep = self.coordinator_endpoint if i == self.num_stages - 1 else self.remote_endpoints[i + 1]
            self.stage_configs.append(StageConfig(
                self.stage_names[i], mc, opt_cfg, next_ep, prev_ep, self.coordinator_endpoint,
                loss_cfg if i == self.num_stages - 1 else None, i == 0, i == self.num_stages - 1))

    def deploy_stages(self, timeout: float = 60.0) -> bool:
        if self.is_deployed:
            print("Stages already deployed")
            return True
        for name, ep in zip(self.stage_names, self.remote_endpoints):
            self.comm.connect(name, ep)
        for name, cfg in zip(self.stage_names, self.stage_configs):
            self._deliver_config(name, cfg)
        if not self.join(CommandType.CONFIG_RECEIVED, self.num_stages, timeout, raise_on_timeout=False):
            print("Not all stages reported ready")
            return False
        self.comm.dequeue_all_messages_by_type(CommandType.CONFIG_RECEIVED)
        self.is_deployed = True
        print(f"Deployed {self.num_stages} pipeline stages")
        return True

    def _deliver_config(self, name: str, cfg: StageConfig):
        self.comm.send_message(Message(name, CommandType.CONFIG_TRANSFER, cfg.dumps(), "coordinator"))

    # -- lifecycle ---------------------------------------------------------------------------------------
    def _broadcast(self, ctype: CommandType, payload=None):
        for name in self.stage_names:
            self.comm.send_message(Message(name, ctype, payload, "coordinator"))

    def start(self):
        self._broadcast(CommandType.TRAIN_MODE)
        print(f"Started all {self.num_stages} pipeline stages")

    def set_training(self, training: bool):
        self._broadcast(CommandType.TRAIN_MODE if training else CommandType.EVAL_MODE)

    def stop(self):
        self._broadcast(CommandType.SHUTDOWN)
        print("Stopped all pipeline stages")

    # -- data path ---------------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, microbatch_id: int, targets: Optional[torch.Tensor] = None):
        if not self.stage_names:
            raise RuntimeError("No stages available for processing")
        self.comm.send_message(Message(self.stage_names[0], CommandType.FORWARD_JOB,
                                       Job(x, microbatch_id, targets), "coordinator"))

    def backward(self, gradient: torch.Tensor, microbatch_id: int):
        if not self.stage_names:
            raise RuntimeError("No stages available for processing")
        self.comm.send_message(Message(self.stage_names[-1], CommandType.BACKWARD_JOB,
                                       Job(gradient, microbatch_id), "coordinator"))

    def compute_loss(self, predictions, targets) -> float:
        if self.loss_function is None:
            raise RuntimeError("Loss function is not set in the coordinator")
        return self.loss_function.compute_loss(predictions, targets)

    def compute_gradient(self, predictions, targets):
        if self.loss_function is None:
            raise RuntimeError("Loss function is not set in the coordinator")
        return self.loss_function.compute_gradient(predictions, targets)

    def join(self, ctype: CommandType, expected: int, timeout: float = 60.0, raise_on_timeout: bool = True) -> bool:
        ok = self.comm.wait_for_count(ctype, expected, timeout)
        self._drain_errors()
        if not ok and raise_on_timeout:
            raise PipelineTimeout(f"timeout waiting for {expected} x {ctype.name} "
                                  f"(got {self.comm.message_count(ctype)}); stage errors: {self.errors[-3:]}")
        return ok

    def _drain_errors(self):
        for m in self.comm.dequeue_all_messages_by_type(CommandType.ERROR_REPORT):
            self.errors.append(str(m.payload))
            print(f"[pipeline error] {m.payload}")
        if self.errors:
            raise RuntimeError("pipeline stage failed: " + self.errors[-1])

    def async_process_batch(self, microbatch_inputs: Sequence[torch.Tensor],
                            microbatch_labels: Sequence[torch.Tensor], timeout: float = 120.0) -> float:
        """Semi-async schedule: inject every microbatch; each one's backward starts as soon as its forward leaves the
        last stage.  Returns the mean microbatch loss (coordinator.hpp:273-326)."""
        m = self.num_microbatches
        if len(microbatch_inputs) != m or len(microbatch_labels) != m:
            raise ValueError("Microbatch size mismatch with coordinator configuration")
        if self.loss_on_last_stage:
            for i in range(m):
                self.forward(microbatch_inputs[i], i, microbatch_labels[i])
            self.join(CommandType.BACKWARD_JOB, m, timeout)
            self.comm.dequeue_all_messages_by_type(CommandType.BACKWARD_JOB)
            self.comm.dequeue_all_messages_by_type(CommandType.FORWARD_JOB)
            return float("nan")  # the loss lives on the last stage; update_parameters() returns it
        for i in range(m):
            self.forward(microbatch_inputs[i], i)
        total, done = 0.0, 0
        deadline = time.monotonic() + timeout
        while done < m:
            self.comm.wait_for_input(0.05)
            self._drain_errors()
            for msg in self.comm.dequeue_all_messages_by_type(CommandType.FORWARD_JOB):
                job = msg.payload
                tgt = microbatch_labels[job.micro_batch_id].to(job.data.device)
                loss, grad, _ = self.loss_function.forward_backward(job.data, tgt)
                total += float(loss)
                self.backward(grad, job.micro_batch_id)
                done += 1
            if time.monotonic() > deadline:
                raise PipelineTimeout("timeout waiting for forward results")
        self.join(CommandType.BACKWARD_JOB, m, timeout)
        self.comm.dequeue_all_messages_by_type(CommandType.BACKWARD_JOB)
        return total / m if m else total

    def sync_process_batch(self, microbatch_inputs, microbatch_labels, timeout: float = 120.0) -> float:
        """GPipe schedule: all forwards, join, loss/gradient on the coordinator, all backwards, join."""
        m = self.num_microbatches
        for i in range(m):
            self.forward(microbatch_inputs[i], i)
        self.join(CommandType.FORWARD_JOB, m, timeout)
        fwd = self.comm.dequeue_all_messages_by_type(CommandType.FORWARD_JOB)
        total = 0.0
        grads = {}
        self.last_correct = 0.0          # correct predictions of this batch (classification losses report them)
        for msg in fwd:
            job = msg.payload
            tgt = microbatch_labels[job.micro_batch_id].to(job.data.device)
            loss, grad, correct = self.loss_function.forward_backward(job.data, tgt)
            total += float(loss)
            if correct is not None:
                self.last_correct += float(correct)
            grads[job.micro_batch_id] = grad
        for i in sorted(grads):
            self.backward(grads[i], i)
        self.join(CommandType.BACKWARD_JOB, m, timeout)
        self.comm.dequeue_all_messages_by_type(CommandType.BACKWARD_JOB)
        return total / m if m else total

    def update_parameters(self, timeout: float = 60.0) -> dict:
        self._broadcast(CommandType.UPDATE_PARAMETERS)
        self.join(CommandType.PARAMETERS_UPDATED, self.num_stages, timeout)
        out = {"loss": float("nan"), "correct": 0.0}
        for m in self.comm.dequeue_all_messages_by_type(CommandType.PARAMETERS_UPDATED):
            p = m.payload or {}
            if p.get("count", 0) > 0:
                out = {"loss": p["loss_sum"] / p["count"], "correct": p["correct_sum"]}
        return out

    # -- monitoring --------------------------------------------------------------------------------------
    def print_profiling_on_all_stages(self, timeout: float = 30.0):
        self._broadcast(CommandType.PRINT_PROFILING)
        self.join(CommandType.PROFILING_PRINTED, self.num_stages, timeout)
        self.comm.dequeue_all_messages_by_type(CommandType.PROFILING_PRINTED)

    def clear_profiling_data(self):
        self._broadcast(CommandType.CLEAR_PROFILING)
        self.join(CommandType.PROFILING_CLEARED, self.num_stages, 30.0)
        self.comm.dequeue_all_messages_by_type(CommandType.PROFILING_CLEARED)

    def fetch_loss_stats(self, timeout: float = 30.0) -> dict:
        """Loss / correct-count accumulated on the last stage since the last fetch or update (used by validation,
        where no optimizer step must happen)."""
        self._broadcast(CommandType.STATUS_REQUEST, "reset_stats")
        self.join(CommandType.STATUS_RESPONSE, self.num_stages, timeout)
        out = {"loss": float("nan"), "correct": 0.0}
        for m in self.comm.dequeue_all_messages_by_type(CommandType.STATUS_RESPONSE):
            p = m.payload or {}
            if p.get("count", 0) > 0:
                out = {"loss": p["loss_sum"] / p["count"], "correct": p["correct_sum"]}
        return out

    def request_status_from_all_stages(self, timeout: float = 30.0):
        self._broadcast(CommandType.STATUS_REQUEST)
        self.join(CommandType.STATUS_RESPONSE, self.num_stages, timeout)
        return [m.payload for m in self.comm.dequeue_all_messages_by_type(CommandType.STATUS_RESPONSE)]

    def health_check(self, timeout: float = 10.0) -> bool:
        self._broadcast(CommandType.HEALTH_CHECK)
        ok = self.join(CommandType.HEALTH_CHECK, self.num_stages, timeout, raise_on_timeout=False)
        self.comm.dequeue_all_messages_by_type(CommandType.HEALTH_CHECK)
        return ok

    def barrier(self, timeout: float = 60.0):
        self._broadcast(CommandType.BARRIER_SYNC)
        self.join(CommandType.BARRIER_SYNC, self.num_stages, timeout)
        self.comm.dequeue_all_messages_by_type(CommandType.BARRIER_SYNC)

    def report_load(self, timeout: float = 30.0):
        self._broadcast(CommandType.REPORT_LOAD)
        self.join(CommandType.LOAD_REPORT, self.num_stages, timeout)
        return {m.sender_id: m.payload for m in self.comm.dequeue_all_messages_by_type(CommandType.LOAD_REPORT)}

    def balance_load(self, timeout: float = 30.0) -> List[Partition]:
        """Re-partition from measured per-stage times (the reference lists this as "Coming Soon")."""
        from ...partitioner import BalancedPartitioner
        loads = self.report_load(timeout)
        layers = self.model.get_layers()
        costs = []
        for i, part in enumerate(self.partitions):
            lt = loads.get(self.stage_names[i])
            t = (lt.avg_forward_time_ms + lt.avg_backward_time_ms) if lt else 1.0
            n = part.end_layer - part.start_layer
            costs.extend([max(t, 1e-6) / n] * n)
        return BalancedPartitioner(costs=costs).get_partitions(layers, self.num_stages)

    # -- parameters / checkpoint -------------------------------------------------------------------------
    def collect_current_parameters(self, timeout: float = 60.0) -> bool:
        """Gather every stage's parameters (and BN running stats) into ``self.model`` in layer order, so that
        ``self.model.save_to_file`` writes one reference-compatible .json/.bin for the whole network."""
        self._broadcast(CommandType.SEND_PARAMS)
        if not self.join(CommandType.PARAMS_TRANSFER, self.num_stages, timeout, raise_on_timeout=False):
            print("Warning: Not all stages sent their parameters within timeout.")
            return False
        msgs = {m.payload["stage"]: m.payload for m in
                self.comm.dequeue_all_messages_by_type(CommandType.PARAMS_TRANSFER)}
        self.model.initialize()
        params = []
        for name in self.stage_names:
            params.extend(msgs[name]["params"])
        self.model.load_parameters(params)
        for i, part in enumerate(self.partitions):
            sub = Sequential("tmp")
            sub.layers = self.model.layers[part.start_layer:part.end_layer]
            got = msgs[self.stage_names[i]]["buffers"]
            for k, b in sub.buffers().items():
                if k in got:
                    b.copy_(got[k])
        return True

    def send_params(self, stage_index: int, timeout: float = 60.0) -> bool:
        """Push the coordinator's copy of a stage's parameters to that stage."""
        part = self.partitions[stage_index]
        sub = Sequential("tmp")
        sub.layers = self.model.layers[part.start_layer:part.end_layer]
        payload = {"params": [p.detach().cpu() for p in sub.parameters()],
                   "buffers": {k: v.detach().cpu() for k, v in sub.buffers().items()}}
        self.comm.send_message(Message(self.stage_names[stage_index], CommandType.LOAD_PARAMS, payload, "coordinator"))
        ok = self.join(CommandType.PARAMS_LOADED, 1, timeout, raise_on_timeout=False)
        self.comm.dequeue_all_messages_by_type(CommandType.PARAMS_LOADED)
        return ok

    def send_all_params(self) -> bool:
        return all(self.send_params(i) for i in range(self.num_stages))

    # -- recovery ----------------------------------------------------------------------------------------
    def replace_stage(self, stage_index: int, endpoint: Endpoint, timeout: float = 60.0,
                      restore_parameters: bool = True) -> bool:
        """Bring a fresh worker at ``endpoint`` into the pipeline in place of stage ``stage_index`` (which crashed or
        was lost): deploy the stage's configuration to it, make its neighbours re-dial it, discard whatever the
        surviving stages kept of the batch that was in flight, and -- with ``restore_parameters`` -- load EVERY stage
        with the coordinator's copy of the parameters (the last ``collect_current_parameters`` / ``save_checkpoint``),
        so that all stages continue from the same snapshot.  Optimizer moments of the replaced stage start from zero.
        The reference has no recovery path (a lost worker ends the run)."""
        if not 0 <= stage_index < self.num_stages:
            raise IndexError("stage index out of range")
        name = self.stage_names[stage_index]
        self.errors.clear()
        for c in CommandType:                       # answers and error reports of the failed batch are void
            if c != CommandType._START and c != CommandType._COUNT:
                self.comm.dequeue_all_messages_by_type(c)
        self.remote_endpoints[stage_index] = endpoint
        self._initialize_topology()
        try:
            self.comm.disconnect(name)
        except Exception:
            pass
        if not self.comm.connect(name, endpoint):
            return False
        self._deliver_config(name, self.stage_configs[stage_index])
        if not self.join(CommandType.CONFIG_RECEIVED, 1, timeout, raise_on_timeout=False):
            print(f"Replacement for {name} did not report ready")
            return False
        self.comm.dequeue_all_messages_by_type(CommandType.CONFIG_RECEIVED)
        survivors = 0
        for i, other in enumerate(self.stage_names):
            if i == stage_index:
                continue
            rewire = {}
            if i == stage_index - 1:
                rewire["next_stage"] = self._endpoint_payload(endpoint)
            if i == stage_index + 1:
                rewire["prev_stage"] = self._endpoint_payload(endpoint)
            self.comm.send_message(Message(other, CommandType.CONFIG_TRANSFER, {"rewire": rewire, "reset": True},
                                           "coordinator"))
            survivors += 1
        if survivors and not self.join(CommandType.CONFIG_RECEIVED, survivors, timeout, raise_on_timeout=False):
            print("Not all surviving stages acknowledged the new topology")
            return False
        self.comm.dequeue_all_messages_by_type(CommandType.CONFIG_RECEIVED)
        if restore_parameters:
            have = bool(self.model.layers) and all(p is not None for p in self.model.parameters())
            if not have:
                print("Warning: no parameter snapshot on the coordinator (collect_current_parameters was never "
                      "called): the replacement starts from its initialisation, the other stages keep their weights")
            elif not self.send_all_params():
                return False
        self.set_training(True)
        print(f"Replaced {name}")
        return True

    def _endpoint_payload(self, endpoint: Endpoint):
        """In-process endpoints carry live objects and travel as they are; the others as JSON."""
        return endpoint if endpoint.kind == "in_process" else endpoint.to_json()

    def save_checkpoint(self, path: str) -> bool:
        if not self.collect_current_parameters():
            return False
        self.model.save_to_file(path, save_state=True)
        return True


class InProcessCoordinator(Coordinator):
    """All stages as threads of this process, one per device (ref in_process_coordinator.hpp)."""

    def __init__(self, model: Sequential, optimizer: Optimizer, stage_devices: Sequence = ("cpu",)):
        super().__init__(model, optimizer)
        self.num_stages = len(stage_devices)
        self.comm = InProcessCommunicator("coordinator")
        self.coordinator_endpoint = Endpoint.in_process(self.comm)
        self.stages: List[PipelineStage] = []
        for i, dev in enumerate(stage_devices):
            c = InProcessCommunicator(f"stage_{i}")
            self.stages.append(PipelineStage(c, f"stage_{i}", device=dev))
            self.remote_endpoints.append(Endpoint.in_process(c))

    def get_stages(self) -> List[PipelineStage]:
        return list(self.stages)

    def _deliver_config(self, name, cfg):
        # in-process endpoints hold live communicator objects: hand the config object over as is
        self.comm.send_message(Message(name, CommandType.CONFIG_TRANSFER, cfg, "coordinator"))

    def deploy_stages(self, timeout: float = 60.0) -> bool:
        for s in self.stages:
            if s._thread is None:
                s.start(threaded=True)
        return super().deploy_stages(timeout)

    def stop(self):
        super().stop()
        for s in self.stages:
            s.stop()


class DistributedCoordinator(Coordinator):
    """Coordinator as rank ``coordinator_rank`` of a torch.distributed job; stage ``i`` is rank ``stage_ranks[i]``
    running a ``NetworkStageWorker``.  Replaces the ASIO/TCP DistributedCoordinator."""

    def __init__(self, model, optimizer, store, control_group, coordinator_rank: int, stage_ranks: Sequence[int],
                 stage_devices: Optional[Sequence[str]] = None):
        super().__init__(model, optimizer)
        from .communicator import DistCommunicator
        self.num_stages = len(stage_ranks)
        self.comm = DistCommunicator("coordinator", coordinator_rank, store, control_group, None, None,
                                     torch.device("cpu"))
        self.coordinator_endpoint = Endpoint.network(coordinator_rank, "cpu")
        devs = stage_devices or ["cpu"] * len(stage_ranks)
        self.remote_endpoints = [Endpoint.network(r, d) for r, d in zip(stage_ranks, devs)]
