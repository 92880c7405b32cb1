"""PipelineStage / NetworkStageWorker (ref include/pipeline/pipeline_stage.hpp:29-308,
network_stage_worker.hpp:25-115): an event loop that dequeues messages by priority and

  FORWARD_JOB        -> model.forward(mb)  -> send to next stage (or, on the last stage with targets attached,
                                              compute loss+gradient locally and start the backward immediately)
  BACKWARD_JOB       -> model.backward(mb) -> send to previous stage (the first stage sends only an ack: the
                                              reference ships dL/dinput back to the coordinator, which discards it)
  UPDATE_PARAMETERS  -> optimizer.update() + clear_gradients -> PARAMETERS_UPDATED
  CONFIG_TRANSFER    -> build model + optimizer (+ loss) from JSON, set_device, initialize, connect peers
  TRAIN_MODE / EVAL_MODE / PRINT_PROFILING / CLEAR_PROFILING / UPDATE_LOAD / SEND_PARAMS / LOAD_PARAMS /
  CHECKPOINT_REQUEST / STATUS_REQUEST / HEALTH_CHECK / BARRIER_SYNC / SHUTDOWN

The last five groups are stubs in the reference (pipeline_stage.hpp:163-179); here they are implemented so that
weights trained through the pipeline can be gathered and checkpointed."""
from __future__ import annotations

import threading
import time
import traceback
from typing import Optional

import torch

from ...nn.loss import LossFactory
from ...nn.optimizers import OptimizerFactory
from ...nn.sequential import Sequential
from .communicator import Communicator
from .protocol import CommandType, Job, LoadTracker, Message, StageConfig


class PipelineStage:
    def __init__(self, communicator: Communicator, name: str = "", model: Optional[Sequential] = None,
                 optimizer=None, loss_fn=None, device=None):
        self.comm = communicator
        self.name = name or communicator.id
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        self.device = torch.device(device) if device is not None else None
        self.is_first = self.is_last = False
        self.should_stop = False
        self.load = LoadTracker()
        self._thread: Optional[threading.Thread] = None
        self._fwd_ms = self._bwd_ms = 0.0
        self._fwd_n = self._bwd_n = 0
        self.loss_sum = 0.0
        self.correct_sum = 0.0
        self.loss_count = 0
        self._hw = None
        if model is not None and optimizer is not None:
            optimizer.attach(model.parameters(), model.gradients())

    def get_stage_id(self) -> str:
        return self.name

    def get_hardware_info(self):
        """This worker's HardwareInfo (ref NetworkStageWorker::get_hardware_info)."""
        if self._hw is None:
            from ...utils.hardware import HardwareInfo
            self._hw = HardwareInfo()
            self._hw.initialize()
        return self._hw

    def is_configured(self) -> bool:
        return self.model is not None and self.optimizer is not None

    # -- lifecycle ---------------------------------------------------------------------------------------
    def start(self, threaded: bool = True):
        self.should_stop = False
        if threaded:
            self._thread = threading.Thread(target=self.message_loop, name=self.name, daemon=True)
            self._thread.start()
        else:
            self.message_loop()

    def stop(self):
        self.should_stop = True
        if self._thread is not None:
            self._thread.join(timeout=10)

    def message_loop(self):
        # the CUDA current device is per thread: a stage thread of a single-process multi-GPU pipeline must select its
        # own GPU, or its kernels would be launched on device 0 with another device's pointers
        if self.device is not None and torch.device(self.device).type == "cuda":
            torch.cuda.set_device(torch.device(self.device))
        while not self.should_stop:
            if not self.comm.wait_for_input(timeout=0.05):
                continue
            while True:
                msg = self.comm.dequeue_input_message()
                if msg is None:
                    break
                try:
                    self.process_message(msg)
                except Exception as e:  # report instead of dying silently (ref pipeline_stage.hpp:276-282)
                    self._send("coordinator", CommandType.ERROR_REPORT,
                               f"{self.name}: {type(e).__name__}: {e}\n{traceback.format_exc()}")
                if self.should_stop:
                    break
        self.comm.flush_output_messages()

    # -- helpers -----------------------------------------------------------------------------------------
    def _send(self, recipient: str, ctype: CommandType, payload=None):
        self.comm.send_message(Message(recipient, ctype, payload, self.name))

    def _timed(self, fn, *a):
        t0 = time.perf_counter()
        out = fn(*a)
        return out, (time.perf_counter() - t0) * 1000.0

    # -- dispatch ----------------------------------------------------------------------------------------
    def process_message(self, msg: Message):
        c = msg.command_type
        if c == CommandType.FORWARD_JOB:
            self._forward_job(msg.payload)
        elif c == CommandType.BACKWARD_JOB:
            self._backward_job(msg.payload)
        elif c == CommandType.UPDATE_PARAMETERS:
            self.optimizer.update(zero_grads=True)
            self._send("coordinator", CommandType.PARAMETERS_UPDATED,
                       {"loss_sum": self.loss_sum, "correct_sum": self.correct_sum, "count": self.loss_count})
            self.loss_sum = self.correct_sum = 0.0
            self.loss_count = 0
        elif c == CommandType.TRAIN_MODE:
            self.model.set_training(True)
        elif c == CommandType.EVAL_MODE:
            self.model.set_training(False)
        elif c == CommandType.SHUTDOWN:
            self.should_stop = True
        elif c == CommandType.CONFIG_TRANSFER:
            self.handle_configuration(msg.payload)
        elif c == CommandType.PRINT_PROFILING:
            if self.model.is_profiling_enabled():
                self.model.print_profiling_summary()
            else:
                print(f"Warning: Profiling not enabled for stage {self.name}")
            self._send("coordinator", CommandType.PROFILING_PRINTED)
        elif c == CommandType.CLEAR_PROFILING:
            self.model.clear_profiling_data()
            self._send("coordinator", CommandType.PROFILING_CLEARED)
        elif c in (CommandType.UPDATE_LOAD, CommandType.REPORT_LOAD):
            self.load.avg_forward_time_ms = self._fwd_ms / max(1, self._fwd_n)
            self.load.avg_backward_time_ms = self._bwd_ms / max(1, self._bwd_n)
            if torch.cuda.is_available() and self.device is not None and self.device.type == "cuda":
                self.load.max_memory_usage_mb = torch.cuda.max_memory_allocated(self.device) / 1048576.0
            else:
                from ...utils.misc import get_memory_usage_kb
                self.load.max_memory_usage_mb = get_memory_usage_kb() / 1024.0
            # CPU utilisation since the previous UPDATE_LOAD (ref pipeline_stage.hpp:199-229 folds HardwareInfo's
            # dynamic info into the LoadTracker the same way)
            if self._hw is None:
                from ...utils.hardware import HardwareInfo
                self._hw = HardwareInfo()
                self._hw.initialize()
            self.load.avg_cpu_utilization = float(self._hw.update_dynamic_info()["cpu_utilization_percent"])
            if c == CommandType.REPORT_LOAD:
                self._send("coordinator", CommandType.LOAD_REPORT, self.load)
        elif c == CommandType.SEND_PARAMS:
            params = [p.detach().to("cpu").contiguous() for p in self.model.parameters()]
            bufs = {k: v.detach().to("cpu") for k, v in self.model.buffers().items()}
            self._send("coordinator", CommandType.PARAMS_TRANSFER, {"stage": self.name, "params": params, "buffers": bufs})
        elif c == CommandType.LOAD_PARAMS:
            self.model.load_parameters(msg.payload["params"])
            for (k, b) in self.model.buffers().items():
                if k in msg.payload.get("buffers", {}):
                    b.copy_(msg.payload["buffers"][k].to(b.device))
            self._send("coordinator", CommandType.PARAMS_LOADED)
        elif c == CommandType.CHECKPOINT_REQUEST:
            self.model.save_to_file(str(msg.payload) + "_" + self.name, save_state=True)
            self._send("coordinator", CommandType.CHECKPOINT_COMPLETE, self.name)
        elif c == CommandType.STATUS_REQUEST:
            self._send("coordinator", CommandType.STATUS_RESPONSE,
                       {"stage": self.name, "layers": self.model.layer_size() if self.model else 0,
                        "training": self.model.is_training() if self.model else None,
                        "loss_sum": self.loss_sum, "correct_sum": self.correct_sum, "count": self.loss_count})
            if msg.payload == "reset_stats":
                self.loss_sum = self.correct_sum = 0.0
                self.loss_count = 0
        elif c == CommandType.HEALTH_CHECK:
            self._send("coordinator", CommandType.HEALTH_CHECK, self.name)
        elif c == CommandType.BARRIER_SYNC:
            if self.device is not None and self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            self._send("coordinator", CommandType.BARRIER_SYNC, self.name)
        elif c == CommandType.ERROR_REPORT:
            print(f"{self.name} received error report from {msg.sender_id}: {msg.payload}")
        else:
            raise RuntimeError(f"Unknown command type {int(c)} received by {self.name}")

    def _forward_job(self, job: Job):
        x = job.data
        out, ms = self._timed(self.model.forward, x, job.micro_batch_id)
        self._fwd_ms += ms
        self._fwd_n += 1
        if self.is_last and job.targets is not None and self.loss_fn is not None:
            # loss lives on the last stage: turn around immediately (semi-async "immediate backward")
            t = job.targets.to(out.device)
            if self.model.is_training():
                loss, grad, correct = self.loss_fn.forward_backward(out, t)
                self._accumulate(loss, correct)
                self._backward_job(Job(grad, job.micro_batch_id))
                self._send("coordinator", CommandType.FORWARD_JOB, Job(None, job.micro_batch_id))
            else:
                loss, _, correct = self.loss_fn.forward_backward(out, t, want_grad=False)
                self._accumulate(loss, correct)
                self._send("coordinator", CommandType.FORWARD_JOB, Job(None, job.micro_batch_id))
        else:
            self._send("next_stage", CommandType.FORWARD_JOB, Job(out, job.micro_batch_id, job.targets))

    def _accumulate(self, loss, correct):
        self.loss_sum += float(loss)
        if correct is not None:
            self.correct_sum += float(correct)
        self.loss_count += 1

    def _backward_job(self, job: Job):
        grad, ms = self._timed(self.model.backward, job.data, job.micro_batch_id)
        self._bwd_ms += ms
        self._bwd_n += 1
        if self.is_first:
            self._send("prev_stage", CommandType.BACKWARD_JOB, Job(None, job.micro_batch_id))
        else:
            self._send("prev_stage", CommandType.BACKWARD_JOB, Job(grad, job.micro_batch_id))

    def _rewire(self, req: dict):
        """Recovery message from the coordinator (``Coordinator.replace_stage``): re-dial the named neighbours and, with
        ``reset``, drop everything that belongs to the batch that was in flight when a stage was lost -- cached
        activations, partially accumulated gradients, loss counters."""
        from .protocol import Endpoint
        for role, ep in (req.get("rewire") or {}).items():
            ep = ep if isinstance(ep, Endpoint) else Endpoint.from_json(ep)
            try:
                self.comm.disconnect(role)
            except Exception:
                pass
            if not self.comm.connect(role, ep):
                raise ConnectionError(f"{self.name}: cannot reach the new {role}")
        if req.get("reset") and self.model is not None:
            self.model.clear_gradients()
            self.model.clear_cache()
            self.loss_sum = self.correct_sum = 0.0
            self.loss_count = 0
        self._send("coordinator", CommandType.CONFIG_RECEIVED, True)

    def handle_configuration(self, payload):
        if isinstance(payload, dict) and ("rewire" in payload or "reset" in payload):
            return self._rewire(payload)
        cfg = payload if isinstance(payload, StageConfig) else StageConfig.from_json(payload)
        self.name = cfg.stage_id
        self.is_first, self.is_last = cfg.is_first, cfg.is_last
        self.model = Sequential.load_from_config(cfg.model_config)
        self.model.compute_input_gradient = not cfg.is_first
        self.optimizer = OptimizerFactory.create_from_config(cfg.optimizer_config)
        self.loss_fn = LossFactory.create_from_config(cfg.loss_config) if cfg.loss_config else None
        if self.device is not None:
            self.model.set_device(self.device)
        seed = cfg.model_config.get("seed")
        if seed is not None:
            self.model.set_seed(int(seed))
        self.model.initialize()
        self.model.enable_profiling(True)
        self.optimizer.attach(self.model.parameters(), self.model.gradients())
        self.comm.connect("coordinator", cfg.coordinator_endpoint)
        self.comm.connect("next_stage", cfg.next_stage_endpoint)
        self.comm.connect("prev_stage", cfg.prev_stage_endpoint)
        self._send("coordinator", CommandType.CONFIG_RECEIVED, True)


class NetworkStageWorker(PipelineStage):
    """One OS process per stage (the reference's ``network_worker <port> [--gpu]``): joins the torch.distributed job,
    builds a DistCommunicator and serves until SHUTDOWN."""

    def __init__(self, rank: int, store, control_group, data_group, data_ranks, use_gpu: bool = False,
                 device_index: int = 0):
        device = torch.device("cuda", device_index) if use_gpu else torch.device("cpu")
        from .communicator import DistCommunicator
        comm = DistCommunicator(f"worker_{rank}", rank, store, control_group, data_group, data_ranks, device)
        super().__init__(comm, f"worker_{rank}", device=device)
