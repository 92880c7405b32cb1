"""Pipeline control protocol (ref include/pipeline/command_type.hpp:20-68, message.hpp:21-166, job.hpp:14-35,
stage_config.hpp:8-35, endpoint.hpp:17-95, load_tracker.hpp:9-22).

Same command vocabulary and the same JSON stage-config schema as the reference.  Differences by design: tensors
never travel inside control messages on the distributed backend (activations/gradients go GPU->GPU over NCCL p2p;
the message only carries shape metadata), and endpoints are ``{rank, device}`` instead of host:port."""
from __future__ import annotations

import enum
import json
from dataclasses import dataclass, field
from typing import Any, Optional

import torch


class CommandType(enum.IntEnum):
    _START = 0
    # core commands
    FORWARD_JOB = 1
    BACKWARD_JOB = 2
    UPDATE_PARAMETERS = 3
    # mode switching
    TRAIN_MODE = 4
    EVAL_MODE = 5
    SHUTDOWN = 6
    # configuration
    CONFIG_TRANSFER = 7
    CONFIG_RECEIVED = 8
    LOAD_PARAMS = 9
    PARAMS_LOADED = 10
    SEND_PARAMS = 11
    PARAMS_TRANSFER = 12
    # status and monitoring
    STATUS_REQUEST = 13
    STATUS_RESPONSE = 14
    PARAMETERS_UPDATED = 15
    HEALTH_CHECK = 16
    # error handling
    ERROR_REPORT = 17
    JOB_FAILURE = 18
    # advanced features
    BARRIER_SYNC = 19
    CHECKPOINT_REQUEST = 20
    CHECKPOINT_COMPLETE = 21
    # load balancing and resource management
    UPDATE_LOAD = 22
    REPORT_LOAD = 23
    LOAD_REPORT = 24
    # profiling and debugging
    PRINT_PROFILING = 25
    PROFILING_PRINTED = 26
    CLEAR_PROFILING = 27
    PROFILING_CLEARED = 28
    _COUNT = 29


@dataclass
class Job:
    """A tensor plus the microbatch it belongs to.  ``targets`` (optional) rides along with forward jobs so the
    LAST stage can compute loss + gradient locally instead of shipping logits back to the coordinator."""
    data: Optional[torch.Tensor]
    micro_batch_id: int = 0
    targets: Optional[torch.Tensor] = None


@dataclass
class LoadTracker:
    avg_forward_time_ms: float = 0.0
    avg_backward_time_ms: float = 0.0
    avg_cpu_utilization: float = 0.0
    max_memory_usage_mb: float = 0.0


@dataclass
class Message:
    recipient_id: str
    command_type: CommandType
    payload: Any = None          # None | Job | str | bool | float | LoadTracker | dict
    sender_id: str = ""

    def has_job(self) -> bool:
        return isinstance(self.payload, Job)

    def has_type(self, kind) -> bool:
        """``msg.has_type(Job)`` / ``has_type(str)`` / ``has_type(bool)`` (ref Message::has_type<T>())."""
        if isinstance(self.payload, bool):          # bool is an int subclass: only has_type(bool) matches it
            return kind is bool
        return isinstance(self.payload, kind)


@dataclass
class Endpoint:
    """Where a pipeline participant lives.  ``in_process`` endpoints carry the communicator object itself (ref
    Endpoint::in_process); ``rank`` endpoints name a torch.distributed rank; ``tcp`` endpoints are the reference's
    ``Endpoint::network(host, port)`` (endpoint.hpp:46-51), served by the native TCP transport."""
    kind: str = "rank"                 # "rank" | "in_process" | "tcp"
    rank: int = -1
    device: str = "cpu"
    communicator: Any = field(default=None, repr=False, compare=False)
    host: str = ""
    port: int = 0

    @staticmethod
    def network(rank: int, device: str = "cpu") -> "Endpoint":
        return Endpoint("rank", int(rank), device)

    @staticmethod
    def tcp(host: str, port: int, device: str = "cpu") -> "Endpoint":
        return Endpoint("tcp", -1, device, None, host, int(port))

    @staticmethod
    def in_process(communicator) -> "Endpoint":
        return Endpoint("in_process", -1, "cpu", communicator)

    # the reference's generic accessors (endpoint.hpp:28-44)
    def communication_type(self) -> str:
        return self.kind

    def get_parameter(self, key: str):
        if key in ("host", "port", "rank", "device", "communicator"):
            v = getattr(self, key)
            return str(v) if key == "port" else v
        raise KeyError(f"Parameter {key} not found")

    def set_parameter(self, key: str, value):
        if key not in ("host", "port", "rank", "device", "communicator"):
            raise KeyError(f"Parameter {key} not found")
        setattr(self, key, int(value) if key in ("port", "rank") else value)

    def to_json(self) -> dict:
        if self.kind == "tcp":     # the reference's schema (endpoint.hpp:59-83): parameters are strings
            return {"communication_type_": "tcp",
                    "parameters_": {"host": self.host, "port": str(self.port), "device": self.device}}
        return {"communication_type": self.kind, "rank": self.rank, "device": self.device}

    @staticmethod
    def from_json(j: dict) -> "Endpoint":
        if "communication_type_" in j:
            p = j.get("parameters_", {})
            return Endpoint(j["communication_type_"], -1, p.get("device", "cpu"), None, p.get("host", ""),
                            int(p.get("port", 0)))
        return Endpoint(j.get("communication_type", "rank"), int(j.get("rank", -1)), j.get("device", "cpu"))


@dataclass
class StageConfig:
    """JSON-serialisable stage description (ref stage_config.hpp:16-23)."""
    stage_id: str
    model_config: dict
    optimizer_config: dict
    next_stage_endpoint: Endpoint
    prev_stage_endpoint: Endpoint
    coordinator_endpoint: Endpoint
    loss_config: Optional[dict] = None     # extension: the last stage owns the loss
    is_first: bool = False
    is_last: bool = False

    def to_json(self) -> dict:
        return {"stage_id": self.stage_id, "model_config": self.model_config,
                "optimizer_config": self.optimizer_config,
                "next_stage_endpoint": self.next_stage_endpoint.to_json(),
                "prev_stage_endpoint": self.prev_stage_endpoint.to_json(),
                "coordinator_endpoint": self.coordinator_endpoint.to_json(),
                "loss_config": self.loss_config, "is_first": self.is_first, "is_last": self.is_last}

    def dumps(self) -> str:
        return json.dumps(self.to_json())

    @staticmethod
    def from_json(j) -> "StageConfig":
        if isinstance(j, str):
            j = json.loads(j)
        return StageConfig(j["stage_id"], j["model_config"], j["optimizer_config"],
                           Endpoint.from_json(j["next_stage_endpoint"]),
                           Endpoint.from_json(j["prev_stage_endpoint"]),
                           Endpoint.from_json(j["coordinator_endpoint"]), j.get("loss_config"),
                           bool(j.get("is_first", False)), bool(j.get("is_last", False)))
