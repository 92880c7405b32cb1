"""Device registry with the reference's vocabulary (ref include/device/device_manager.hpp:9-74, device_type.hpp):
devices are keyed ``"CPU:0"`` / ``"GPU:i"``; ``getCPU()`` / ``getGPU(i)`` return torch devices.  Contexts, flows and
tasks of the reference (cudaMalloc wrappers, named streams, task handles) map onto torch's caching allocator and CUDA
streams/events: ``get_flow(device, name)`` hands out one named ``torch.cuda.Stream`` per device."""
from __future__ import annotations

import enum
from typing import Dict, List

import torch


class DeviceType(enum.Enum):
    CPU = "CPU"
    GPU = "GPU"


class DeviceManager:
    _instance = None

    def __init__(self):
        self.devices: Dict[str, torch.device] = {}
        self._flows: Dict[tuple, "torch.cuda.Stream"] = {}
        self.default_device_id = "CPU:0"
        self.discover_devices()

    @classmethod
    def get_instance(cls) -> "DeviceManager":
        if cls._instance is None:
            cls._instance = DeviceManager()
        return cls._instance

    def discover_devices(self):
        self.devices = {"CPU:0": torch.device("cpu")}
        if torch.cuda.is_available():
            for i in range(torch.cuda.device_count()):
                self.devices[f"GPU:{i}"] = torch.device("cuda", i)

    def get_device(self, key) -> torch.device:
        if isinstance(key, DeviceType):
            # deterministic: the lowest index of that type (the reference returns an arbitrary map entry)
            for k in sorted(self.devices):
                if k.startswith(key.value):
                    return self.devices[k]
            raise RuntimeError("No device of the specified type found")
        if key not in self.devices:
            raise RuntimeError("Device with the given ID not found")
        return self.devices[key]

    def get_available_device_ids(self) -> List[str]:
        return sorted(self.devices)

    def has_device(self, key: str) -> bool:
        return key in self.devices

    def set_default_device(self, key):
        if isinstance(key, DeviceType):
            key = next(k for k in sorted(self.devices) if k.startswith(key.value))
        if key not in self.devices:
            raise RuntimeError("Device with the given ID not found")
        self.default_device_id = key

    def get_flow(self, device: torch.device, name: str = "default"):
        if device.type != "cuda":
            return None
        k = (device.index, name)
        if k not in self._flows:
            self._flows[k] = torch.cuda.current_stream(device) if name == "default" else torch.cuda.Stream(device)
        return self._flows[k]


def initializeDefaultDevices():
    DeviceManager.get_instance().discover_devices()


def getCPU() -> torch.device:
    return torch.device("cpu")


def getGPU(index: int = 0) -> torch.device:
    m = DeviceManager.get_instance()
    key = f"GPU:{index}"
    if not m.has_device(key):
        raise RuntimeError("Requested GPU index not found")
    return m.get_device(key)
