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


class Device:
    """The reference's ``Device`` surface (include/device/device.hpp:10-40) over a ``torch.device``: identity, memory
    figures, raw allocations (returned as byte tensors -- their ``data_ptr()`` is the address), host<->device copies and
    named flows.  Everything inside the library works with ``torch.device`` directly (``.torch_device``)."""

    def __init__(self, device_type: DeviceType, index: int, manager: "DeviceManager"):
        self._type, self._id, self._manager = device_type, int(index), manager
        self.torch_device = torch.device("cpu") if device_type == DeviceType.CPU else torch.device("cuda", self._id)

    def device_type(self) -> DeviceType:
        return self._type

    def getID(self) -> int:
        return self._id

    def getName(self) -> str:
        if self._type == DeviceType.GPU:
            return torch.cuda.get_device_name(self._id)
        import platform
        try:
            with open("/proc/cpuinfo") as f:
                for line in f:
                    if line.startswith("model name"):
                        return line.split(":", 1)[1].strip()
        except OSError:
            pass
        return platform.processor() or "CPU"

    def getTotalMemory(self) -> int:
        if self._type == DeviceType.GPU:
            return int(torch.cuda.get_device_properties(self._id).total_memory)
        return _meminfo().get("MemTotal", 0)

    def getAvailableMemory(self) -> int:
        if self._type == DeviceType.GPU:
            return int(torch.cuda.mem_get_info(self._id)[0])
        return _meminfo().get("MemAvailable", 0)

    def allocateMemory(self, size: int) -> torch.Tensor:
        return torch.empty(int(size), dtype=torch.uint8, device=self.torch_device)

    def allocateAlignedMemory(self, size: int, alignment: int = 64) -> torch.Tensor:
        """A byte tensor whose address is a multiple of ``alignment`` (over-allocate and slice)."""
        raw = torch.empty(int(size) + int(alignment), dtype=torch.uint8, device=self.torch_device)
        off = (-raw.data_ptr()) % int(alignment)
        return raw[off:off + int(size)]

    def deallocateMemory(self, buf):          # ownership is the tensor's: dropping the reference frees it
        del buf

    deallocateAlignedMemory = deallocateMemory

    def copyToDevice(self, dest: torch.Tensor, src: torch.Tensor, size: int = None):
        n = dest.numel() if size is None else int(size)
        dest.reshape(-1)[:n].copy_(src.reshape(-1)[:n], non_blocking=True)

    def copyToHost(self, dest: torch.Tensor, src: torch.Tensor, size: int = None):
        n = dest.numel() if size is None else int(size)
        dest.reshape(-1)[:n].copy_(src.reshape(-1)[:n])

    def createFlow(self, flow_id: str):
        self._manager.get_flow(self.torch_device, flow_id)

    def getFlow(self, flow_id: str = "default"):
        return self._manager.get_flow(self.torch_device, flow_id)

    def __eq__(self, other):
        if isinstance(other, Device):
            return self.torch_device == other.torch_device
        return self.torch_device == other

    def __hash__(self):
        return hash(self.torch_device)

    def __repr__(self):
        return f"Device({self._type.value}:{self._id})"


def _meminfo():
    out = {}
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                k, _, v = line.partition(":")
                if k in ("MemTotal", "MemAvailable"):
                    out[k] = int(v.split()[0]) * 1024
    except OSError:
        pass
    return out


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


    # -- the reference's spellings (include/device/device_manager.hpp:11-67) -------------------------------------
    @classmethod
    def getInstance(cls) -> "DeviceManager":
        return cls.get_instance()

    def discoverDevices(self):
        self.discover_devices()

    def getDevice(self, key) -> Device:
        """A ``Device`` wrapper (``get_device`` returns the bare ``torch.device``)."""
        td = self.get_device(key)
        return Device(DeviceType.CPU if td.type == "cpu" else DeviceType.GPU, td.index or 0, self)

    def getAvailableDeviceIDs(self) -> List[str]:
        return self.get_available_device_ids()

    def hasDevice(self, key: str) -> bool:
        return self.has_device(key)

    def setDefaultDevice(self, key):
        self.set_default_device(key)

    def addDevice(self, device, key: str = None):
        """Register an extra device (a ``Device`` or ``torch.device``) under ``key`` (default ``TYPE:index``)."""
        td = device.torch_device if isinstance(device, Device) else torch.device(device)
        key = key or (f"GPU:{td.index or 0}" if td.type == "cuda" else f"CPU:{td.index or 0}")
        self.devices[key] = td

    def removeDevice(self, key: str):
        self.devices.pop(key, None)
        if self.default_device_id == key:
            self.default_device_id = "CPU:0"

    def clearDevices(self):
        self.devices.clear()
        self._flows.clear()


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
