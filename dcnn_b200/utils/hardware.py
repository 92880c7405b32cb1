"""Hardware introspection and thread/NUMA affinity (ref include/utils/hardware_info.hpp:127-362,
thread_affinity.hpp:44,156).  The reference spends ~2700 lines on CPU topology because its hot path is the CPU; here the
hot path is the GPU, so this module keeps what a launcher needs: CPU model/cores/RAM/NUMA, per-GPU properties (torch +
NVML when present), and pinning each rank to the CPU cores of its GPU's NUMA node."""
from __future__ import annotations

import os
import platform
from typing import Dict, List, Optional


class HardwareInfo:
    def __init__(self):
        self.initialized = False
        self.info: Dict[str, object] = {}

    def initialize(self) -> bool:
        self.info = {"cpu": self._cpu(), "memory": self._memory(), "numa": self._numa(), "gpus": self._gpus()}
        self.initialized = True
        return True

    @staticmethod
    def _cpu():
        model, mhz, flags = "", 0.0, []
        try:
            with open("/proc/cpuinfo") as f:
                for line in f:
                    k, _, v = line.partition(":")
                    k, v = k.strip(), v.strip()
                    if k == "model name" and not model:
                        model = v
                    elif k == "cpu MHz" and not mhz:
                        mhz = float(v)
                    elif k == "flags" and not flags:
                        flags = v.split()
        except OSError:
            pass
        return {"vendor_model": model or platform.processor(), "logical_cores": os.cpu_count() or 1,
                "usable_cores": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count(),
                "mhz": mhz, "avx2": "avx2" in flags, "avx512f": "avx512f" in flags, "fma": "fma" in flags}

    @staticmethod
    def _memory():
        out = {}
        try:
            with open("/proc/meminfo") as f:
                for line in f:
                    k, _, v = line.partition(":")
                    if k in ("MemTotal", "MemAvailable", "MemFree"):
                        out[k] = int(v.split()[0]) * 1024
        except OSError:
            pass
        return out

    @staticmethod
    def _numa():
        nodes = {}
        base = "/sys/devices/system/node"
        if os.path.isdir(base):
            for d in sorted(os.listdir(base)):
                if d.startswith("node") and d[4:].isdigit():
                    try:
                        with open(os.path.join(base, d, "cpulist")) as f:
                            nodes[int(d[4:])] = f.read().strip()
                    except OSError:
                        pass
        return nodes

    @staticmethod
    def _gpus():
        out = []
        try:
            import torch
            for i in range(torch.cuda.device_count()):
                p = torch.cuda.get_device_properties(i)
                out.append({"index": i, "name": p.name, "sms": p.multi_processor_count,
                            "memory_bytes": p.total_memory, "cc": f"{p.major}.{p.minor}"})
        except Exception:
            pass
        return out

    def get_usable_cores(self) -> int:
        return int(self.info.get("cpu", {}).get("usable_cores", 1))

    def print_info(self):
        if not self.initialized:
            self.initialize()
        c, m = self.info["cpu"], self.info["memory"]
        print("=== Hardware Information ===")
        print(f"CPU: {c['vendor_model']}  logical cores {c['logical_cores']} (usable {c['usable_cores']}), "
              f"{c['mhz']:.0f} MHz, AVX2 {c['avx2']}, AVX-512 {c['avx512f']}")
        if m:
            print(f"RAM: total {m.get('MemTotal', 0) / 2**30:.1f} GiB, available {m.get('MemAvailable', 0) / 2**30:.1f} GiB")
        for n, cpus in self.info["numa"].items():
            print(f"NUMA node {n}: cpus {cpus}")
        for g in self.info["gpus"]:
            print(f"GPU {g['index']}: {g['name']}  {g['sms']} SMs  {g['memory_bytes'] / 2**30:.0f} GiB  sm_{g['cc']}")


def _parse_cpulist(s: str) -> List[int]:
    out = []
    for part in s.split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            out.extend(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return out


class ThreadAffinity:
    """Pin the calling process to a CPU set (the reference pins to P-/E-cores; here: to the NUMA node of the rank's GPU)."""

    def __init__(self, hw: Optional[HardwareInfo] = None):
        self.hw = hw or HardwareInfo()
        if not self.hw.initialized:
            self.hw.initialize()

    @staticmethod
    def gpu_numa_node(gpu_index: int) -> Optional[int]:
        try:
            import torch
            bus = torch.cuda.get_device_properties(gpu_index).pci_bus_id  # type: ignore[attr-defined]
        except Exception:
            return None
        p = f"/sys/bus/pci/devices/{str(bus).lower()}/numa_node"
        try:
            with open(p) as f:
                n = int(f.read().strip())
            return n if n >= 0 else None
        except OSError:
            return None

    def set_cores(self, cores: List[int]) -> bool:
        if not hasattr(os, "sched_setaffinity") or not cores:
            return False
        try:
            os.sched_setaffinity(0, set(cores) & os.sched_getaffinity(0) or os.sched_getaffinity(0))
            return True
        except OSError:
            return False

    def pin_to_gpu_node(self, gpu_index: int) -> bool:
        node = self.gpu_numa_node(gpu_index)
        if node is None or node not in self.hw.info["numa"]:
            return False
        return self.set_cores(_parse_cpulist(self.hw.info["numa"][node]))

    def print_affinity_info(self):
        if hasattr(os, "sched_getaffinity"):
            print(f"Process {os.getpid()} may run on cores: {sorted(os.sched_getaffinity(0))}")
