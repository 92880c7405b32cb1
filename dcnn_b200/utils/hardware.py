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
        self.info = {"cpu": self._cpu(), "topology": self._topology(), "caches": self._caches(),
                     "memory": self._memory(), "numa": self._numa(), "cgroup": self._cgroup(), "gpus": self._gpus()}
        self._prev_stat = self._read_proc_stat()
        self.initialized = True
        return True

    # ---- CPU topology: sockets / physical cores / SMT siblings / per-core max frequency (P- vs E-core classes) ----------
    @staticmethod
    def _read(path, default=None):
        try:
            with open(path) as f:
                return f.read().strip()
        except OSError:
            return default

    @classmethod
    def _topology(cls):
        base = "/sys/devices/system/cpu"
        cores = {}
        for d in sorted(os.listdir(base)) if os.path.isdir(base) else []:
            if not (d.startswith("cpu") and d[3:].isdigit()):
                continue
            cpu = int(d[3:])
            t = os.path.join(base, d, "topology")
            pkg = cls._read(os.path.join(t, "physical_package_id"))
            core = cls._read(os.path.join(t, "core_id"))
            sib = cls._read(os.path.join(t, "thread_siblings_list")) or str(cpu)
            fmax = cls._read(os.path.join(base, d, "cpufreq", "cpuinfo_max_freq"))
            cores[cpu] = {"package": int(pkg) if pkg not in (None, "") else 0,
                          "core": int(core) if core not in (None, "") else cpu,
                          "siblings": _parse_cpulist(sib), "max_khz": int(fmax) if fmax and fmax.isdigit() else 0}
        phys = {(c["package"], c["core"]) for c in cores.values()}
        fm = sorted({c["max_khz"] for c in cores.values() if c["max_khz"] > 0})
        # hybrid parts expose two frequency classes: the faster one is the performance class (ref hardware_info.cpp
        # classifies P-/E-cores the same way); on a uniform part every core is a performance core
        threshold = (fm[0] + fm[-1]) / 2 if len(fm) > 1 and fm[-1] > 1.15 * fm[0] else 0
        pcores = sorted(c for c, v in cores.items() if v["max_khz"] >= threshold)
        ecores = sorted(c for c, v in cores.items() if threshold and v["max_khz"] < threshold)
        return {"logical": len(cores) or (os.cpu_count() or 1), "physical": len(phys) or (os.cpu_count() or 1),
                "sockets": len({c["package"] for c in cores.values()}) or 1,
                "threads_per_core": max((len(c["siblings"]) for c in cores.values()), default=1),
                "performance_cores": pcores, "efficiency_cores": ecores, "per_cpu": cores}

    @classmethod
    def _caches(cls):
        out = {}
        base = "/sys/devices/system/cpu/cpu0/cache"
        for d in sorted(os.listdir(base)) if os.path.isdir(base) else []:
            if not d.startswith("index"):
                continue
            lvl, typ = cls._read(os.path.join(base, d, "level")), cls._read(os.path.join(base, d, "type"))
            size = cls._read(os.path.join(base, d, "size"))
            if lvl and size:
                out[f"L{lvl}{'i' if typ == 'Instruction' else 'd' if typ == 'Data' else ''}"] = size
        return out

    @classmethod
    def _cgroup(cls):
        """Container limits (cgroup v2): what the scheduler will actually give this process."""
        out = {}
        cpu_max = cls._read("/sys/fs/cgroup/cpu.max")
        if cpu_max and not cpu_max.startswith("max"):
            try:
                quota, period = cpu_max.split()
                out["cpu_limit_cores"] = int(quota) / int(period)
            except ValueError:
                pass
        mem_max = cls._read("/sys/fs/cgroup/memory.max")
        if mem_max and mem_max.isdigit():
            out["memory_limit_bytes"] = int(mem_max)
        return out

    # ---- dynamic state (ref update_dynamic_info / LoadTracker inputs) --------------------------------------------------
    @staticmethod
    def _read_proc_stat():
        try:
            with open("/proc/stat") as f:
                parts = f.readline().split()[1:]
            vals = [int(v) for v in parts]
            idle = vals[3] + (vals[4] if len(vals) > 4 else 0)
            return sum(vals), idle
        except (OSError, ValueError, IndexError):
            return 0, 0

    def update_dynamic_info(self) -> Dict[str, object]:
        """CPU utilisation since the previous call, load average, memory in use, thermal zones, per-GPU utilisation /
        memory / temperature / power / SM clock (NVML when available)."""
        if not self.initialized:
            self.initialize()
        total, idle = self._read_proc_stat()
        pt, pi = self._prev_stat
        self._prev_stat = (total, idle)
        util = 100.0 * (1.0 - (idle - pi) / (total - pt)) if total > pt else 0.0
        mem = self._memory()
        dyn = {"cpu_utilization_percent": max(0.0, min(100.0, util)),
               "load_average": list(os.getloadavg()) if hasattr(os, "getloadavg") else [],
               "memory_used_percent": 100.0 * (1 - mem.get("MemAvailable", 0) / mem["MemTotal"]) if mem.get("MemTotal") else 0.0,
               "thermal_zones_c": self._thermal(), "gpus": self._gpus_dynamic()}
        self.info["dynamic"] = dyn
        return dyn

    @classmethod
    def _thermal(cls):
        out = {}
        base = "/sys/class/thermal"
        for d in sorted(os.listdir(base)) if os.path.isdir(base) else []:
            if d.startswith("thermal_zone"):
                t, name = cls._read(os.path.join(base, d, "temp")), cls._read(os.path.join(base, d, "type"), d)
                if t and t.lstrip("-").isdigit():
                    out[f"{name}:{d[12:]}"] = int(t) / 1000.0
        return out

    @staticmethod
    def _gpus_dynamic():
        out = []
        try:
            import pynvml
            pynvml.nvmlInit()
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u, m = pynvml.nvmlDeviceGetUtilizationRates(h), pynvml.nvmlDeviceGetMemoryInfo(h)
                out.append({"index": i, "utilization_percent": u.gpu, "memory_used_bytes": m.used,
                            "temperature_c": pynvml.nvmlDeviceGetTemperature(h, pynvml.NVML_TEMPERATURE_GPU),
                            "power_w": pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                            "sm_mhz": pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)})
        except Exception:
            pass
        return out

    def get_performance_cores(self) -> List[int]:
        return list(self.info.get("topology", {}).get("performance_cores", []))

    def get_efficiency_cores(self) -> List[int]:
        return list(self.info.get("topology", {}).get("efficiency_cores", []))

    def to_json(self) -> str:
        import json
        if not self.initialized:
            self.initialize()
        slim = dict(self.info)
        slim["topology"] = {k: v for k, v in self.info["topology"].items() if k != "per_cpu"}
        return json.dumps(slim, indent=2, default=str)

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
        t = self.info["topology"]
        print(f"Topology: {t['sockets']} socket(s), {t['physical']} physical cores, {t['threads_per_core']} thread(s)/core; "
              f"{len(t['performance_cores'])} performance / {len(t['efficiency_cores'])} efficiency cores")
        if self.info["caches"]:
            print("Caches: " + ", ".join(f"{k} {v}" for k, v in self.info["caches"].items()))
        if self.info["cgroup"]:
            print(f"Container limits: {self.info['cgroup']}")
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

    def set_performance_cores(self) -> bool:
        """ref ThreadAffinity P-core mode: the fast class of a hybrid CPU (every core on a uniform part)."""
        return self.set_cores(self.hw.get_performance_cores())

    def set_efficiency_cores(self, max_cores: int = 0) -> bool:
        """ref `network_worker --ecore [--max-ecores N]`: keep a CPU-side worker off the performance cores.  On a part
        without an efficiency class this falls back to the LAST cores of the allowed set (away from core 0, where the
        launcher and interrupt handlers tend to live)."""
        cores = self.hw.get_efficiency_cores()
        if not cores:
            allowed = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else []
            cores = allowed[len(allowed) // 2:] or allowed
        if max_cores > 0:
            cores = cores[:max_cores]
        return self.set_cores(cores)

    def pin_to_gpu_node(self, gpu_index: int) -> bool:
        node = self.gpu_numa_node(gpu_index)
        if node is None or node not in self.hw.info["numa"]:
            return False
        return self.set_cores(_parse_cpulist(self.hw.info["numa"][node]))

    def print_affinity_info(self):
        if hasattr(os, "sched_getaffinity"):
            print(f"Process {os.getpid()} may run on cores: {sorted(os.sched_getaffinity(0))}")
