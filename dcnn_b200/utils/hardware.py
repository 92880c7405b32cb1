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

    # ---- the reference's accessors (include/utils/hardware_info.hpp) -------------------------------------------------
    def _need(self):
        if not self.initialized:
            self.initialize()
        return self.info

    def is_initialized(self) -> bool:
        return self.initialized

    def _cpuinfo_field(self, *names) -> str:
        try:
            with open("/proc/cpuinfo") as f:
                for line in f:
                    k, _, v = line.partition(":")
                    if k.strip() in names:
                        return v.strip()
        except OSError:
            pass
        return ""

    def get_vendor(self) -> str:
        return self._cpuinfo_field("vendor_id") or platform.machine()

    def get_model_name(self) -> str:
        return str(self._need()["cpu"]["vendor_model"])

    def get_architecture(self) -> str:
        return platform.machine()

    def get_family(self) -> int:
        v = self._cpuinfo_field("cpu family")
        return int(v) if v.isdigit() else 0

    def get_model(self) -> int:
        v = self._cpuinfo_field("model")
        return int(v) if v.isdigit() else 0

    def get_stepping(self) -> int:
        v = self._cpuinfo_field("stepping")
        return int(v) if v.isdigit() else 0

    def get_physical_cores(self) -> int:
        return int(self._need()["topology"]["physical"])

    def get_logical_cores(self) -> int:
        return int(self._need()["topology"]["logical"])

    def get_sockets(self) -> int:
        return int(self._need()["topology"]["sockets"])

    def supports_hyperthreading(self) -> bool:
        return int(self._need()["topology"]["threads_per_core"]) > 1

    def _flag(self, name: str) -> bool:
        return name in self._cpuinfo_field("flags").split()

    def supports_avx(self) -> bool:
        return self._flag("avx")

    def supports_avx2(self) -> bool:
        return bool(self._need()["cpu"]["avx2"])

    def supports_avx512(self) -> bool:
        return bool(self._need()["cpu"]["avx512f"])

    def supports_fma(self) -> bool:
        return bool(self._need()["cpu"]["fma"])

    def supports_sse4_2(self) -> bool:
        return self._flag("sse4_2")

    def supports_aes(self) -> bool:
        return self._flag("aes")

    def supports_sha(self) -> bool:
        return self._flag("sha_ni")

    def supports_bmi1(self) -> bool:
        return self._flag("bmi1")

    def supports_bmi2(self) -> bool:
        return self._flag("bmi2")

    def _khz(self, which: str) -> List[int]:
        per = self._need()["topology"]["per_cpu"]
        if which == "max":
            return [c["max_khz"] for c in per.values() if c["max_khz"] > 0]
        out = []
        for cpu in per:
            v = self._read(f"/sys/devices/system/cpu/cpu{cpu}/cpufreq/{which}")
            if v and v.isdigit():
                out.append(int(v))
        return out

    def get_max_frequency_mhz(self) -> float:
        v = self._khz("max")
        return max(v) / 1000.0 if v else float(self._need()["cpu"]["mhz"])

    def get_min_frequency_mhz(self) -> float:
        v = self._khz("cpuinfo_min_freq")
        return min(v) / 1000.0 if v else float(self._need()["cpu"]["mhz"])

    def get_base_frequency_mhz(self) -> float:
        v = self._khz("base_frequency")
        return max(v) / 1000.0 if v else float(self._need()["cpu"]["mhz"])

    def get_cache_line_size(self) -> int:
        v = self._read("/sys/devices/system/cpu/cpu0/cache/index0/coherency_line_size")
        return int(v) if v and v.isdigit() else 64

    def get_memory_hierarchy(self) -> Dict[str, object]:
        return dict(self._need()["caches"])

    def get_total_memory_bytes(self) -> int:
        return int(self._memory().get("MemTotal", 0))

    def get_available_memory_bytes(self) -> int:
        return int(self._memory().get("MemAvailable", 0))

    def get_used_memory_bytes(self) -> int:
        m = self._memory()
        return int(m.get("MemTotal", 0) - m.get("MemAvailable", 0))

    def get_memory_utilization_percent(self) -> float:
        m = self._memory()
        return 100.0 * (1.0 - m.get("MemAvailable", 0) / m["MemTotal"]) if m.get("MemTotal") else 0.0

    def get_load_averages(self) -> List[float]:
        return list(os.getloadavg()) if hasattr(os, "getloadavg") else []

    def update_utilization(self) -> float:
        return float(self.update_dynamic_info()["cpu_utilization_percent"])

    def get_overall_utilization(self) -> float:
        dyn = self.info.get("dynamic") or self.update_dynamic_info()
        return float(dyn["cpu_utilization_percent"])

    def get_thermal_info(self) -> Dict[str, float]:
        return self._thermal()

    def is_containerized(self) -> bool:
        if os.path.exists("/.dockerenv") or os.environ.get("KUBERNETES_SERVICE_HOST"):
            return True
        cg = self._read("/proc/1/cgroup", "") or ""
        return any(w in cg for w in ("docker", "kubepods", "containerd", "lxc")) or bool(self._need()["cgroup"])

    def is_virtualized(self) -> bool:
        return self._flag("hypervisor")

    def get_container_cpu_limit(self) -> float:
        """CPUs granted by the cgroup quota (0 = unlimited)."""
        return float(self._need()["cgroup"].get("cpu_limit_cores", 0.0) or 0.0)

    def get_optimal_thread_count(self) -> int:
        """Physical cores that this process may use (affinity mask and cgroup quota respected)."""
        usable = self.get_usable_cores()
        per_core = max(1, int(self._need()["topology"]["threads_per_core"]))
        n = max(1, usable // per_core if usable >= per_core else usable)
        limit = self.get_container_cpu_limit()
        return max(1, min(n, int(limit))) if limit >= 1 else n

    def get_recommended_cpu_affinity(self, num_threads: int = 0) -> List[int]:
        """One logical CPU per physical core, performance cores first."""
        topo = self._need()["topology"]
        allowed = set(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else set(topo["per_cpu"])
        seen, out = set(), []
        for cpu in topo["performance_cores"] + topo["efficiency_cores"]:
            c = topo["per_cpu"].get(cpu)
            if c is None or cpu not in allowed:
                continue
            key = (c["package"], c["core"])
            if key not in seen:
                seen.add(key)
                out.append(cpu)
        return out[:num_threads] if num_threads > 0 else out

    def is_suitable_for_heavy_workload(self) -> bool:
        return self.get_physical_cores() >= 4 and self.get_total_memory_bytes() >= 8 * 2**30

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

    def has_efficiency_cores(self) -> bool:
        return bool(self.hw.get_efficiency_cores())

    def get_performance_core_count(self) -> int:
        return len(self.hw.get_performance_cores())

    def get_efficiency_core_count(self) -> int:
        return len(self.hw.get_efficiency_cores())

    def set_current_thread_affinity(self, cores: List[int]) -> bool:
        """Pin the CALLING thread only (``set_cores`` pins the process)."""
        try:
            import threading
            os.sched_setaffinity(threading.get_native_id(), set(cores))
            return True
        except (AttributeError, OSError, ValueError):
            return False

    def pin_to_gpu_node(self, gpu_index: int) -> bool:
        node = self.gpu_numa_node(gpu_index)
        if node is None or node not in self.hw.info["numa"]:
            return False
        return self.set_cores(_parse_cpulist(self.hw.info["numa"][node]))

    def print_affinity_info(self):
        if hasattr(os, "sched_getaffinity"):
            print(f"Process {os.getpid()} may run on cores: {sorted(os.sched_getaffinity(0))}")
