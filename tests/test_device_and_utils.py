"""DeviceManager / hardware / misc utility tests (the reference's device_manager_test.cpp kind) plus env parsing."""
import os

import pytest
import torch

from dcnn_b200.utils import env
from dcnn_b200.utils.device import DeviceManager, DeviceType, getCPU, initializeDefaultDevices
from dcnn_b200.utils.misc import MetricsLogger, benchmark, get_memory_usage_kb


def test_device_manager_discovers_cpu_and_is_a_singleton():
    initializeDefaultDevices()
    dm = DeviceManager.get_instance()
    assert dm is DeviceManager.get_instance()
    ids = dm.get_available_device_ids()
    assert "CPU:0" in ids and dm.has_device("CPU:0") and not dm.has_device("GPU:99")
    assert dm.get_device("CPU:0") == torch.device("cpu") == getCPU()
    n_gpu = sum(1 for i in ids if i.startswith("GPU:"))
    assert n_gpu == (torch.cuda.device_count() if torch.cuda.is_available() else 0)
    with pytest.raises(Exception):
        dm.get_device("GPU:99")
    assert DeviceType.CPU.name == "CPU" and DeviceType.GPU.name == "GPU"
    flow = dm.get_flow(torch.device("cpu"), "default")
    assert flow is None or flow is not False              # CPU flows are no-ops


def test_env_file_parsing_types_and_precedence(tmp_path):
    f = tmp_path / "cfg.env"
    f.write_text('# comment\nEPOCHS=7\nLR_INITIAL=0.25\nNAME="quoted value"\nFLAG=true\nBROKEN LINE\nEMPTY=\n')
    env.clear_loaded()
    try:
        assert env.load_env_file(str(f))
        assert env.get_env("EPOCHS", 1) == 7 and env.get_env("LR_INITIAL", 0.0) == 0.25
        assert env.get_env("NAME", "") == "quoted value" and env.get_env("FLAG", False) is True
        assert env.get_env("MISSING_KEY_XYZ", 42) == 42
        assert env.get_env("NAME", 3) == 3                 # unparsable as int -> default
        env.override("EPOCHS", 9)
        assert env.get_env("EPOCHS", 1) == 9 and os.environ["EPOCHS"] == "9"
        assert not env.load_env_file(str(tmp_path / "nope.env"))
    finally:
        env.clear_loaded()
        for k in ("EPOCHS", "LR_INITIAL", "NAME", "FLAG", "EMPTY"):
            os.environ.pop(k, None)


def test_misc_helpers(tmp_path, capsys):
    assert get_memory_usage_kb() > 0
    calls = []
    ms = benchmark("noop", lambda: calls.append(1), runs=3)
    assert len(calls) == 4 and ms >= 0 and "noop" in capsys.readouterr().out
    log = MetricsLogger(str(tmp_path / "a" / "m.jsonl"), rank=2)
    rec = log.log(step=1, loss=0.5)
    assert rec["rank"] == 2 and rec["step"] == 1 and (tmp_path / "a" / "m.jsonl").read_text().count("\n") == 1
    assert MetricsLogger(None).log(x=1)["x"] == 1         # no file: still returns the record


def test_hardware_info_topology_dynamic_and_affinity():
    import json
    import time
    from dcnn_b200.utils.hardware import HardwareInfo, ThreadAffinity, _parse_cpulist
    hw = HardwareInfo()
    assert hw.initialize()
    t = hw.info["topology"]
    assert t["logical"] >= t["physical"] >= 1 and t["sockets"] >= 1 and t["threads_per_core"] >= 1
    assert sorted(hw.get_performance_cores() + hw.get_efficiency_cores()) == sorted(t["per_cpu"]) or not t["per_cpu"]
    assert hw.get_usable_cores() >= 1 and "MemTotal" in hw.info["memory"]
    time.sleep(0.05)
    dyn = hw.update_dynamic_info()
    assert 0.0 <= dyn["cpu_utilization_percent"] <= 100.0 and 0.0 <= dyn["memory_used_percent"] <= 100.0
    blob = json.loads(hw.to_json())
    assert "per_cpu" not in blob["topology"] and "dynamic" in blob
    assert _parse_cpulist("0-2,5, 7-8") == [0, 1, 2, 5, 7, 8]
    if hasattr(os, "sched_getaffinity"):
        before = os.sched_getaffinity(0)
        try:
            ta = ThreadAffinity(hw)
            assert ta.set_efficiency_cores(1) and len(os.sched_getaffinity(0)) == 1
            os.sched_setaffinity(0, before)
            assert ta.set_performance_cores() and os.sched_getaffinity(0) <= before
            assert not ta.set_cores([])
        finally:
            os.sched_setaffinity(0, before)


def test_mem_pool_best_fit_and_temp_buffer():
    """ref include/nn/mem_pool.hpp: free blocks keyed by capacity, smallest sufficient block wins, TempBuffer returns
    its block on scope exit."""
    import torch
    from dcnn_b200.nn import MemPool, TempBuffer
    pool = MemPool()
    a, b = pool.get([4, 8]), pool.get([100])
    assert a.shape == (4, 8) and b.shape == (100,) and pool.misses == 2
    pa, pb = a.data_ptr(), b.data_ptr()
    pool.release(a)
    pool.release(b)
    assert pool.free_blocks() == 2 and pool.free_bytes() == (32 + 100) * 4
    c = pool.get([5, 5])                  # 25 elements: the 32-element block, not the 100-element one
    assert c.data_ptr() == pa and c.shape == (5, 5) and pool.hits == 1
    d = pool.get([64])
    assert d.data_ptr() == pb
    e = pool.get([1000])                  # nothing fits any more
    assert pool.misses == 3 and e.numel() == 1000
    pool.release(c)
    with TempBuffer(pool, [2, 3]) as t:
        assert t.data_ptr() == pa and pool.free_blocks() == 0
        t.fill_(1.0)
    assert pool.free_blocks() == 1
    buf = TempBuffer(pool, [8])
    buf.release()
    buf.release()                         # idempotent
    assert pool.free_blocks() == 1
    pool.release(torch.zeros(3, 3))       # foreign contiguous tensors of the pool's dtype are accepted
    assert pool.free_blocks() == 2
    pool.clear()
    assert pool.free_blocks() == 0


def test_matrix_container():
    """ref include/matrix/matrix.hpp: value semantics, element-wise operators, matrix product, shape errors."""
    import pytest
    import torch
    from dcnn_b200.utils.matrix import Matrix
    a = Matrix(2, 3, [1, 2, 3, 4, 5, 6])
    b = Matrix(2, 3)
    b.fill(1.0)
    c = a + b
    assert c(1, 2) == 7.0 and a(1, 2) == 6.0 and (c - b).data().equal(a.data())
    c -= b
    c += b
    c *= 2.0
    c /= 4.0
    assert torch.allclose(c.data(), (a.data() + 1) / 2)
    p = a * Matrix(3, 2, [1, 0, 0, 1, 1, 1])
    assert p.rows() == 2 and p.cols() == 2 and torch.allclose(p.data(), torch.tensor([[4., 5.], [10., 11.]]))
    assert torch.allclose((a * 2.0).data(), (a / 0.5).data())
    r = a.reshape(3, 2)
    r[0, 0] = 9.0
    assert a(0, 0) == 1.0 and r.size() == 6          # reshape and clone copy
    with pytest.raises(ValueError):
        a + Matrix(3, 2)
    with pytest.raises(ValueError):
        a * Matrix(2, 2)
    with pytest.raises(ZeroDivisionError):
        a / 0.0
    m = Matrix(4, 4)
    m.fill_random_uniform(0.5, seed=1)
    assert float(m.data().abs().max()) <= 0.5
    m.resize(2, 8)
    assert m.rows() == 2 and float(m.data().abs().sum()) == 0.0


def test_reference_style_device_and_hardware_accessors():
    """The reference's spellings: DeviceManager::getInstance / getDevice / addDevice ..., Device::getName /
    getTotalMemory / allocateAlignedMemory / copyToDevice / getFlow, HardwareInfo getters, ThreadAffinity counts."""
    import torch
    from dcnn_b200.utils.device import Device, DeviceManager, DeviceType
    from dcnn_b200.utils.hardware import HardwareInfo, ThreadAffinity
    dm = DeviceManager.getInstance()
    dm.discoverDevices()
    assert dm.hasDevice("CPU:0") and "CPU:0" in dm.getAvailableDeviceIDs()
    cpu = dm.getDevice("CPU:0")
    assert isinstance(cpu, Device) and cpu.getID() == 0 and cpu.device_type() == DeviceType.CPU and cpu.getName()
    assert cpu == torch.device("cpu") and cpu == dm.getDevice(DeviceType.CPU)
    assert cpu.getTotalMemory() >= cpu.getAvailableMemory() > 0
    buf = cpu.allocateAlignedMemory(1000, 256)
    assert buf.numel() == 1000 and buf.data_ptr() % 256 == 0
    src = torch.arange(100, dtype=torch.uint8)
    cpu.copyToDevice(buf, src, 100)
    back = torch.zeros(100, dtype=torch.uint8)
    cpu.copyToHost(back, buf, 100)
    assert torch.equal(back, src) and cpu.getFlow("default") is None
    cpu.deallocateAlignedMemory(buf)
    dm.addDevice(torch.device("cpu"), "CPU:7")
    assert dm.hasDevice("CPU:7")
    dm.setDefaultDevice("CPU:7")
    dm.removeDevice("CPU:7")
    assert not dm.hasDevice("CPU:7") and dm.default_device_id == "CPU:0"
    dm.clearDevices()
    assert dm.getAvailableDeviceIDs() == []
    dm.discoverDevices()

    hw = HardwareInfo()
    assert not hw.is_initialized()
    assert hw.get_logical_cores() >= hw.get_physical_cores() >= 1 and hw.is_initialized()
    assert hw.get_sockets() >= 1 and hw.get_model_name() and hw.get_architecture()
    assert hw.get_total_memory_bytes() >= hw.get_used_memory_bytes() >= 0
    assert 0.0 <= hw.get_memory_utilization_percent() <= 100.0
    assert hw.get_cache_line_size() in (32, 64, 128, 256)
    assert isinstance(hw.supports_avx2(), bool) and (not hw.supports_avx512() or hw.supports_avx2())
    assert hw.get_max_frequency_mhz() >= 0 and hw.get_optimal_thread_count() >= 1
    rec = hw.get_recommended_cpu_affinity()
    assert len(rec) == len(set(rec)) and 1 <= len(rec) <= hw.get_logical_cores()
    assert len(hw.get_recommended_cpu_affinity(1)) == 1
    assert 0.0 <= hw.update_utilization() <= 100.0 and 0.0 <= hw.get_overall_utilization() <= 100.0
    assert isinstance(hw.is_containerized(), bool) and isinstance(hw.is_virtualized(), bool)
    ta = ThreadAffinity(hw)
    assert ta.get_performance_core_count() + ta.get_efficiency_core_count() >= 1
    assert ta.has_efficiency_cores() == (ta.get_efficiency_core_count() > 0)
    import os
    before = os.sched_getaffinity(0)
    assert ta.set_current_thread_affinity(sorted(before))
