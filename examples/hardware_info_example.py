#!/usr/bin/env python
"""Print CPU / memory / NUMA / GPU information and the thread-affinity view (examples/hardware_info_example.cpp)."""
from _common import sys  # noqa: F401

from dcnn_b200.utils.device import DeviceManager, DeviceType, initializeDefaultDevices
from dcnn_b200.utils.hardware import HardwareInfo, ThreadAffinity

if __name__ == "__main__":
    hw = HardwareInfo()
    hw.initialize()
    hw.print_info()
    ThreadAffinity(hw).print_affinity_info()
    initializeDefaultDevices()
    m = DeviceManager.get_instance()
    print("Available devices:", m.get_available_device_ids())
    print("Default CPU device:", m.get_device(DeviceType.CPU))
