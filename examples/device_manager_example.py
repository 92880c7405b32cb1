#!/usr/bin/env python
"""Device manager walk-through -- counterpart of the reference's examples/device_manager_example.cpp: enumerate the
devices, show their memory, allocate / fill / check / free a buffer on the CPU, allocate 1 GB on the first GPU, move a
buffer across with the device's own copy calls and run work on a named flow (stream)."""
from _common import sys  # noqa: F401

import torch

from dcnn_b200.utils.device import DeviceManager, DeviceType, initializeDefaultDevices


def main():
    print("=== Device Manager Example ===")
    print("\nInitializing devices...")
    initializeDefaultDevices()
    manager = DeviceManager.getInstance()
    ids = manager.getAvailableDeviceIDs()
    print(f"\nFound {len(ids)} device(s):")
    print("Available devices:", ids)
    for key in ids:
        d = manager.getDevice(key)
        kind = "CPU" if d.device_type() == DeviceType.CPU else "GPU"
        print(f"  Device {key}: {d.getName()} (Type: {kind})")
        print(f"    Memory - Total: {d.getTotalMemory() // (1 << 20)} MB, Available: {d.getAvailableMemory() // (1 << 20)} MB")

    cpu = manager.getDevice("CPU:0")
    print("\nTesting allocation on CPU device...")
    size = 1 << 20
    buf = cpu.allocateMemory(size)
    print(f"  Successfully allocated {size} bytes")
    buf.fill_(0xAA)
    ok = int(buf[0]) == 0xAA and int(buf[-1]) == 0xAA
    print("  Memory access test " + ("passed" if ok else "failed"))
    aligned = cpu.allocateAlignedMemory(4096, 256)
    print(f"  Aligned allocation: address % 256 = {aligned.data_ptr() % 256}")
    cpu.deallocateMemory(buf)
    print("  Successfully deallocated memory")

    gpus = [k for k in ids if manager.getDevice(k).device_type() == DeviceType.GPU]
    if not gpus:
        print("\nNo GPU devices available for testing")
    else:
        key = gpus[0]
        gpu = manager.getDevice(key)
        print(f"\nTesting allocation on GPU device {key}...")
        big = gpu.allocateMemory(1 << 30)
        print(f"  Successfully allocated {1 << 30} bytes on GPU")
        gpu.deallocateMemory(big)
        del big
        print("  Successfully deallocated GPU memory")
        # host -> device -> host through the device's copy calls, on a named flow
        src = torch.arange(1 << 16, dtype=torch.uint8).pin_memory() if torch.cuda.is_available() else None
        dev_buf, back = gpu.allocateMemory(1 << 16), torch.empty(1 << 16, dtype=torch.uint8)
        gpu.createFlow("copies")
        with torch.cuda.stream(gpu.getFlow("copies")):
            gpu.copyToDevice(dev_buf, src)
            dev_buf.add_(1)
        gpu.getFlow("copies").synchronize()
        gpu.copyToHost(back, dev_buf)
        print("  Round trip over flow 'copies': " + ("passed" if torch.equal(back, src + 1) else "failed"))
    print("\n=== Example completed successfully ===")


if __name__ == "__main__":
    main()
