#!/usr/bin/env python
"""Device-memory demo -- counterpart of the reference's examples/device_ptr_example.cu (`device_ptr<T[]>`: owning pointer
with device, size, capacity and grow-only ensure()).  Here device memory is a torch tensor on a DeviceManager device;
the demo shows allocation, grow-only reuse, host<->device copies from pinned memory, a hand-written kernel running on the
buffer through the C ABI (dcnn_fill / dcnn_add), and -- with 2+ GPUs -- a raw cudaMalloc'd buffer of the kind the
data-parallel path shares between processes over CUDA IPC."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200.utils.device import DeviceManager  # noqa: E402


class GrowOnlyBuffer:
    """ensure(n) semantics of device_ptr: capacity never shrinks, contents are not preserved on growth."""

    def __init__(self, device):
        self.device, self.t = device, torch.empty(0, device=device)

    def ensure(self, n):
        if self.t.numel() < n:
            self.t = torch.empty(n, device=self.device)
        return self.t[:n]

    @property
    def capacity(self):
        return self.t.numel()


def main():
    dm = DeviceManager.get_instance()
    print("devices:", dm.get_available_device_ids())
    dev = dm.get_device("GPU:0") if dm.has_device("GPU:0") else dm.get_device("CPU:0")
    buf = GrowOnlyBuffer(dev)
    a = buf.ensure(1 << 20)
    print(f"ensure(1M): capacity {buf.capacity}")
    buf.ensure(1 << 10)
    print(f"ensure(1K): capacity {buf.capacity} (grow-only)")
    host = torch.arange(1 << 20, dtype=torch.float32)
    if dev.type == "cuda":
        host = host.pin_memory()
    a.copy_(host, non_blocking=True)
    if dev.type == "cuda":
        from dcnn_b200._lib import F, LL, P, check, lib
        b = torch.empty_like(a)
        s = P(torch.cuda.current_stream().cuda_stream)
        check(lib().dcnn_fill(P(b.data_ptr()), LL(b.numel()), F(2.0), s), "dcnn_fill")
        out = torch.empty_like(a)
        check(lib().dcnn_add(P(a.data_ptr()), P(b.data_ptr()), P(out.data_ptr()), LL(a.numel()), s), "dcnn_add")
        torch.cuda.synchronize()
        assert torch.equal(out.cpu(), host + 2.0)
        print("dcnn_fill + dcnn_add on the buffer: OK")
        import ctypes
        ptr = ctypes.c_void_p()
        check(lib().dcnn_dev_malloc(ctypes.byref(ptr), LL(1 << 20)), "dcnn_dev_malloc")
        print(f"cudaMalloc'd symmetric-buffer candidate at 0x{ptr.value:x} (exported to peers with dcnn_ipc_get_handle)")
        check(lib().dcnn_dev_free(ptr), "dcnn_dev_free")
    else:
        assert torch.equal(a, host)
        print("CPU buffer round trip: OK")


if __name__ == "__main__":
    main()
