#!/usr/bin/env python
"""Loopback throughput and round-trip latency of the native TCP transport (csrc/host/tcp_transport.cpp) through the
public ``TcpCommunicator`` API, i.e. including the message codec and the one copy into the received tensor.
The reference has no benchmark for its TcpCommunicator; the sizes are the pipeline's: control messages (a few bytes),
a ResNet stage boundary at microbatch 32 (32x256x16x16 floats = 8 MiB) and a large one (64 MiB)."""
import sys
import time

import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))

from dcnn_b200.parallel.pipeline import CommandType, Endpoint, Job, Message, TcpCommunicator  # noqa: E402


def main():
    a, b = TcpCommunicator("a", "127.0.0.1", 0, 2), TcpCommunicator("b", "127.0.0.1", 0, 2)
    a.connect("b", Endpoint.tcp("127.0.0.1", b.port))
    b.connect("a", Endpoint.tcp("127.0.0.1", a.port))
    # round trip of an empty control message
    n = 2000
    t0 = time.perf_counter()
    for _ in range(n):
        a.send_message(Message("b", CommandType.HEALTH_CHECK))
        b.wait_for_count(CommandType.HEALTH_CHECK, 1, 10.0)
        b.dequeue_message_by_type(CommandType.HEALTH_CHECK)
        b.send_message(Message("a", CommandType.HEALTH_CHECK))
        a.wait_for_count(CommandType.HEALTH_CHECK, 1, 10.0)
        a.dequeue_message_by_type(CommandType.HEALTH_CHECK)
    rtt = (time.perf_counter() - t0) / n
    print(f"control message round trip: {rtt * 1e6:.1f} us")
    for shape, reps in (((32, 256, 16, 16), 40), ((64, 256, 32, 32), 8)):
        x = torch.randn(*shape)
        mib = x.numel() * 4 / 2**20
        a.send_message(Message("b", CommandType.FORWARD_JOB, Job(x, 0)))      # warm the buffer pool
        b.wait_for_count(CommandType.FORWARD_JOB, 1, 60.0)
        b.dequeue_all_messages_by_type(CommandType.FORWARD_JOB)
        t0 = time.perf_counter()
        for i in range(reps):
            a.send_message(Message("b", CommandType.FORWARD_JOB, Job(x, i)))
        b.wait_for_count(CommandType.FORWARD_JOB, reps, 120.0)
        dt = time.perf_counter() - t0
        b.dequeue_all_messages_by_type(CommandType.FORWARD_JOB)
        print(f"{mib:6.1f} MiB jobs: {reps * mib / 1024 / dt:.2f} GiB/s ({dt / reps * 1e3:.2f} ms per job)")
    print("stats", a.stats())
    a.close()
    b.close()


if __name__ == "__main__":
    main()
