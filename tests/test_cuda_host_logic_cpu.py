"""Host-side decisions of dcnn_b200/ops/cuda.py that need no GPU: the one-entry cache of packed stem activations and
the routing of reduction-bound Dense products."""
import torch

from dcnn_b200.ops import cuda as K


def test_packed_stem_cache_hits_once_for_the_same_unmodified_tensor(monkeypatch):
    calls = []
    monkeypatch.setattr(K, "_pack_taps", lambda x, *a: (calls.append(x), torch.zeros(1))[1])
    monkeypatch.setattr(K, "_PACK_CACHE", None)
    geom = (3, 3, (1, 1), (1, 1), 8, 8, 32)
    x = torch.randn(2, 3, 8, 8)
    fwd = K._pack_taps_cached(x, *geom, store=True)            # the layer's forward call packs and remembers
    assert K._pack_taps_cached(x, *geom, store=False) is fwd   # its weight gradient reuses ...
    assert len(calls) == 1
    assert K._pack_taps_cached(x, *geom, store=False) is not fwd   # ... once: the entry is dropped after use
    assert len(calls) == 2

    K._pack_taps_cached(x, *geom, store=True)
    other = torch.randn(2, 3, 8, 8)                            # another tensor object (even at a recycled address)
    n = len(calls)
    K._pack_taps_cached(other, *geom, store=False)
    assert len(calls) == n + 1

    K._pack_taps_cached(x, *geom, store=True)
    x.add_(1.0)                                                # modified in place since the forward call
    n = len(calls)
    K._pack_taps_cached(x, *geom, store=False)
    assert len(calls) == n + 1

    K._pack_taps_cached(x, *geom, store=True)
    n = len(calls)
    K._pack_taps_cached(x, 5, 5, (1, 1), (2, 2), 8, 8, 96, store=False)   # same tensor, another layer geometry
    assert len(calls) == n + 1


def test_only_degenerate_dense_products_take_the_split_k_route():
    assert K._reduction_bound(64, 128, 2097152)          # the reference's gemm_benchmark shape
    assert K._reduction_bound(8, 12, 65536)
    assert not K._reduction_bound(256, 200, 2048)        # the ResNet-50 classifier
    assert not K._reduction_bound(4096, 2048, 65536)     # many output tiles: the forward kernel fills the machine
    assert not K._reduction_bound(6, 12, 65536)          # rows not a multiple of 4: the wgrad kernel cannot take it
