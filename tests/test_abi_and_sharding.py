"""CPU-side checks: the C-ABI library exports every symbol the header declares, the product has no CPU
path, and the batch-sharding host logic works across 2 ranks (gloo)."""
import ctypes
import os
import re

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "soda_b200.h")).read()
    return re.findall(r"SODA_API\s+[\w\s\*]+?\b(soda_\w+)\s*\(", src)


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 32 and "soda_conv_lif_fwd" in names and "soda_bin_events" in names
    lib = ctypes.CDLL(os.path.join(ROOT, "snn_for_object_detection_b200", "libsoda_b200.so"))
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/soda_b200.h but not exported"
    lib.soda_abi_version.restype = ctypes.c_int
    assert lib.soda_abi_version() == 2


def test_ctypes_binding_covers_the_header():
    from snn_for_object_detection_b200 import _lib

    for n in _declared_symbols():
        assert hasattr(_lib.lib, n)


def test_cpu_tensors_are_rejected_everywhere():
    """No CPU fallback: every op refuses a CPU tensor instead of computing on the host."""
    import snn_for_object_detection_b200 as pkg

    x = torch.zeros(2, 32, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.lif_scan(x, 2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.pool2d(x, 2, 2, "S")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.pack_conv(torch.zeros(8, 32, 3, 3), 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.bin_events(torch.zeros(16, dtype=torch.uint8), torch.zeros(2, dtype=torch.int64),
                           torch.zeros(1, dtype=torch.int64), 16000, 2, 8, 8)


def test_product_never_imports_the_oracle():
    pkg_dir = os.path.join(ROOT, "snn_for_object_detection_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


def test_shard_range_partitions_exactly():
    from snn_for_object_detection_b200.sharding import shard_range

    for n in (0, 1, 5, 8, 40, 1023):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from snn_for_object_detection_b200 import sharding

    n_total, A = 5, 7
    samples = [torch.full((3,), float(i)) for i in range(n_total)]
    mine = sharding.shard_samples(samples)
    lo, hi = sharding.shard_range(n_total, rank, world)
    assert [int(s[0]) for s in mine] == list(range(lo, hi))
    # per-rank "predictions" = f(sample id); gathered result must be in batch order on every rank
    cls = torch.stack([torch.full((A, 3), float(i)) for i in range(lo, hi)])
    bbox = torch.stack([torch.full((A, 4), float(10 * i)) for i in range(lo, hi)])
    g_cls, g_bbox = sharding.gather_predictions(cls, bbox, n_total)
    ok = g_cls.shape == (n_total, A, 3) and all(float(g_cls[i, 0, 0]) == i and float(g_bbox[i, 0, 0]) == 10 * i
                                                 for i in range(n_total))
    # gradient all-reduce: mean over ranks, written back in place
    p1, p2 = torch.nn.Parameter(torch.zeros(3)), torch.nn.Parameter(torch.zeros(2, 2))
    p1.grad, p2.grad = torch.full((3,), float(rank + 1)), torch.full((2, 2), float(10 * (rank + 1)))
    sharding.allreduce_flat_gradients([p1, p2])
    mean = sum(range(1, world + 1)) / world
    ok = ok and torch.allclose(p1.grad, torch.full((3,), mean)) and torch.allclose(p2.grad, torch.full((2, 2), 10 * mean))
    q.put((rank, bool(ok)))
    dist.barrier()
    dist.destroy_process_group()


def test_sharding_two_ranks_gloo():
    world, port = 2, 29000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_spike_count_bounds_pick_the_conv_operand_scheme():
    """Host logic of the exact-input tags (ops.mark_exact): bounds add over Residual sums and 2x2 sum pools, and the
    conv flags follow them - <= 256 (8 significant bits): bf16-only scheme; <= 2048: tf32 without the lo(a) half."""
    from snn_for_object_detection_b200 import _lib, ops

    a, b, c = torch.zeros(1), torch.zeros(1), torch.zeros(1)
    assert ops.exact_bound(a) is None and ops._exact_flags(a) == 0
    ops.mark_exact(a)
    ops.mark_exact(b, 4)
    assert ops.exact_bound(a) == 1 and ops.sum_bound(a, b) == 5 and ops.sum_bound(a, c) is None
    both = _lib.SODA_CONV_FLAG_INPUT_EXACT_TF32 | _lib.SODA_CONV_FLAG_INPUT_EXACT_BF16
    assert ops._exact_flags(a) == both and ops._exact_flags(ops.mark_exact(c, 256)) == both
    assert ops._exact_flags(ops.mark_exact(c, 257)) == _lib.SODA_CONV_FLAG_INPUT_EXACT_TF32
    assert ops._exact_flags(ops.mark_exact(c, 4096)) == 0 and not ops.is_exact(c)
    assert ops.exact_bound(ops.mark_exact(torch.zeros(1), None)) is None


def test_convert_sync_batchnorm_marks_every_batchnorm():
    """Host side of the SyncBatchNorm exchange (reference config.yaml:76): every generated BatchNorm is switched."""
    from snn_for_object_detection_b200 import BlockGen, sharding
    from snn_for_object_detection_b200.modules.common import SodaBatchNorm2d
    from snn_for_object_detection_b200.modules.layer_gen import LIF, Conv, Norm

    block = BlockGen(2, [Conv(8), Norm(), LIF(), Conv(8), Norm(), LIF()])
    norms = [m for m in block.modules() if isinstance(m, SodaBatchNorm2d)]
    assert len(norms) == 2 and not any(getattr(m, "sync", False) for m in norms)
    assert sharding.convert_sync_batchnorm(block) is block
    assert all(m.sync for m in norms)
