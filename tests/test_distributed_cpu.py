"""world_size-2 gloo tests of the N > 1 path: contiguous frequency shards, one
all-reduce of [loss, grad]; the per-shard evaluator is the CPU oracle here, the CUDA path
on the GPU box (tests/test_gpu_parity.py::test_two_shards_equal_one)."""
import os
import socket

import numpy as np
import pytest

from conftest import PARAMS, get_case


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, freqs, ref, queue):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from plate_inverse_problem_b200 import parallel
    case = get_case("tiny")
    F = freqs.size

    def local(lo, hi):
        L, g, _ = case.op.loss_and_grad(freqs[lo:hi], PARAMS, ref[lo:hi], "MSE")
        return L * (hi - lo) / F, g * 1.0      # oracle means over its own shard; rescale to the sweep

    # the oracle's gradient is of the shard-mean loss: rescale it the same way
    def local_scaled(lo, hi):
        L, g = local(lo, hi)
        return L, g * (hi - lo) / F

    val, grad = parallel.sharded_value_and_grad(local_scaled, F, PARAMS.size)
    fr = parallel.sharded_gather(lambda lo, hi: case.op.forward(freqs[lo:hi], PARAMS), F)
    lo, hi = parallel.shard_range(F)
    # the product's own sharded sweeps (Problem.getFRFunction, SymmetricProblem.getFRFunction: complex FRF as an
    # all-reduce of interleaved doubles) with the native call replaced by the oracle
    from types import SimpleNamespace
    from plate_inverse_problem_b200.Problem import Problem
    from plate_inverse_problem_b200.ProblemSymm import SymmetricProblem
    calls = []

    def fake_forward(fs, params, return_u=False):
        calls.append(len(fs))
        y = np.stack([case.op.forward(fs, params) * (1.0 + 0.5j)], axis=1)       # [n, 1] complex
        return np.abs(y[:, 0]), y
    for cls in (Problem, SymmetricProblem):
        prob = object.__new__(cls)
        prob._st = SimpleNamespace(device=0)
        prob._forward_raw = fake_forward
        out = prob.getFRFunction()(freqs, PARAMS)
        want = case.op.forward(freqs, PARAMS) * (1.0 + 0.5j)
        assert calls[-1] == hi - lo                       # this rank solved only its own block
        assert np.allclose(out, np.abs(want) if cls is Problem else want, rtol=1e-12)
        assert out.dtype == (np.float64 if cls is Problem else np.complex128)
    queue.put((rank, val, grad, fr, lo, hi))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_sharding_matches_single_process():
    import torch.multiprocessing as mp
    case = get_case("tiny")
    freqs = np.linspace(40.0, 600.0, 7)          # odd count: ragged shards
    ref = case.op.forward(freqs, PARAMS) * 1.1
    L1, g1, _ = case.op.loss_and_grad(freqs, PARAMS, ref, "MSE")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, freqs, ref, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    assert (res[0][4], res[0][5], res[1][4], res[1][5]) == (0, 3, 3, 7)
    for _, val, grad, fr, _, _ in res:
        assert abs(val - L1) <= 1e-12 * abs(L1)
        assert np.max(np.abs(grad - g1)) <= 1e-10 * np.max(np.abs(g1))
        assert np.allclose(fr, case.op.forward(freqs, PARAMS), rtol=1e-12)


def test_shard_range_covers_everything():
    from plate_inverse_problem_b200.parallel import shard_range
    for n in (0, 1, 7, 400, 2001):
        for w in (1, 2, 4, 8):
            blocks = [shard_range(n, r, w) for r in range(w)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
