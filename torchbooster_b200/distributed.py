"""Rank helpers, sampler sharding and the one cross-rank exchange of the hot path.

Drop-in for ``torchbooster/distributed.py`` (same names, arguments, defaults and error
behaviour: ``get_rank`` ``:24-28``, ``get_local_rank`` ``:31-38``, ``gather`` ``:41-56``,
``is_primary`` ``:58-60``, ``synchronize`` ``:63-68``, ``get_world_size`` ``:71-75``,
``data_sampler`` ``:78-98``, ``find_free_port`` ``:101-107``, ``launch`` ``:110-153``,
``job`` ``:156-205``): one process per GPU, NCCL over NVLink for the collective.

Addition: ``all_reduce_metrics`` -- the reference has no all_reduce at all; each
rank would otherwise only see its own shard's metrics.  The packed int64 metric buffer
(``MetricState.buffer``, 8 MB at 1000 classes) and the float64 loss pair are summed with one
``all_reduce`` each; on an NVSwitch box this is latency-bound, so it is issued once per
evaluation interval, not per step.
"""
from __future__ import annotations

import logging
import os
import socket
from typing import Any, Callable, List

import torch
from torch import distributed as dist
from torch import multiprocessing as mp
from torch.utils.data import Dataset, DistributedSampler, RandomSampler, Sampler, SequentialSampler

LOCAL_PROCESS_GROUP = None

# The reference hard-codes NCCL and requires CUDA (``distributed.py:168-169,175``).  CPU-only
# test rigs may switch the transport to gloo; the product default stays NCCL.
BACKEND = "nccl"
REQUIRE_CUDA = True


def _active() -> bool:
    return dist.is_available() and dist.is_initialized()


def get_rank() -> int:
    """Rank of this process, 0 outside a process group."""
    return dist.get_rank() if _active() else 0


def get_local_rank() -> int:
    """Rank inside this machine's group, 0 outside a process group."""
    if not _active():
        return 0
    group = LOCAL_PROCESS_GROUP
    if not group:                       # same complaint as the reference (distributed.py:36-37)
        raise ValueError("LOCAL_PROCESS_GROUP is None")
    return dist.get_rank(group=group)


def gather(tensor: torch.Tensor, tensor_list: List[torch.Tensor] = None):
    """Gather ``tensor`` from every rank into ``tensor_list`` on rank 0 (``None`` elsewhere)."""
    dist.gather(tensor, gather_list=tensor_list if is_primary() else None, dst=0)


def is_primary() -> bool:
    return get_rank() == 0


def synchronize() -> None:
    """Barrier over all ranks; a no-op for a single process."""
    if not _active() or dist.get_world_size() == 1:
        return
    dist.barrier()


def get_world_size() -> int:
    return dist.get_world_size() if _active() else 1


def data_sampler(dataset: Dataset, shuffle: bool, distributed: bool) -> Sampler:
    """``DistributedSampler(dataset, shuffle=True)`` when distributed (the reference ignores
    ``shuffle`` there, ``distributed.py:96``), else Random / Sequential sampler.

    The distributed shard is ``randperm(N, seed 0 + epoch)`` head-padded to a multiple of the
    world size, rank r taking ``[r::world]`` (``torch/utils/data/distributed.py:107-141``):
    independent shards, no exchange step.
    """
    if distributed:
        return DistributedSampler(dataset, shuffle=True)
    return RandomSampler(dataset) if shuffle else SequentialSampler(dataset)


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as sock:
        sock.bind(("", 0))
        return sock.getsockname()[1]


def launch(
    fn: Callable,
    n_gpu_per_machine: int,
    n_machine: int = 1,
    machine_rank: int = 0,
    dist_url: str = None,
    args: tuple = (),
) -> None:
    """Run ``fn(*args)`` in-process for a world of one, else once per GPU via ``mp.spawn``."""
    ranks = n_machine * n_gpu_per_machine
    if ranks == 1:                      # a world of one never spawns
        fn(*args)
        return None

    logging.info("Launching in distributed mode")
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    if dist_url == "auto":
        if n_machine != 1:
            raise ValueError("dist_url='auto' no supported in multi-machine jobs")      # message of distributed.py:146-147
        dist_url = "tcp://127.0.0.1:%d" % find_free_port()
    spawn_args = (fn, ranks, n_gpu_per_machine, machine_rank, dist_url, args)
    mp.spawn(job, nprocs=n_gpu_per_machine, args=spawn_args, daemon=False)


def job(
    local_rank: int,
    fn: Callable,
    world_size: int,
    n_gpu_per_machine: int,
    machine_rank: int = 0,
    dist_url: str = None,
    args: tuple = (),
) -> None:
    """Per-process entry: join the NCCL group, pin the GPU, build per-machine groups, run ``fn``."""
    global LOCAL_PROCESS_GROUP

    per = n_gpu_per_machine
    if REQUIRE_CUDA and not torch.cuda.is_available():
        raise OSError("CUDA is not available on this machine")           # distributed.py:168-169
    try:
        dist.init_process_group(backend=BACKEND, init_method=dist_url, world_size=world_size,
                                rank=machine_rank * per + local_rank)
    except Exception:
        raise OSError("NCC groups failed to initialize")                 # distributed.py:181-182 (sic)
    synchronize()
    if REQUIRE_CUDA:
        have = torch.cuda.device_count()
        if per > have:
            raise ValueError(f"Asked for {per} gpus bu got {have} available")   # distributed.py:186-189 (sic)
        torch.cuda.set_device(local_rank)
    if LOCAL_PROCESS_GROUP is not None:
        raise ValueError("torch.distributed.LOCAL_PROCESS_GROUP is not None")
    # every rank creates every machine's group (new_group is collective) and keeps its own
    machines = [dist.new_group(list(range(m * per, m * per + per))) for m in range(world_size // per)]
    LOCAL_PROCESS_GROUP = machines[machine_rank]
    fn(*args)


# ---------------------------------------------------------------------------------------------
# addition: the metric exchange
# ---------------------------------------------------------------------------------------------
def all_reduce_sum(tensor: torch.Tensor, group=None) -> torch.Tensor:
    """In-place ``all_reduce(SUM)``; a no-op outside a process group or for a world of one."""
    if _active() and dist.get_world_size(group) > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)
    return tensor


def all_reduce_metrics(state, group=None, transport: str = "int64") -> None:
    """Sum a ``MetricState`` over ranks: one all_reduce of the packed buffer ``[n, correct_1, correct_k, invalid,
    cm...]`` and one of the float64 ``[loss_sum, loss_count]``.  Integer sums are exact; the float64 pair is summed in
    NCCL's (run-to-run deterministic) order.

    ``transport="int64"`` sends the state as it is (8 MB at 1000 classes).  ``transport="int32"`` sends the bins as
    int32 and the four header values as 21-bit limbs (``tb_metric_state_pack_i32``): half the bytes, the same totals,
    as long as the ranks together have scored fewer than 2**31 rows since the last reset.  Every rank checks its own
    share of that bound (``rows_fed * world < 2**31``) and raises instead of wrapping; all ranks of one job must pass
    the same ``transport``.
    """
    if transport not in ("int64", "int32"):
        raise ValueError(f"transport must be 'int64' or 'int32', not {transport!r}")
    if not (_active() and dist.get_world_size(group) > 1):
        return
    if transport == "int32" and state.buffer.is_cuda:
        world = dist.get_world_size(group)
        if state.rows_fed * world >= 2 ** 31:
            raise ValueError(f"int32 transport: this rank scored {state.rows_fed} rows, {world} such ranks could pass 2**31 "
                             "per bin; use transport='int64'")
        wire = state.pack_wire()
        dist.all_reduce(wire, op=dist.ReduceOp.SUM, group=group)
        state.unpack_wire()
        state.rows_fed *= world
    else:
        all_reduce_sum(state.buffer, group=group)
    all_reduce_sum(state.loss, group=group)
