"""Collectives for engines spread over several GPUs (include/gtsam_b200.h, gtb_set_distributed).

The library takes three collectives as a C callback; in a GTSAM build they are ncclReduce / ncclBroadcast /
ncclAllReduce on the engine stream.  Here the plumbing is torch.distributed (backend "nccl", one process per GPU).
`ThreadComm` is the same contract between engines that live in one process (threads, one GPU): it exists so that the
fan-in / broadcast logic can be parity-tested on a single-GPU box.
"""
import threading
import traceback

from . import capi


class _DeviceDoubles:
    """View of `count` doubles at a raw device pointer (CUDA array interface, no copy)."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f8", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def _tensor(ptr, count, device):
    import torch
    return torch.as_tensor(_DeviceDoubles(ptr, count), device=device)


def torch_comm(device, group=None):
    """capi.COMM_FN running the collectives with torch.distributed, ordered on the engine's stream."""
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", device)

    def fn(user, op, buf, count, stream):
        try:
            t = _tensor(buf, count, dev)
            with torch.cuda.stream(torch.cuda.ExternalStream(stream, device=dev)):
                if op == capi.COMM_REDUCE_SUM_TO_0:
                    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM, group=group)
                elif op == capi.COMM_BROADCAST_FROM_0:
                    dist.broadcast(t, src=0, group=group)
                elif op == capi.COMM_ALLREDUCE_SUM:
                    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
                elif op == capi.COMM_ALLREDUCE_MIN:
                    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
                else:
                    return 2
            return 0
        except Exception:   # an exception must not cross the C boundary
            traceback.print_exc()
            return 1

    return capi.COMM_FN(fn)


class ThreadComm:
    """The same collectives between `world` engines driven by `world` threads of one process on one GPU."""

    def __init__(self, world, device=0):
        self.world = world
        self.device = device
        self.barrier = threading.Barrier(world)
        self.slots = [None] * world
        self.calls = {0: 0, 1: 0, 2: 0, 3: 0}
        self.bytes = 0

    def callback(self, rank):
        import torch
        dev = torch.device("cuda", self.device)

        def fn(user, op, buf, count, stream):
            try:
                torch.cuda.ExternalStream(stream, device=dev).synchronize()
                self.slots[rank] = (buf, count, op)
                self.barrier.wait()
                if rank == 0:
                    assert all(s[1] == count and s[2] == op for s in self.slots), self.slots
                    ts = [_tensor(s[0], count, dev) for s in self.slots]
                    if op == capi.COMM_REDUCE_SUM_TO_0:
                        for t in ts[1:]:
                            ts[0] += t
                    elif op == capi.COMM_BROADCAST_FROM_0:
                        for t in ts[1:]:
                            t.copy_(ts[0])
                    elif op in (capi.COMM_ALLREDUCE_SUM, capi.COMM_ALLREDUCE_MIN):
                        acc = ts[0].clone()
                        for t in ts[1:]:
                            acc = acc + t if op == capi.COMM_ALLREDUCE_SUM else torch.minimum(acc, t)
                        for t in ts:
                            t.copy_(acc)
                    torch.cuda.synchronize(dev)
                    self.calls[op] += 1
                    self.bytes += 8 * count
                self.barrier.wait()
                return 0
            except Exception:
                traceback.print_exc()
                self.barrier.abort()
                return 1

        return capi.COMM_FN(fn)
