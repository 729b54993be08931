"""``horovod.torch`` subset over ``torch.distributed`` (see ../__init__.py)."""
import os

import torch
import torch.distributed as dist

Average = 'average'
Sum = 'sum'


def init():
    if dist.is_initialized():
        return
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29544')
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', rank)) % torch.cuda.device_count())
    backend = 'cuda:nccl,cpu:gloo' if use_cuda else 'gloo'
    dist.init_process_group(backend=backend, rank=rank, world_size=world)


def size():
    return dist.get_world_size() if dist.is_initialized() else 1


def rank():
    return dist.get_rank() if dist.is_initialized() else 0


def local_rank():
    return int(os.environ.get('LOCAL_RANK', rank()))


def mpi_threads_supported():
    return True


def allgather(tensor, name=None):
    """Concatenates the ranks' tensors along dim 0 (Horovod semantics)."""
    if size() == 1:
        return tensor.clone()
    tensor = tensor.contiguous()
    out = tensor.new_empty((tensor.size(0) * size(),) + tuple(tensor.shape[1:]))
    dist.all_gather_into_tensor(out, tensor)
    return out


class _Handle(object):
    def __init__(self, tensor, work, average):
        self.tensor, self.work, self.average = tensor, work, average


def allreduce_async(tensor, average=None, name=None, op=None):
    op = op if op is not None else (Average if average in (None, True) else Sum)
    out = tensor.clone()
    work = dist.all_reduce(out, async_op=True) if size() > 1 else None
    return _Handle(out, work, op == Average)


def synchronize(handle):
    if handle.work is not None:
        handle.work.wait()
    if handle.average:
        handle.tensor.div_(size())
    return handle.tensor


def allreduce(tensor, average=None, name=None, op=None):
    return synchronize(allreduce_async(tensor, average=average, name=name, op=op))


def broadcast_parameters(params, root_rank=0):
    if size() == 1:
        return
    items = params.items() if isinstance(params, dict) else params
    for _, p in sorted(items):
        if torch.is_tensor(p):
            dist.broadcast(p.data, src=root_rank)
