"""Stand-in for mpi4py (no MPI on this box): ``MPI.COMM_WORLD`` with the three methods the reference uses
(``Get_size``, ``Barrier``, ``gather``), over ``torch.distributed``.  See horovod/__init__.py."""
import torch.distributed as dist


class _Comm(object):
    def Get_size(self):
        return dist.get_world_size() if dist.is_initialized() else 1

    def Get_rank(self):
        return dist.get_rank() if dist.is_initialized() else 0

    def Barrier(self):
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.barrier()

    def gather(self, obj, root=0):
        if not dist.is_initialized() or dist.get_world_size() == 1:
            return [obj]
        out = [None] * dist.get_world_size() if dist.get_rank() == root else None
        dist.gather_object(obj, out, dst=root)
        return out


class _MPI(object):
    COMM_WORLD = _Comm()


MPI = _MPI()
