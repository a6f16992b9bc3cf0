import os
import sys


class _World:
    def __init__(self):
        self._size = int(os.environ.get("WORLD_SIZE", "1"))
        self._rank = int(os.environ.get("RANK", "0"))
        self._dist = None
        self._group = None
        if self._size > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                dist.init_process_group("gloo")
            self._dist = dist
            # a dedicated gloo group: the default group may be NCCL when the caller initialised it
            self._group = dist.new_group(backend="gloo")

    def Get_rank(self):
        return self._rank

    def Get_size(self):
        return self._size

    def gather(self, obj, root=0):
        if self._size == 1:
            return [obj]
        out = [None] * self._size if self._rank == root else None
        self._dist.gather_object(obj, out, dst=root, group=self._group)
        return out

    def bcast(self, obj, root=0):
        if self._size == 1:
            return obj
        box = [obj]
        self._dist.broadcast_object_list(box, src=root, group=self._group)
        return box[0]

    def Barrier(self):
        if self._size > 1:
            self._dist.barrier(group=self._group)

    barrier = Barrier

    def Abort(self, errorcode=1):
        sys.stdout.flush()
        os._exit(errorcode)


class _Lazy:
    """COMM_WORLD is created on first use so importing the module has no side effects."""
    _w = None

    def __getattr__(self, name):
        if _Lazy._w is None:
            _Lazy._w = _World()
        return getattr(_Lazy._w, name)


COMM_WORLD = _Lazy()
