"""NVLS (NVLink-SHARP multicast) memory for the cross-GPU aggregation: every rank binds one allocation to a shared multicast
object; ``multimem.ld_reduce`` on the multicast address returns the sum over all GPUs' copies, computed in the NVSwitch
(csrc/fedb_nvls.cu, the ``nvls`` branch of k_fedavg_p2p).  The multicast handle is a POSIX file descriptor: rank 0 creates it and
hands it to the other ranks over a Unix-domain socket (SCM_RIGHTS); torch.distributed only carries the socket path and the
barriers around the set-up.
"""
from __future__ import annotations

import ctypes
import os
import socket
import tempfile
from typing import Optional

import torch

from ..ops import _native as nat


def supported(device: Optional[int] = None) -> bool:
    lib = nat.lib(required=False)
    if lib is None or not torch.cuda.is_available() or not hasattr(lib, "fedb_nvls_supported"):
        return False
    return lib.fedb_nvls_supported(torch.cuda.current_device() if device is None else device) == 1


class NvlsBuffer:
    """``nbytes`` of device memory on every rank, bound to one multicast object.  ``uc_ptr`` addresses this rank's copy,
    ``mc_ptr`` the multicast view (for multimem.* instructions only).  Collective: every rank must construct it."""

    def __init__(self, nbytes: int, rank: int, world: int):
        import torch.distributed as dist
        self.lib = nat.lib(required=True)
        self.rank, self.world = rank, world
        dev = torch.cuda.current_device()
        mc = ctypes.c_uint64(0)
        size = ctypes.c_int64(0)
        if rank == 0:
            fd = ctypes.c_int32(-1)
            nat.check(self.lib.fedb_nvls_create(world, nbytes, ctypes.byref(mc), ctypes.byref(fd), ctypes.byref(size)))
            path = os.path.join(tempfile.gettempdir(), f"fedb200_nvls_{os.getpid()}_{id(self) & 0xFFFF}.sock")
            if os.path.exists(path):
                os.unlink(path)
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(path)
            srv.listen(world)
            meta = [path, int(size.value)]
            dist.broadcast_object_list(meta, src=0)
            for _ in range(world - 1):
                conn, _addr = srv.accept()
                socket.send_fds(conn, [b"fd"], [int(fd.value)])
                conn.recv(1)                      # the peer has imported it
                conn.close()
            srv.close()
            os.unlink(path)
            os.close(int(fd.value))
        else:
            meta = [None, None]
            dist.broadcast_object_list(meta, src=0)
            path, size.value = meta[0], int(meta[1])
            cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            cli.settimeout(60)
            cli.connect(path)
            _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
            nat.check(self.lib.fedb_nvls_import(int(fds[0]), ctypes.byref(mc)))
            cli.send(b"k")
            cli.close()
            os.close(int(fds[0]))
        self.mc_handle, self.nbytes = int(mc.value), int(size.value)
        nat.check(self.lib.fedb_nvls_add_device(self.mc_handle, dev))
        dist.barrier()                            # binding needs every device added
        mem = ctypes.c_uint64(0)
        uc, mcp = ctypes.c_void_p(), ctypes.c_void_p()
        nat.check(self.lib.fedb_nvls_bind_map(self.mc_handle, dev, self.nbytes, ctypes.byref(mem), ctypes.byref(uc), ctypes.byref(mcp)))
        self.mem_handle, self.uc_ptr, self.mc_ptr = int(mem.value), int(uc.value), int(mcp.value)
        dist.barrier()                            # every copy is bound and zeroed before anyone reduces

    def close(self) -> None:
        if self.mc_handle:
            self.lib.fedb_nvls_release(self.mc_handle, self.mem_handle, self.uc_ptr, self.mc_ptr, self.nbytes)
            self.mc_handle = 0
