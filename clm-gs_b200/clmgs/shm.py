"""A host table shared by every rank of one node and mapped into every rank's GPU.

Multi-GPU design (new -- the reference is single-process): the ranks of one node share the node's
host memory, so the offloaded SH-parameter table needs no owner-to-renderer hop.  It lives in ONE anonymous shared
memory file (memfd: backed by RAM like /dev/shm, but not limited by the size of that mount), every rank maps the
whole table and registers the mapping with CUDA (pinned + device-mapped).  Any GPU then reads the rows its cameras
need straight from host DRAM over ITS OWN PCIe link with the zero-copy row kernels -- the same kernels as on one
GPU -- while each row is still updated only by its owner's host Adam (the ranks synchronise at step boundaries).
Gradient tables and Adam moments stay private to the owner.
"""
import ctypes
import mmap
import os

import numpy as np
import torch

_HOST_REGISTER_PORTABLE, _HOST_REGISTER_MAPPED = 1, 2


class SharedHostTable:
    """float32 [n_rows, cols] in node-shared host memory.  `tensor` is a CPU torch tensor over the mapping whose
    data_ptr() is valid in device code of this process after register_cuda()."""

    def __init__(self, n_rows, cols, group=None, use_dist=True):
        import torch.distributed as dist
        self.n_rows, self.cols = int(n_rows), int(cols)
        self.nbytes = max(self.n_rows * self.cols * 4, mmap.PAGESIZE)
        self.nbytes = (self.nbytes + mmap.PAGESIZE - 1) // mmap.PAGESIZE * mmap.PAGESIZE
        multi = use_dist and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        rank = dist.get_rank(group) if multi else 0
        self._fd = -1
        if rank == 0:
            self._fd = os.memfd_create("clmgs_sh_table", 0)
            os.ftruncate(self._fd, self.nbytes)
            path = ["/proc/%d/fd/%d" % (os.getpid(), self._fd)]
        else:
            path = [None]
        if multi:
            dist.broadcast_object_list(path, src=0, group=group)
            if rank != 0:
                self._fd = os.open(path[0], os.O_RDWR)   # same node, same user: a second handle on the same file
        self._map = mmap.mmap(self._fd, self.nbytes, flags=mmap.MAP_SHARED, prot=mmap.PROT_READ | mmap.PROT_WRITE)
        if multi:
            dist.barrier(group=group)                    # everyone holds the file before rank 0 could go away
        arr = np.frombuffer(self._map, dtype=np.float32, count=self.n_rows * self.cols).reshape(self.n_rows, self.cols)
        self.array = arr
        self.tensor = torch.from_numpy(arr)
        self._registered = False

    def register_cuda(self):
        """cudaHostRegister(portable | mapped) of the whole mapping in THIS process (each rank does it for its own
        device).  Unified addressing: the host address is the device address."""
        if self._registered:
            return
        rt = torch.cuda.cudart()
        ptr = self.tensor.data_ptr()
        err = rt.cudaHostRegister(ptr, self.nbytes, _HOST_REGISTER_PORTABLE | _HOST_REGISTER_MAPPED)
        if int(err) != 0:
            raise RuntimeError("cudaHostRegister of the shared SH table (%d bytes) failed: %s" % (self.nbytes, err))
        self._registered = True

    def close(self):
        if self._registered:
            try:
                torch.cuda.cudart().cudaHostUnregister(self.tensor.data_ptr())
            except Exception:
                pass
            self._registered = False
        self.tensor = self.array = None
        try:
            self._map.close()
        except BufferError:   # numpy views still alive: the mapping goes with the process
            pass
        if self._fd >= 0:
            os.close(self._fd)
            self._fd = -1
