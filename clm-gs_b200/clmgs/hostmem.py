"""Pinned host tables backed by transparent huge pages.

The offloaded state is four [N, 48] fp32 tables (19.6 GB each at 102 M Gaussians) that the host Adam and the GPU's
zero-copy row kernels touch at random 192-byte rows.  With 4 KiB pages every such access is also a TLB miss whose
page-table entry is itself a DRAM access (the page tables of 78 GB are ~150 MB); with 2 MiB pages the tables' page
tables fit the CPU caches.  cudaHostAlloc does not use huge pages, so: anonymous mmap + madvise(MADV_HUGEPAGE), then
cudaHostRegister (portable | mapped) -- under unified addressing the host address is valid in device code.
Falls back to torch's pinned allocator when huge pages or the registration are not available.
"""
import mmap
import os
import weakref

import numpy as np
import torch

_HUGE = 2 << 20
_STATE = dict(enabled=os.environ.get("CLMGS_HUGE_PAGES", "1") != "0", regions=0, fallbacks=0)


def _unregister(ptr):
    try:
        torch.cuda.cudart().cudaHostUnregister(ptr)
    except Exception:
        pass


def pinned_tensor(shape, dtype=torch.float32, zero=False):
    """CPU tensor of `shape` in pinned, device-mapped host memory (zero-filled if `zero`)."""
    shape = (shape,) if isinstance(shape, int) else tuple(int(x) for x in shape)
    if not torch.cuda.is_available():
        return torch.zeros(shape, dtype=dtype) if zero else torch.empty(shape, dtype=dtype)
    itemsize = torch.empty((), dtype=dtype).element_size()
    nbytes = int(np.prod(shape)) * itemsize
    if _STATE["enabled"] and nbytes >= 8 * _HUGE and hasattr(mmap, "MADV_HUGEPAGE"):
        try:
            size = (nbytes + _HUGE - 1) // _HUGE * _HUGE
            m = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
            m.madvise(mmap.MADV_HUGEPAGE)
            npdt = torch.empty((), dtype=dtype).numpy().dtype
            arr = np.frombuffer(m, dtype=npdt, count=int(np.prod(shape))).reshape(shape)
            t = torch.from_numpy(arr)
            err = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), size, 1 | 2)   # portable | mapped
            if int(err) != 0:
                raise RuntimeError("cudaHostRegister failed: %s" % (err,))
            weakref.finalize(arr, _unregister, t.data_ptr())
            _STATE["regions"] += 1
            return t            # fresh anonymous memory reads as zero
        except Exception:
            _STATE["fallbacks"] += 1
    return (torch.zeros if zero else torch.empty)(shape, dtype=dtype, pin_memory=True)


def huge_page_stats():
    """(regions allocated through the huge-page path, fallbacks, AnonHugePages of this process in GB)."""
    gb = None
    try:
        for ln in open("/proc/self/smaps_rollup"):
            if ln.startswith("AnonHugePages"):
                gb = round(int(ln.split()[1]) / 1e6, 2)
    except Exception:
        pass
    return dict(regions=_STATE["regions"], fallbacks=_STATE["fallbacks"], anon_huge_gb=gb)
