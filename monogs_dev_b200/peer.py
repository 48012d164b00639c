"""Peer-mapped parameter / gradient arenas for the data-parallel mapping window.

Every rank allocates its packed parameter arena and gradient arena in memory that all ranks of
the node map into their address space (torch's symmetric-memory allocator: CUDA VMM handles
exchanged once at rendezvous — plumbing only).  `b2gs_adam_step_peer` (csrc/adam.cu) then does
all_reduce(grad) + Adam + zero_grad as ONE kernel over NVLink loads / stores, bracketed by two
flag barriers; no NCCL call is left on the data path of a step.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib


class PeerArena:
    def __init__(self, device, group=None, multicast=True):
        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise RuntimeError("PeerArena needs an initialised process group")
        self._symm = symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if self.world > 16:
            raise RuntimeError("b2gs_adam_step_peer supports at most 16 ranks")
        self.device = torch.device(device)
        try:                                   # older torch wants an explicit opt-in per group
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass
        self._sig = symm.empty(64, dtype=torch.int32, device=self.device)
        self._sig.zero_()
        self._sig_h = symm.rendezvous(self._sig, self.group)
        self.fail = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.want_multicast = bool(multicast)
        self.multicast = False           # set by alloc(): the arenas have an NVLS multicast mapping
        self._arena = self._arena_h = None
        self._struct = None
        self.total = 0

    def alloc(self, total):
        """(flat, flat_grad) of `total` floats each, zero-filled, mapped by every rank.  Collective."""
        n = (int(total) + 63) // 64 * 64
        self._arena = self._symm.empty(2 * n, dtype=torch.float32, device=self.device)
        self._arena.zero_()
        self._arena_h = self._symm.rendezvous(self._arena, self.group)
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)               # every rank's zeros have landed before anyone reads
        self.total = n
        W = self.world
        self._pb = (C.c_void_p * W)(*[int(p) for p in self._arena_h.buffer_ptrs])
        self._gb = (C.c_void_p * W)(*[int(p) + 4 * n for p in self._arena_h.buffer_ptrs])
        self._sb = (C.c_void_p * W)(*[int(p) for p in self._sig_h.buffer_ptrs])
        mc = 0
        if self.want_multicast:
            try:
                mc = int(self._arena_h.multicast_ptr or 0) if self._arena_h.has_multicast_support else 0
            except Exception:
                mc = 0
        self.multicast = mc != 0
        self._struct = _lib.B2gsPeerArena(W, self.rank, self._pb, self._gb, self._sb, self.fail.data_ptr(),
                                          mc if mc else None, (mc + 4 * n) if mc else None)
        return self._arena[:total], self._arena[n:n + total]

    def set_fail_word(self, word):
        """let the barrier's time-out flag live in caller memory (one int32 device word), e.g. next to
        the overflow flags a window reads back every step"""
        self.fail = word
        if self._struct is not None:
            self._struct.fail_flag = word.data_ptr()

    def struct(self):
        if self._struct is None:
            raise RuntimeError("PeerArena.alloc() has not been called")
        return C.byref(self._struct)

    def check(self):
        """host-side: did a barrier time out (a peer died / fell out of step)?"""
        if int(self.fail.item()):
            raise RuntimeError("libb2gs peer barrier timed out")
