"""Peer-memory arena for the node-local data-parallel step (csrc/peer.cu; include/arcb200.h "data-parallel step over
peer memory").  One process per GPU: each rank allocates one arena with the library's own cudaMalloc (so that the
allocation is an IPC-exportable base pointer, independent of torch's caching allocator), the 64-byte handles travel
through torch.distributed's host channel once at start-up, and every rank maps the others' arenas.  After that the
step's gradient exchange is plain loads/stores over NVLink from our kernels; NCCL is not on the data path.

Replaces DDP's gradient all-reduce + per-rank Adam of the reference (common/trainer/basic_trainer.py:192-202)."""
import ctypes as C

import torch

from . import _lib

FLAG_BYTES = 256          # 16 uint64 barrier slots + padding
ALIGN = 256


class _DevMem:
    """exposes a raw device range to torch through __cuda_array_interface__ (no copy, no ownership)"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = dict(shape=(nbytes,), typestr='|u1', data=(ptr, False), version=2)


def _round_up(x, a):
    return (x + a - 1) // a * a


class PeerArena:
    """layout (same on every rank): [flags | region_0 | region_1 | ...], regions given as {name: nbytes}"""

    def __init__(self, regions, device, rank, world, exchange):
        """exchange(bytes) -> list of every rank's bytes, in rank order (e.g. torch.distributed.all_gather_object)"""
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.offsets, off = {}, FLAG_BYTES
        for name, nbytes in regions.items():
            self.offsets[name] = (off, nbytes)
            off = _round_up(off + nbytes, ALIGN)
        self.nbytes = off
        with torch.cuda.device(self.device):
            base, handle = C.c_void_p(), (C.c_ubyte * 64)()
            alloc_err = None
            try:
                _lib.call('arcb200_peer_alloc', self.nbytes, C.byref(base), handle)
            except RuntimeError as e:                       # still take part in the exchange: the others are waiting in it
                alloc_err = e
            self.base = base.value
            handles = exchange(None if alloc_err is not None else bytes(handle))
            assert len(handles) == world
            if alloc_err is not None or any(h is None for h in handles):
                raise RuntimeError('peer arena allocation failed on a rank: %s' % (alloc_err or 'see that rank'))
            self.bases = []
            try:
                for w, h in enumerate(handles):
                    if w == rank:
                        self.bases.append(self.base)
                        continue
                    p = C.c_void_p()
                    _lib.call('arcb200_peer_open', (C.c_ubyte * 64).from_buffer_copy(h), C.byref(p))
                    self.bases.append(p.value)
            except RuntimeError:
                for w, b in enumerate(self.bases):          # unmap what was mapped before the failure
                    if w != rank and b:
                        _lib.call('arcb200_peer_close', b)
                self.bases = None
                raise
        self._bytes = torch.as_tensor(_DevMem(self.base, self.nbytes), device=self.device)
        self.epoch = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._flag_ptrs = (C.c_void_p * world)(*self.bases)

    def local(self, name, dtype):
        """torch view of this rank's own region"""
        off, nbytes = self.offsets[name]
        return self._bytes[off:off + nbytes].view(dtype)

    def pointers(self, name, byte_offset=0):
        """ctypes array of every rank's pointer to the region (index = rank), as mapped in this process"""
        off = self.offsets[name][0] + byte_offset
        return (C.c_void_p * self.world)(*[b + off for b in self.bases])

    def barrier(self, timeout_ms=0):
        """flag barrier between the ranks, on the current stream (nothing blocks on the host).  A rank that waits longer than
        the timeout (default ARCB200_PEER_TIMEOUT_MS or 10 minutes) ends the job: the kernel traps, see arcb200_peer_barrier"""
        if getattr(self, 'host_barrier', None) is not None:
            # ranks that share ONE device (a test on a 1-GPU box): kernels of different processes are time-sliced, so a kernel
            # spinning on another rank's flag may never see it run - synchronise through the host instead
            torch.cuda.current_stream(self.device).synchronize()
            self.host_barrier()
            return
        _lib.call('arcb200_peer_barrier', self._flag_ptrs, self.world, self.rank, self.epoch.data_ptr(),
                  self.status.data_ptr(), int(timeout_ms), _lib.current_stream())

    def check(self):
        """host-side check after a synchronisation: raises if a barrier wait timed out (a rank fell out of step)"""
        if int(self.status.item()) != 0:
            raise RuntimeError('peer barrier timed out on rank %d: the ranks are not executing the same steps' % self.rank)

    def close(self, sync_ranks=None):
        """Unmap the peers' arenas and free the own one.  sync_ranks: a callable that returns once EVERY rank has unmapped
        (e.g. torch.distributed.barrier) - the own allocation is only released after it, so that no peer still has it
        mapped; without it the own arena is left to the process exit."""
        if getattr(self, 'bases', None) is None:
            return
        torch.cuda.synchronize(self.device)
        self._bytes = None
        for w, b in enumerate(self.bases):
            if w != self.rank and b:
                _lib.call('arcb200_peer_close', b)
        self.bases = None
        if sync_ranks is not None:
            sync_ranks()
            if self.base:
                _lib.call('arcb200_peer_free', self.base)
                self.base = None


class SymmArena(PeerArena):
    """The same arena on torch's symmetric memory (torch.distributed._symmetric_memory: cuMemCreate + fabric handles exchanged
    through the process group's store): every rank's pointer to every arena as before, plus - when the fabric supports it - ONE
    multicast address per region that reaches all arenas at once (multimem.ld_reduce / multimem.st, csrc/peer.cu
    adam_mc16_kernel).  Torch is plumbing here: allocation and handle exchange; the kernels are the library's."""

    def __init__(self, regions, device, rank, world, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.offsets, off = {}, FLAG_BYTES
        for name, nbytes in regions.items():
            self.offsets[name] = (off, nbytes)
            off = _round_up(off + nbytes, ALIGN)
        self.nbytes = off
        with torch.cuda.device(self.device):
            self._bytes = symm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            self._hdl = symm.rendezvous(self._bytes, group if group is not None else dist.group.WORLD)
            self._bytes.zero_()
            torch.cuda.synchronize(self.device)
        self.bases = [int(b) for b in self._hdl.buffer_ptrs]
        self.base = self.bases[rank]
        assert self.base == self._bytes.data_ptr()
        self.mc_base = int(self._hdl.multicast_ptr or 0)
        self.epoch = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._flag_ptrs = (C.c_void_p * world)(*self.bases)

    def mc_pointer(self, name, byte_offset=0):
        """multicast address of the region (0 when the fabric has no multicast)"""
        return self.mc_base + self.offsets[name][0] + byte_offset if self.mc_base else 0

    def close(self, sync_ranks=None):
        """torch owns the allocation and the mappings: synchronise, let every rank finish, drop the references"""
        if getattr(self, 'bases', None) is None:
            return
        torch.cuda.synchronize(self.device)
        if sync_ranks is not None:
            sync_ranks()
        self.bases, self._hdl, self._bytes = None, None, None


def adam_step_mc16(mc_grad16, mc_half_out, lo, hi, p, m, v, step, lr, scale=128.0, beta1=0.9, beta2=0.999, eps=1e-8,
                   weight_decay=0.0, grad_scale=1.0, step_dev=None):
    """mc_grad16 / mc_half_out: multicast addresses from SymmArena.mc_pointer"""
    _lib.call('arcb200_adam_step_mc16', mc_grad16, mc_half_out, int(lo), int(hi), p.data_ptr(), m.data_ptr(), v.data_ptr(),
              float(scale), int(step), float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
              _lib.ptr(step_dev), _lib.current_stream())


def adam_step_peer(grads, half_out, world, lo, hi, p, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8,
                   weight_decay=0.0, grad_scale=1.0, step_dev=None):
    """grads / half_out: ctypes pointer arrays from PeerArena.pointers (half_out may be None)"""
    _lib.call('arcb200_adam_step_peer', grads, half_out, int(world), int(lo), int(hi), p.data_ptr(), m.data_ptr(),
              v.data_ptr(), int(step), float(lr), float(beta1), float(beta2), float(eps), float(weight_decay),
              float(grad_scale), _lib.ptr(step_dev), _lib.current_stream())


def grad_to_half(g32, g16, keep_lo, keep_hi, scale=128.0):
    """fp16 loss-scaled copy of the gradient outside [keep_lo, keep_hi); clears g32 there"""
    _lib.call('arcb200_grad_to_half', g32.data_ptr(), g16.data_ptr(), g32.numel(), int(keep_lo), int(keep_hi), float(scale),
              _lib.current_stream())


def adam_step_peer16(g_own, grads16, half_out, world, rank, lo, hi, p, m, v, step, lr, scale=128.0, beta1=0.9, beta2=0.999,
                     eps=1e-8, weight_decay=0.0, grad_scale=1.0, step_dev=None):
    """grads16 / half_out: ctypes pointer arrays from PeerArena.pointers"""
    _lib.call('arcb200_adam_step_peer16', g_own.data_ptr(), grads16, half_out, int(world), int(rank), int(lo), int(hi),
              p.data_ptr(), m.data_ptr(), v.data_ptr(), float(scale), int(step), float(lr), float(beta1), float(beta2), float(eps),
              float(weight_decay), float(grad_scale), _lib.ptr(step_dev), _lib.current_stream())


def slice_bounds(n, rank, world, quantum=1024):
    """[lo, hi) of rank's slice of an n-element vector; slices start on multiples of `quantum` elements"""
    per = _round_up((n + world - 1) // world, quantum)
    return min(n, rank * per), min(n, (rank + 1) * per)
