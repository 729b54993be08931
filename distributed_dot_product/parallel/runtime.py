# -*- coding: utf-8 -*-
"""Process bootstrap and the symmetric heap behind the fused collective kernels.

Replaces the reference's import-time ``hvd.init()`` + MPI world
(reference: distributed_dot_product/utils/comm.py:3-10, module.py:19) with

* a lazy, torchrun-style bootstrap (``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` / ``MASTER_ADDR``),
  one process per GPU; ``torch.distributed`` is used only for rendezvous, handle exchange and the
  CPU/gloo plumbing path;
* a **symmetric heap**: every rank allocates the same layout with the native allocator, exchanges
  CUDA IPC handles once, and maps every peer's heap, so kernels address remote workspaces and flag
  pads directly over NVLink / NVSwitch;
* a monotonically increasing call sequence number shared by all ranks (SPMD call order), used by the
  kernels as the flag epoch: flags are never reset, workspaces are double buffered by ``seq & 1``.
"""
import ctypes
import os
import threading

import torch
import torch.distributed as dist

from distributed_dot_product import _native

_state = {
    'initialized': False,
    'world': 1,
    'rank': 0,
    'local_rank': 0,
    'group': None,
    'owns_pg': False,
}
_lock = threading.RLock()


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def init(backend=None, device=None):
    """Initialises the runtime (idempotent).  Called lazily by every public entry point.

    * ``torch.distributed`` already initialised -> reuse the default group;
    * ``WORLD_SIZE > 1`` in the environment (torchrun) -> ``init_process_group`` with NCCL when CUDA
      is available, gloo otherwise, and bind this process to ``cuda:LOCAL_RANK``;
    * otherwise a single-process world of size 1.
    """
    with _lock:
        if _state['initialized']:
            return
        if dist.is_available() and dist.is_initialized():
            _state['world'] = dist.get_world_size()
            _state['rank'] = dist.get_rank()
            _state['local_rank'] = _env_int('LOCAL_RANK', _state['rank'])
        elif _env_int('WORLD_SIZE', 1) > 1:
            use_cuda = torch.cuda.is_available()
            _state['local_rank'] = _env_int('LOCAL_RANK', _env_int('RANK', 0))
            if use_cuda:
                torch.cuda.set_device(_state['local_rank'] % torch.cuda.device_count())
            if backend is None:
                backend = 'cuda:nccl,cpu:gloo' if use_cuda else 'gloo'
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            os.environ.setdefault('MASTER_PORT', '29533')
            kwargs = {}
            if use_cuda:
                kwargs['device_id'] = torch.device('cuda', torch.cuda.current_device())
            try:
                dist.init_process_group(backend=backend, **kwargs)
            except TypeError:
                dist.init_process_group(backend=backend)
            _state['owns_pg'] = True
            _state['world'] = dist.get_world_size()
            _state['rank'] = dist.get_rank()
        _state['initialized'] = True


def shutdown():
    """Releases the symmetric heap and, if this module created it, the process group."""
    with _lock:
        heap = _Heap._instance
        if heap is not None:
            heap.release()
            _Heap._instance = None
        if _state['owns_pg'] and dist.is_initialized():
            dist.destroy_process_group()
        _state.update(initialized=False, world=1, rank=0, local_rank=0, owns_pg=False)


def get_world_size():
    init()
    return _state['world']


def get_rank():
    init()
    return _state['rank']


def get_local_rank():
    init()
    return _state['local_rank']


def barrier():
    """Host-side barrier (the kernels themselves synchronise with device-side flags)."""
    init()
    if _state['world'] > 1:
        dist.barrier()


def _align(x, a):
    return (x + a - 1) // a * a


class _Heap(object):
    """One symmetric allocation per rank:

        [ done counters : 256 B ][ exit counter : 256 B ]
        [ parity 0 : workspace | flags ][ parity 1 : workspace | flags ]

    Offsets are identical on every rank because every rank sizes it from the same (SPMD) request.
    """
    _instance = None
    HEADER = 512
    GRANULE = 64 << 20

    def __init__(self):
        self.ws_cap = 0
        self.flag_cap = 0
        self.base = None          # own allocation (int address)
        self.peers = []           # addresses of every rank's heap as mapped here (own included)
        self.seq = 0
        self.comm = [None, None]  # prebuilt ctypes structs per parity
        self.generation = 0
        self.backend = 'ipc'      # 'vmm' (cuMemCreate + POSIX descriptors) or 'ipc' (cudaMalloc + CUDA IPC)
        self.mc_base = None       # multicast alias of the whole heap (vmm back end, NVSwitch multicast)
        self.heap_note = ''       # why a better back end was not used (driver error text)
        self._stream = None       # the one stream the collective kernels are enqueued on
        self._status_host = None  # pinned mirror of the status word (see next_comm)

    @classmethod
    def get(cls):
        if cls._instance is None:
            cls._instance = _Heap()
        return cls._instance

    def release(self):
        if self.base is None:
            return
        lib = _native.load()
        torch.cuda.synchronize()
        if _state['world'] > 1 and dist.is_initialized():
            dist.barrier()
        if self.backend == 'vmm':
            if self.mc_base:
                lib.ddp_vmm_release(ctypes.c_void_p(self.mc_base))
            for r, p in enumerate(self.peers):
                if p:
                    lib.ddp_vmm_release(ctypes.c_void_p(p))   # own mapping included
        else:
            for r, p in enumerate(self.peers):
                if r != _state['rank'] and p:
                    lib.ddp_heap_close(ctypes.c_void_p(p))
            lib.ddp_heap_free(ctypes.c_void_p(self.base))
        self.mc_base = None
        self.base = None
        self.peers = []
        self.ws_cap = self.flag_cap = 0
        self._stream = None

    def ensure(self, ws_bytes, n_flags):
        """Collective when it has to grow: every rank passes the same sizes in the same call."""
        if ws_bytes <= self.ws_cap and n_flags <= self.flag_cap:
            return
        lib = _native.load()
        world, rank = _state['world'], _state['rank']
        new_ws = max(_align(int(ws_bytes * 1.25), self.GRANULE), self.ws_cap)
        new_flags = max(_align(int(n_flags * 1.25), 1 << 16), self.flag_cap, 1 << 16)
        self.release()
        parity_bytes = new_ws + _align(new_flags * 4, 256)
        total = self.HEADER + 2 * parity_bytes
        self._allocate(total)
        torch.cuda.synchronize()
        dist.barrier()
        self.ws_cap, self.flag_cap = new_ws, new_flags
        self.seq = 0
        self.generation += 1
        for parity in (0, 1):
            cm = _native.Comm()
            cm.world, cm.rank, cm.phase_mask = world, rank, 3
            off = self.HEADER + parity * parity_bytes
            for r in range(world):
                cm.ws[r] = self.peers[r] + off
                cm.flags[r] = self.peers[r] + off + new_ws
                cm.done[r] = self.peers[r]
            cm.exit_counter = self.base + 256
            cm.status = self.base + 384
            cm.mc_ws = (self.mc_base + off) if self.mc_base else None
            self.comm[parity] = cm

    #: poll the sticky status word every this many collective calls (stream-ordered copy, no device sync)
    POLL_EVERY = 64

    # ------------------------------------------------------------------ allocation back ends
    def _allocate(self, total):
        """Allocates this rank's heap and maps every peer's.  Back ends (``DDP_HEAP`` = auto | vmm | ipc):

        * ``vmm``: CUDA virtual-memory-management allocations (``cuMemCreate``) exchanged as POSIX file descriptors
          over unix-domain sockets, plus - where the driver and the fabric offer it and ``DDP_MULTICAST`` != 0 - one
          multicast object over all ranks' heaps, whose address the all-gather pushers store to (``multimem.st``);
        * ``ipc``: ``cudaMalloc`` + CUDA IPC handles (round-1 path), also the automatic fall-back.
        The choice is collective: every rank reports whether its step worked and all ranks take the same branch."""
        lib = _native.load()
        world, rank = _state['world'], _state['rank']
        want = os.environ.get('DDP_HEAP', 'auto')
        self.backend, self.mc_base, self.heap_note = 'ipc', None, ''
        if want in ('auto', 'vmm'):
            fd_ok, mc_ok, gran = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_size_t(0)
            rc = lib.ddp_vmm_query(ctypes.byref(fd_ok), ctypes.byref(mc_ok), ctypes.byref(gran))
            ok = rc == 0 and fd_ok.value != 0
            note = '' if ok else ('ddp_vmm_query: %s' % (_native.last_error() if rc else 'no POSIX descriptor export'))
            if self._all_ok(ok):
                size = _align(total, gran.value)
                ptr, fd = ctypes.c_void_p(), ctypes.c_int(-1)
                rc = lib.ddp_vmm_alloc(size, ctypes.byref(ptr), ctypes.byref(fd))
                if rc:
                    note = _native.last_error()
                if self._all_ok(rc == 0):
                    fds = self._exchange_fds(fd.value)
                    peers, good = [], True
                    for r in range(world):
                        if r == rank:
                            peers.append(ptr.value)
                            continue
                        q = ctypes.c_void_p()
                        if lib.ddp_vmm_open(fds[r], size, ctypes.byref(q)):
                            good, note = False, _native.last_error()
                            peers.append(None)
                        else:
                            peers.append(q.value)
                    for r, f in enumerate(fds):
                        if f is not None and f >= 0:
                            os.close(f)
                    if self._all_ok(good):
                        self.backend, self.base, self.peers = 'vmm', ptr.value, peers
                        self.vmm_size = size
                        if mc_ok.value and world > 1 and os.environ.get('DDP_MULTICAST', '1') != '0':
                            self._bind_multicast(size)
                        return
                    for q in peers:
                        if q:
                            lib.ddp_vmm_release(ctypes.c_void_p(q))
                elif rc == 0:
                    os.close(fd.value)
                    lib.ddp_vmm_release(ptr)
            self.heap_note = note
            if want == 'vmm':
                raise RuntimeError('DDP_HEAP=vmm: the VMM heap could not be set up on every rank (%s)' % note)
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        _native.check(lib.ddp_heap_alloc(total, ctypes.byref(ptr), handle))
        self.base = ptr.value
        handles = [None] * world
        dist.all_gather_object(handles, bytes(handle.raw))
        self.peers = []
        for r in range(world):
            if r == rank:
                self.peers.append(self.base)
            else:
                p = ctypes.c_void_p()
                buf = ctypes.create_string_buffer(handles[r], 64)
                _native.check(lib.ddp_heap_open(buf, ctypes.byref(p)))
                self.peers.append(p.value)

    @staticmethod
    def _all_ok(ok):
        """True iff ``ok`` holds on every rank (collective)."""
        if _state['world'] == 1:
            return bool(ok)
        flags = [None] * _state['world']
        dist.all_gather_object(flags, bool(ok))
        return all(flags)

    def _exchange_fds(self, fd):
        """All-to-all of one file descriptor per rank (SCM_RIGHTS over abstract unix-domain sockets; the ranks of
        a heap live on one node).  Returns the list of received descriptors, ``None`` at this rank's own index."""
        import socket
        import struct
        world, rank = _state['world'], _state['rank']
        fds = [None] * world
        if world == 1:
            os.close(fd)
            return fds
        self._fd_round = getattr(self, '_fd_round', 0) + 1
        tag = '%s-%s-%d' % (os.environ.get('MASTER_PORT', '0'), os.environ.get('TORCHELASTIC_RUN_ID', 'x'),
                            self._fd_round)
        name = lambda r: '\0ddp-b200-heap-%s-%d' % (tag, r)
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name(rank))
        srv.listen(world)
        dist.barrier()
        for r in range(world):
            if r == rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(name(r))
            socket.send_fds(c, [struct.pack('i', rank)], [fd])
            c.close()
        for _ in range(world - 1):
            conn, _addr = srv.accept()
            msg, got, _flags, _a = socket.recv_fds(conn, 4, 1)
            fds[struct.unpack('i', msg)[0]] = got[0]
            conn.close()
        srv.close()
        os.close(fd)
        dist.barrier()
        return fds

    def _bind_multicast(self, size):
        """One multicast object over all ranks' heaps (collective; any failure leaves ``mc_base`` unset and the
        pushers on unicast stores, with the driver's error in ``heap_note``)."""
        lib = _native.load()
        world, rank = _state['world'], _state['rank']
        obj, fd = ctypes.c_ulonglong(0), ctypes.c_int(-1)
        ok = True
        if rank == 0:
            ok = lib.ddp_mc_create(size, world, ctypes.byref(obj), ctypes.byref(fd)) == 0
            if not ok:
                self.heap_note = _native.last_error()
        if not self._all_ok(ok):
            self.heap_note = self.heap_note or 'rank 0 could not create the multicast object'
            return
        send = fd.value if rank == 0 else os.open(os.devnull, os.O_RDONLY)
        fds = self._exchange_fds(send)
        if rank != 0:
            ok = lib.ddp_mc_import(fds[0], ctypes.byref(obj)) == 0
        for f in fds:
            if f is not None and f >= 0:
                os.close(f)
        if ok:
            ok = lib.ddp_mc_add_device(obj.value) == 0
        if not ok:
            self.heap_note = _native.last_error()
        if not self._all_ok(ok):      # also the barrier "every device added" that binding needs
            return
        mc = ctypes.c_void_p()
        ok = lib.ddp_mc_bind(obj.value, ctypes.c_void_p(self.base), size, ctypes.byref(mc)) == 0
        if not ok:
            self.heap_note = _native.last_error()
        if self._all_ok(ok):
            self.mc_base = mc.value
        elif ok:
            lib.ddp_vmm_release(mc)

    def next_comm(self, chunk_rows):
        """The communication block of the next collective call (sequence number = epoch of its flags).

        The epoch / parity protocol assumes that every collective kernel of this process is enqueued on ONE
        stream, in the same order on every rank; a second stream would let call ``seq`` overtake ``seq - 1`` and
        reuse a workspace parity that is still being read.  The first call fixes the stream, later calls from
        another stream raise.  Every ``POLL_EVERY`` calls the status word is copied to pinned host memory in
        stream order and the value of the PREVIOUS poll is examined, so a peer that timed out turns into an
        exception within at most 2 * POLL_EVERY calls without a single device synchronisation."""
        stream = torch.cuda.current_stream().cuda_stream
        if self._stream is None:
            self._stream = stream
        elif stream != self._stream:
            raise RuntimeError('distributed-dot-product: collective kernels must be enqueued on one stream '
                               '(first call used stream %#x, this one %#x)' % (self._stream, stream))
        self.seq += 1
        if self.seq % self.POLL_EVERY == 0:
            self._poll_status(stream)
        cm = self.comm[self.seq & 1]
        cm.seq = self.seq
        cm.chunk_rows = int(chunk_rows)
        cm.gather_ws = None
        return cm

    def _poll_status(self, stream):
        lib = _native.load()
        if self._status_host is None:
            self._status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        elif int(self._status_host[0]) != 0:   # written by the copy enqueued POLL_EVERY calls ago
            raise RuntimeError('distributed-dot-product: a fused kernel timed out waiting for a peer '
                               '(wait id %d); results since then are invalid' % int(self._status_host[0]))
        _native.check(lib.ddp_read_u32_async(ctypes.c_void_p(self.base + 384),
                                             ctypes.c_void_p(self._status_host.data_ptr()),
                                             ctypes.c_void_p(stream)))


def heap():
    init()
    return _Heap.get()


def check_status():
    """Failure detection: raises if a fused kernel on this rank ever gave up waiting for a peer.

    The kernels never hang the GPU: a wait on a peer flag that exceeds 20 s records which wait it was
    (1/2 = gathered rows for nt/all, 3/4 = a peer's done counter, 5 = reduce-scatter partials) in a sticky
    status word of the symmetric heap and carries on with invalid data.  Call this (it synchronises the
    device) at step boundaries of long jobs, or after an unexpected result, to turn that into an exception.
    """
    if not torch.cuda.is_available() or not _native.available():
        return 0
    lib = _native.load()
    torch.cuda.synchronize()
    site = max(lib.ddp_debug_timeout_site(), lib.ddp_debug_timeout_site_rowops(), lib.ddp_debug_timeout_site_attn())
    if site > 0:
        raise RuntimeError('distributed-dot-product: an in-kernel pipeline (mbarrier) wait gave up after 4 s (site %d: '
                           '1 smem slot, 2 TMEM buffer, 3 operands, 4 accumulator, 5 P chunk, 6 softmax row, 11-59 attention kernels); results are '
                           'invalid - please report this shape' % site)
    hp = _Heap._instance
    if hp is None or hp.base is None:
        return 0
    host = ctypes.c_uint32(0)
    rc = lib.ddp_read_u32(ctypes.c_void_p(hp.base + 384), ctypes.byref(host))
    _native.check(rc)
    if host.value:
        raise RuntimeError('distributed-dot-product: a fused kernel timed out waiting for a peer '
                           '(wait id %d); results since then are invalid' % host.value)
    return 0
