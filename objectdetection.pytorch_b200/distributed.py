"""Multi-GPU use of the hot path: one process per GPU, images sharded along dim 0.

The reference's only multi-GPU mechanism is single-process nn.DataParallel, which
scatters the batch, runs FocalLoss per replica and gathers the two 1-element
losses to GPU 0 for a .mean() (train_ksevendet.py:299-302, 358-361).  Here every
rank runs the match + loss kernels on its own shard and the four local sums
{sum cls_i, sum reg_i, sum num_pos_i, B_local} are combined with ONE all-reduce
(NCCL over NVLink on the GPU box, gloo in the CPU tests).  Gradients are
pre-scaled by the global batch inside the loss kernel, so the backward of the
path needs no collective.  The eval side needs no collective at all.
"""
from __future__ import annotations

import weakref

import torch
import torch.distributed as dist


def shard_range(batch: int, rank: int, world: int):
    """Images [lo, hi) owned by `rank` (equal contiguous shards, like DataParallel's scatter)."""
    if batch % world != 0:
        raise ValueError(f"batch {batch} must divide evenly over {world} ranks")
    per = batch // world
    return rank * per, (rank + 1) * per


def combine_loss_sums(sums: torch.Tensor, group=None) -> torch.Tensor:
    """sums = [sum cls_i, sum reg_i, sum num_pos_i, B_local] -> all-reduced copy (in place)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return sums


def mean_losses(sums: torch.Tensor):
    """(cls_mean[1], reg_mean[1]) of losses.py:168 from (all-reduced) sums."""
    m = sums[:2] / sums[3]
    return m[0:1], m[1:2]


_SHARDED = {}      # (process group, device, kwargs) -> FocalLoss: ONE module (and one set of peer mailboxes) per group


def sharded_focal_loss(classifications, regressions, anchors, annotations, group=None, **loss_kwargs):
    """FocalLoss over a batch whose dim 0 is already this rank's shard.  The module - and with it the
    peer mailboxes of the fused exchange (a cudaMalloc, IPC handles and two host collectives to set
    up) - is built once per (group, device, kwargs) and reused by every later call."""
    from .losses import FocalLoss
    ref = classifications[0] if isinstance(classifications, (list, tuple)) else classifications
    key = (id(group) if group is not None else None, ref.device, tuple(sorted((k, repr(v)) for k, v in loss_kwargs.items())))
    mod = _SHARDED.get(key)
    if mod is None:
        mod = _SHARDED[key] = FocalLoss(process_group=group, sharded=True, **loss_kwargs)
    return mod(classifications, regressions, anchors, annotations)


def _release_mailboxes(lib, device, opened, local):
    """PeerMailbox finalizer (also close()): unmap the peers' mailboxes, free the local one."""
    try:
        with torch.cuda.device(device):
            torch.cuda.synchronize()
            while opened:
                lib.k7_peer_mailbox_close(opened.pop())
            if local[0]:
                lib.k7_peer_mailbox_destroy(local[0])
                local[0] = None
    except Exception:                      # interpreter / context teardown: nothing left to release
        pass


class PeerMailbox:
    """Peer-mapped mailboxes for the exchange fused into k7_loss (include/k7det.h: K7PeerReduce).

    Every rank owns a zeroed mailbox of 2*world slots on its GPU (its own cudaMalloc allocation, so
    that it can be exported with the CUDA IPC API), the 64-byte handles travel once through
    torch.distributed (all_gather_object: plumbing), and every rank maps the others' mailboxes over
    NVLink.  After that a training step needs no collective launch: the last block of the loss row
    kernel posts the four batch sums into every mailbox and adds what it receives in rank order.
    One process per GPU on ONE node; raises RuntimeError when the IPC mapping is not possible (the
    caller then keeps the NCCL all-reduce)."""

    def __init__(self, group=None, device=None, timeout_ms: int = 10000):
        import ctypes as C
        from . import _lib as L
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerMailbox needs an initialised torch.distributed process group")
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > L.K7_MAX_PEERS:
            raise RuntimeError(f"PeerMailbox supports up to {L.K7_MAX_PEERS} ranks")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.timeout_ms = int(timeout_ms)
        self.seq = 0
        self._lib = L.lib()
        self._local = C.c_void_p()
        self._opened = []
        self._local_box = [None]               # what the finalizer frees (it must not hold `self`)
        self._finalizer = weakref.finalize(self, _release_mailboxes, self._lib, self.device, self._opened, self._local_box)
        self._finalizer.atexit = False         # at interpreter exit the CUDA context goes away as a whole
        handle = (C.c_ubyte * 64)()
        err = None
        with torch.cuda.device(self.device):
            try:
                L.check(self._lib.k7_peer_mailbox_create(self.world, C.byref(self._local), handle), "k7_peer_mailbox_create")
                self._local_box[0] = self._local
            except RuntimeError as e:                      # still take part in the collectives below
                err = str(e)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, (self.rank, bytes(handle), err), group=group)
            self._ptrs = [None] * self.world
            if all(g[2] is None for g in gathered):
                for r, hb, _ in gathered:
                    if r == self.rank:
                        self._ptrs[r] = self._local.value
                        continue
                    p = C.c_void_p()
                    try:
                        L.check(self._lib.k7_peer_mailbox_open((C.c_ubyte * 64).from_buffer_copy(hb), C.byref(p)),
                                "k7_peer_mailbox_open")
                        self._opened.append(p)
                        self._ptrs[r] = p.value
                    except RuntimeError as e:
                        err = str(e)
                        break
            else:
                err = next(g[2] for g in gathered if g[2] is not None)
            # everybody must agree before the first post
            flags = [None] * self.world
            dist.all_gather_object(flags, err, group=group)
        bad = next((f for f in flags if f is not None), None)
        if bad is not None:
            self.close()
            raise RuntimeError("peer mailboxes unavailable: " + bad)
        self._struct = L.K7PeerReduce()
        self._struct.world, self._struct.rank, self._struct.timeout_ms = self.world, self.rank, self.timeout_ms
        # the call counter lives on the device and is advanced by the kernel: a captured CUDA graph can be replayed
        self._calls = torch.zeros(2, dtype=torch.int32, device=self.device)      # [calls posted, calls collected]
        self._struct.seq_counter = self._calls.data_ptr()
        self.collect_stream = torch.cuda.Stream(device=self.device)               # where deferred collections wait
        self.last_collect = None                                                   # (event, captured?) of the last deferred collection
        for r in range(self.world):
            self._struct.mailbox[r] = self._ptrs[r]

    def next(self):
        """The K7PeerReduce of the next k7_loss call.  The call number is counted on the device
        (K7PeerReduce.seq_counter: every rank makes the same calls), so nothing per-call is frozen into a
        captured CUDA graph; `seq` mirrors the number of calls ENQUEUED from this object."""
        self.seq += 1
        self._struct.seq = self.seq
        return self._struct

    def post_only(self):
        """The K7PeerReduce of a k7_loss call that only POSTS (FocalLoss.start): a copy of the struct with
        post_only set."""
        import ctypes as C
        from . import _lib as L
        st = L.K7PeerReduce()
        C.memmove(C.byref(st), C.byref(self.next()), C.sizeof(L.K7PeerReduce))
        st.post_only = 1
        return st

    def collect(self, after: "torch.cuda.Event") -> "torch.Tensor":
        """Enqueue k7_peer_collect on the collection stream, behind `after` (recorded behind the posting k7_loss):
        returns the [4] tensor the global sums will be in; self.last_collect is recorded behind it."""
        from . import _lib as L
        import ctypes as C
        with torch.cuda.stream(self.collect_stream):
            self.collect_stream.wait_event(after)
            sums = torch.empty(4, dtype=torch.float32, device=self.device)
            with torch.cuda.device(self.device):
                L.check(self._lib.k7_peer_collect(C.byref(self._struct), L.ptr(sums), L.stream_ptr(self.device)),
                        "k7_peer_collect")
            ev = torch.cuda.Event()
            ev.record(self.collect_stream)
        # (event, recorded under stream capture?): an event of the capture must not be waited for outside it and
        # vice versa; replays are ordered by their launches and join their collection before they end
        self.last_collect = (ev, bool(torch.cuda.is_current_stream_capturing()))
        return sums

    def close(self):
        """Unmap and free now (idempotent; also runs when the object is garbage collected)."""
        _release_mailboxes(self._lib, self.device, self._opened, self._local_box)
        self._local = None
