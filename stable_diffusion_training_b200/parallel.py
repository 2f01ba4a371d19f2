"""Data-parallel gradient mean for the denoising train step: one process per GPU, bucketed NCCL all-reduce over
NVLink 5 / NVSwitch, issued on a communication stream while backward still runs.

Replaces what the reference gets implicitly from GSPMD: the batch is sharded on the "data_parallel" mesh axis and
everything else is replicated (/root/reference/training_utils.py:35-37, 835-840, 852-931), so XLA inserts an all-reduce
(mean) of the fp32 gradients and of the scalar loss. Here:

  * the per-rank loss gradient is pre-scaled by 1/world (train_step passes loss_grad_scale), so a SUM all-reduce of the
    gradient arena yields the mean over the global batch — identical to the 1-rank result on the concatenated batch;
  * the arena is cut into contiguous buckets along its own order, which inside each segment is the order in which
    backward finishes the leaves (training_utils.ParamArena), so buckets become ready front to back;
  * Graph.backward() reports finished leaves (unet.Graph.backward(on_grads=...)); when the last leaf of a bucket is in,
    an event is recorded on the compute stream, the communication stream waits on it and enqueues the all-reduce;
  * finish() flushes what is left, reduces the scalar loss, and makes the compute stream wait for the communication
    stream — the optimizer (replicated, bit-identical on every rank) then runs on reduced gradients.

No collective is issued when world_size == 1. The class works on any torch tensor + process group (the CPU tests run
it with gloo, world_size 2); streams/events are used only for CUDA tensors.
"""
from __future__ import annotations

import os
import sys
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

DEFAULT_BUCKET_BYTES = int(os.environ.get("SDB_DP_BUCKET_MB", "256")) << 20  # r4h: 8 GPUs 43.4 ms (256 MB) vs 45.8 (64 MB)


class NcclComm:
    """A second NCCL communicator over the same ranks, driven directly through the C API (ctypes on the libnccl that
    torch ships): ncclAllReduce is enqueued on OUR communication stream with no framework bookkeeping around it — no
    work objects, no watchdog events — which is what makes the collectives capturable into the step's CUDA graph
    (CompiledTrainStep then replays ONE graph per step instead of a chain of segments with eager collectives between
    them). torch.distributed stays the plumbing: rendezvous, the unique-id broadcast, barriers."""

    _F32, _SUM = 7, 0  # ncclFloat32, ncclSum (nccl.h)

    def __init__(self, group=None):
        import ctypes as C
        lib = None
        errs = []
        cands = []
        try:
            import nvidia.nccl as _n
            cands += [os.path.join(d, "lib", "libnccl.so.2") for d in list(_n.__path__)]
        except Exception as e:  # pragma: no cover
            errs.append(repr(e))
        for cand in cands + ["libnccl.so.2"]:
            try:
                lib = C.CDLL(cand)
                break
            except OSError as e:
                errs.append(str(e))
        if lib is None:
            raise RuntimeError(f"NcclComm: libnccl.so.2 not loadable ({errs})")

        class UniqueId(C.Structure):
            _fields_ = [("internal", C.c_byte * 128)]

        self._C, self._lib = C, lib
        lib.ncclGetErrorString.restype = C.c_char_p
        lib.ncclGetErrorString.argtypes = [C.c_int]
        lib.ncclGetUniqueId.argtypes = [C.POINTER(UniqueId)]
        lib.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
        lib.ncclAllReduce.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.ncclCommDestroy.argtypes = [C.c_void_p]
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        uid = UniqueId()
        if self.rank == 0:
            self._check(lib.ncclGetUniqueId(C.byref(uid)), "ncclGetUniqueId")
        box = [bytes(uid.internal)]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        C.memmove(C.byref(uid), box[0], 128)
        self.comm = C.c_void_p()
        self._check(lib.ncclCommInitRank(C.byref(self.comm), self.world, uid, self.rank), "ncclCommInitRank")

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise RuntimeError(f"{what} failed: {self._lib.ncclGetErrorString(rc).decode()}")

    def all_reduce_sum_f32(self, t: torch.Tensor, stream: "torch.cuda.Stream") -> None:
        """In-place SUM all-reduce of an fp32 CUDA tensor, enqueued on `stream` (capturable)."""
        assert t.dtype == torch.float32 and t.is_cuda and t.is_contiguous()
        self._check(self._lib.ncclAllReduce(t.data_ptr(), t.data_ptr(), t.numel(), self._F32, self._SUM, self.comm,
                                            stream.cuda_stream), "ncclAllReduce")

    def destroy(self) -> None:
        if self.comm:
            self._lib.ncclCommDestroy(self.comm)
            self.comm = None


def build_buckets(ranges: Dict[str, Tuple[int, int]], bucket_elems: int, tail_bucket_elems: Optional[int] = None,
                  tail_span_elems: int = 0) -> List[Tuple[int, int, List[str]]]:
    """Cut the flat arena into contiguous [start, end) buckets on leaf boundaries, in arena order; every bucket holds
    at least `bucket_elems` elements except possibly the last. Leaves larger than a bucket get their own. The last
    `tail_span_elems` elements of the arena (the gradients backward finishes last) are cut into smaller buckets of
    `tail_bucket_elems`, so that little communication is left exposed between the end of backward and the optimizer
    while the bulk moves in few large buckets (few cross-rank synchronisations)."""
    leaves = sorted(ranges.items(), key=lambda kv: kv[1][0])
    out: List[Tuple[int, int, List[str]]] = []
    cur: List[str] = []
    start = None
    end = 0
    total_end = max((e for _, (_, e) in leaves), default=0)
    for name, (s, e) in leaves:
        if start is None:
            start = s
        cur.append(name)
        end = e
        want = bucket_elems
        if tail_bucket_elems and start >= total_end - tail_span_elems:
            want = tail_bucket_elems
        elif tail_bucket_elems and end > total_end - tail_span_elems:
            want = 0  # close the last large bucket where the tail begins
        if end - start >= want:
            out.append((start, end, cur))
            cur, start = [], None
    if cur:
        out.append((start, end, cur))
    # buckets tile the range they span: the alignment padding between leaves rides along with the bucket before it
    fixed = [(s, out[i + 1][0] if i + 1 < len(out) else e, names) for i, (s, e, names) in enumerate(out)]
    return fixed


class PeerAllReduce:
    """Gradient all-reduce over NVLink / NVSwitch PEER MEMORY with no communication kernel on the SMs.

    The gradient arena lives in symmetric memory (torch.distributed._symmetric_memory: every rank maps every other
    rank's arena). A bucket [s, e) is cut into `world` slices; rank r owns slice r:
        barrier                      every rank has finished writing the bucket
        pull   (copy engines)        slice r of every peer's arena  ->  local staging buffers
        reduce (sdb_sum_slices_f32)  arena[slice r] = sum over ranks in rank order          (our kernel, HBM-bound)
        push   (copy engines)        arena[slice r]  ->  slice r of every peer's arena
    and one barrier after the last bucket. Bytes on the wire per rank equal a ring all-reduce's (2 (N-1)/N of the
    arena), but they are moved by the DMA engines: NCCL's all-reduce kernels hold 16-32 SMs for the whole transfer, and a
    persistent one-CTA-per-SM GEMM launched meanwhile runs as two waves (the CTAs that find their SM taken start when
    the first ones finish) — that, not the exposed tail, was the data-parallel overhead of round 1 (3-6 ms per step).
    Only rank r ever computes slice r, so every rank ends up with identical bits (the replicated-state contract of
    the reference: PartitionSpec() for parameters and optimizer state, training_utils.py:852-895)."""

    def __init__(self, numel: int, device, group=None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > 9:
            raise RuntimeError("PeerAllReduce: at most 9 ranks (one NVSwitch domain)")
        self.numel = numel
        self.buf = symm.empty(numel, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, self.group)
        self.peer = [self.buf if p == self.rank else self.hdl.get_buffer(p, (numel,), torch.float32, 0)
                     for p in range(self.world)]
        self.stage: Optional[torch.Tensor] = None
        self.direct = os.environ.get("SDB_DP_PEER_DIRECT", "0") == "1"
        # NVLS variant (SDB_DP_ALLREDUCE=nvls): the arena is also bound to an NVSwitch multicast object; the exchange is
        # ONE kernel per bucket slice (multimem.ld_reduce through the switch + multimem.st to every rank) instead of
        # pull / sum / push: each gradient crosses this rank's HBM once in and once out (no staging copies)
        self.mc_ptr = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        self.nvls = os.environ.get("SDB_DP_ALLREDUCE", "peer") == "nvls"
        if self.nvls and self.mc_ptr == 0:
            if self.rank == 0:
                sys.stderr.write("[sdb dp] SDB_DP_ALLREDUCE=nvls: no multicast support on this system, using the "
                                 "copy-engine exchange\n")
            self.nvls = False
        self.nvls_ctas = int(os.environ.get("SDB_DP_NVLS_CTAS", "0"))

    def _slices(self, s: int, e: int):
        n = e - s
        per = -(-n // self.world)
        per = -(-per // 4) * 4
        return [(min(e, s + q * per), min(e, s + (q + 1) * per)) for q in range(self.world)]

    def prepare(self, max_bucket_elems: int) -> None:
        if self.nvls:
            return   # no staging buffers
        per = -(-(-(-max_bucket_elems // self.world)) // 4) * 4
        self.stage = torch.empty((max(self.world - 1, 1), per), dtype=torch.float32, device=self.buf.device)

    def all_reduce_range(self, s: int, e: int) -> None:
        """SUM all-reduce of arena[s:e] on the CURRENT stream (s, e multiples of 4)."""
        import ctypes as C
        from . import _lib
        W, r = self.world, self.rank
        a, b = self._slices(s, e)[r]
        n = b - a
        self.hdl.barrier(channel=0, timeout_ms=30000)
        if self.nvls:
            if n > 0:
                _lib.check(_lib.lib().sdb_nvls_allreduce_f32(self.mc_ptr + 4 * a, n, self.nvls_ctas,
                                                             _lib.stream_ptr(self.buf.device)), "sdb_nvls_allreduce_f32")
            return
        order = [(r + k) % W for k in range(1, W)]           # start with the next rank: spreads the NVLink load
        if n > 0:
            if self.direct:
                # the reduction kernel loads the peers' slices over NVLink itself (no staging copy: 2 x less local
                # HBM traffic; the kernel's CTAs need no shared memory and co-reside with the GEMM CTAs)
                srcs = [self.peer[p][a:b].data_ptr() for p in range(W) if p != r]
            else:
                for p in order:
                    self.stage[p if p < r else p - 1][:n].copy_(self.peer[p][a:b])
                srcs = [self.stage[i].data_ptr() for i in range(W - 1)]
            ptrs = (C.c_void_p * 8)(*srcs + [None] * (9 - W))
            _lib.check(_lib.lib().sdb_sum_slices_f32(self.buf.data_ptr() + 4 * a, ptrs, W, r, n,
                                                     _lib.stream_ptr(self.buf.device)), "sdb_sum_slices_f32")
            for p in order:
                self.peer[p][a:b].copy_(self.buf[a:b])

    def finish(self) -> None:
        """After the last bucket: every rank's pushes have landed everywhere."""
        self.hdl.barrier(channel=0, timeout_ms=30000)


class GradReducer:
    def __init__(self, grad_flat: torch.Tensor, ranges: Dict[str, Tuple[int, int]], *, group=None,
                 bucket_bytes: int = DEFAULT_BUCKET_BYTES, total_elems: Optional[int] = None,
                 tail_bucket_bytes: int = 0, tail_span_bytes: int = 0):
        self.grad = grad_flat
        self.group = group
        self.world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        es = grad_flat.element_size()
        self.buckets = build_buckets(ranges, max(1, bucket_bytes // es),
                                     tail_bucket_elems=max(1, tail_bucket_bytes // es) if tail_bucket_bytes else None,
                                     tail_span_elems=tail_span_bytes // es)
        if total_elems is not None and self.buckets:  # trailing padding of the arena belongs to the last bucket
            s, e, names = self.buckets[-1]
            self.buckets[-1] = (s, max(e, total_elems), names)
        self.bucket_of = {n: i for i, (_, _, names) in enumerate(self.buckets) for n in names}
        self.cuda = grad_flat.is_cuda
        self.comm_stream = torch.cuda.Stream(device=grad_flat.device) if self.cuda else None
        # direct NCCL communicator (capturable collectives): CUDA tensors, more than one rank, unless SDB_DP_NCCL=torch
        self.nccl: Optional[NcclComm] = None
        if self.cuda and self.world_size > 1 and os.environ.get("SDB_DP_NCCL", "direct") != "torch":
            self.nccl = NcclComm(group)
        # peer-memory all-reduce (copy engines + our reduction kernel): attach() allocates the arena in symmetric memory
        self.peer: Optional[PeerAllReduce] = None
        self._pending: List[int] = []
        self._next = 0            # buckets are issued strictly in index order (identical on every rank)
        self._ready: List[bool] = []
        self.issued_order: List[int] = []
        self.bytes_reduced = 0
        # CUDA-graph capture support (training_utils.CompiledTrainStep): collectives are NOT captured. While
        # `capture` is set, ready buckets are only recorded; every `capture_cut_every` buckets the capturer is asked to
        # close the current graph segment, and at replay time the recorded buckets are all-reduced eagerly on the
        # communication stream right after that segment has been enqueued (replay_segment_buckets / replay_finish).
        self.capture = None          # callable(force: bool) -> None that cuts the graph segment
        self.capture_cut_every = int(os.environ.get("SDB_DP_CUT", "8"))
        self._deferred: List[int] = []
        self.before_issue = None     # callable() run before a bucket is handed to the collective / the capturer

    # ---- one step ----
    def begin(self, _graph=None) -> None:
        self._pending = [len(names) for _, _, names in self.buckets]
        self._ready = [False] * len(self.buckets)
        self._next = 0
        self.issued_order = []

    def on_param_grad(self, names: Sequence[str]) -> None:
        for n in names:
            b = self.bucket_of.get(n)
            if b is None:
                continue
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._ready[b] = True
        self._issue_ready()

    def _issue_ready(self, force: bool = False) -> None:
        while self._next < len(self.buckets) and (force or self._ready[self._next]):
            self._all_reduce_bucket(self._next)
            self._next += 1

    def _all_reduce_bucket(self, b: int) -> None:
        s, e, _ = self.buckets[b]
        self.issued_order.append(b)
        if self.capture is not None:
            # graph capture: the bucket is reduced eagerly after the segment that completes it has been replayed, so the
            # second (parameter-gradient) stream only has to be joined where a segment is cut, not at every bucket
            self._deferred.append(b)
            # the buckets of a segment are reduced after the whole segment: keep the LAST segments short (2 buckets) so
            # little communication is left exposed between the end of backward and the optimizer
            remaining = len(self.buckets) - (b + 1)
            every = self.capture_cut_every if remaining >= self.capture_cut_every else min(2, self.capture_cut_every)
            if len(self._deferred) >= every:
                if self.before_issue is not None:
                    self.before_issue()
                self.capture(False)
            return
        if self.before_issue is not None:
            self.before_issue()  # e.g. unet.Graph.join: parameter gradients are produced on a second stream
        if self.world_size == 1:
            return
        view = self.grad[s:e]
        self.bytes_reduced += view.numel() * view.element_size()
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.grad.device))
            self.comm_stream.wait_event(ev)
            self._reduce_on_comm_stream(view)
        else:
            dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group)

    def _reduce_on_comm_stream(self, t: torch.Tensor) -> None:
        if self.peer is not None and t.data_ptr() != 0 and self._in_arena(t):
            s = (t.data_ptr() - self.grad.data_ptr()) // 4
            with torch.cuda.stream(self.comm_stream):
                self.peer.all_reduce_range(s, s + t.numel())
            self._peer_dirty = True
            return
        if self.nccl is not None:
            self.nccl.all_reduce_sum_f32(t, self.comm_stream)
        else:
            with torch.cuda.stream(self.comm_stream):
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)

    def _in_arena(self, t: torch.Tensor) -> bool:
        a0 = self.grad.data_ptr()
        return a0 <= t.data_ptr() < a0 + self.grad.numel() * 4

    def _peer_finish(self) -> None:
        if self.peer is not None and getattr(self, "_peer_dirty", False):
            with torch.cuda.stream(self.comm_stream):
                self.peer.finish()
            self._peer_dirty = False

    @property
    def graph_capturable(self) -> bool:
        """True when the collectives may be captured into the step's CUDA graph (direct NCCL communicator)."""
        return self.nccl is not None

    def finish(self, loss: torch.Tensor) -> torch.Tensor:
        """Flush the remaining buckets, average the loss over ranks, join the communication stream."""
        self._issue_ready(force=True)
        if self.before_issue is not None:
            self.before_issue()
        if self.capture is not None:
            self.capture(True)  # the replayer reduces the remaining buckets and the loss, then joins the streams
            return loss
        if loss is None:  # a second model's gradients (the text encoder): buckets only
            if self.world_size > 1 and self.cuda:
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.grad.device))
                self.comm_stream.wait_event(ev)
                self._peer_finish()
                torch.cuda.current_stream(self.grad.device).wait_stream(self.comm_stream)
            return None
        if self.world_size > 1:
            if self.cuda:
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.grad.device))
                self.comm_stream.wait_event(ev)
                self._peer_finish()
                self._reduce_on_comm_stream(loss.view(1))
                torch.cuda.current_stream(self.grad.device).wait_stream(self.comm_stream)
                _scale_f32(loss, 1.0 / self.world_size)
            else:
                dist.all_reduce(loss, op=dist.ReduceOp.SUM, group=self.group)
                loss.mul_(1.0 / self.world_size)
        return loss


def _scale_f32(t: torch.Tensor, f: float) -> None:
    """t *= f on the current stream. (0-d loss: one torch elementwise kernel; not on the compute path.)"""
    t.mul_(f)


def _replay_buckets(red: GradReducer, buckets: Sequence[int]) -> None:
    if red.world_size == 1 or not buckets:
        return
    dev = red.grad.device
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(dev))
    red.comm_stream.wait_event(ev)
    for b in buckets:
        s, e, _ = red.buckets[b]
        red._reduce_on_comm_stream(red.grad[s:e])


GradReducer.replay_segment_buckets = _replay_buckets


def _replay_finish(red: GradReducer, buckets: Sequence[int], loss: torch.Tensor) -> None:
    """Eager tail of a replayed backward: remaining buckets, loss mean, join of the communication stream."""
    if red.world_size == 1:
        return
    _replay_buckets(red, buckets)
    dev = red.grad.device
    debug = os.environ.get("SDB_DP_DEBUG") == "1"
    ev = torch.cuda.Event(enable_timing=debug)
    ev.record(torch.cuda.current_stream(dev))
    red.comm_stream.wait_event(ev)
    red._peer_finish()
    red._reduce_on_comm_stream(loss.view(1))
    with torch.cuda.stream(red.comm_stream):
        loss.mul_(1.0 / red.world_size)
        if debug:  # exposed communication: end of this rank's backward -> last all-reduce complete (printed one step late)
            prev = getattr(red, "_dbg_events", None)
            if prev is not None and dist.get_rank(red.group) == 0:
                prev[1].synchronize()
                print(f"[sdb dp] exposed all-reduce tail {prev[0].elapsed_time(prev[1]):.3f} ms", file=sys.stderr)
            ev1 = torch.cuda.Event(enable_timing=True)
            ev1.record(red.comm_stream)
            red._dbg_events = (ev, ev1)
    torch.cuda.current_stream(dev).wait_stream(red.comm_stream)


GradReducer.replay_finish = _replay_finish


def _replicate_state(state, group, sync: str) -> None:
    """The reference replicates parameters and optimizer state from ONE host copy (jax.device_put with
    PartitionSpec(), training_utils.py:446-483). Here every rank builds its own state, so attach() makes that true:
    sync="broadcast" copies rank 0's masters / Lion codes / scales / fp32 momentum / EMA to every rank;
    sync="verify" only compares checksums and raises on a mismatch; sync="none" trusts the caller."""
    if sync == "none" or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    a = state.arena
    bufs = [a.master, a.codes, a.inv_scales, a.mu] + ([a.ema] if a.ema is not None else [])
    if sync == "broadcast":
        for t in bufs:
            dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        a.refresh_shadow()
        return
    if sync != "verify":
        raise ValueError(f"parallel.attach: sync must be 'broadcast', 'verify' or 'none', got {sync!r}")
    sums = torch.stack([t.view(torch.uint8).to(torch.int64).sum() if t.dtype == torch.int8 else
                        t.view(torch.int32).to(torch.int64).sum() for t in bufs])
    lo, hi = sums.clone(), sums.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    if not torch.equal(lo, hi):
        raise RuntimeError("parallel.attach: parameters / optimizer state differ between ranks "
                           "(use sync='broadcast' or load identical states)")


def attach(state, group=None, bucket_bytes: int = DEFAULT_BUCKET_BYTES, sync: str = "broadcast") -> GradReducer:
    """Give a TrainState its data-parallel reducer (train_step then shards nothing itself: each rank is handed its own
    slice of the global batch, as the reference's PartitionSpec("data_parallel", None) does). `sync` makes the
    replicated-state contract hold at attach time (see _replicate_state)."""
    a = state.arena
    if a.device.type == "cuda" and torch.cuda.current_device() != a.device.index:
        raise RuntimeError(f"parallel.attach: the state lives on {a.device}, the current device is "
                           f"cuda:{torch.cuda.current_device()} (call torch.cuda.set_device(local_rank) first)")
    _replicate_state(state, group, sync)
    peer = None
    mode = os.environ.get("SDB_DP_ALLREDUCE", "peer")
    if a.device.type == "cuda" and dist.is_initialized() and dist.get_world_size(group) > 1 and mode in ("peer", "nvls"):
        # the gradient arena moves into symmetric memory (peers read / write it directly); every view is rebound
        peer = PeerAllReduce(a.n_total, a.device, group)
        a.grad = peer.buf
        a.g.update({n: a._view(a.grad, n) for n in a.shapes})   # same dict object the UNet holds as G
        state._grad_tree = None
    ranges = {}
    for n, shape in a.shapes.items():
        numel = 1
        for d in shape:
            numel *= int(d)
        ranges[n] = (a.offset[n], a.offset[n] + numel)
    tail_mb = int(os.environ.get("SDB_DP_TAIL_MB", "32"))
    state.dp = GradReducer(a.grad, ranges, group=group, bucket_bytes=bucket_bytes, total_elems=a.n_total,
                           tail_bucket_bytes=tail_mb << 20, tail_span_bytes=(2 * tail_mb) << 20)
    if peer is not None:
        state.dp.peer = peer
        peer.prepare(max(e - s for s, e, _ in state.dp.buckets))
    return state.dp


def init_process_group_from_env(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """torchrun-style rendezvous (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT). Returns
    (rank, local_rank, world_size); a single process without those variables is world_size 1 and no group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, local, world
