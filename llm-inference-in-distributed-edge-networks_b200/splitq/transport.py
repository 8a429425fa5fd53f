"""Payload transports: the device hop between the two halves of a split model.

The reference only simulates the hop (in-place quant->dequant,
Experiments/Pythia-70M/pythia_model.py:176-188); here the packed payload really moves.

  LocalTransport   both halves on one device: device-to-device copy (keeps the hop in the timing)
  PeerTransport    one process driving two devices: cudaMemcpyPeerAsync over NVLink (C ABI)
  PairExchange     one process per GPU (torch.distributed): rank r sends its payload to its pair
                   partner r^1 and receives the partner's, in one grouped NCCL send/recv, so both
                   NVLink directions carry a payload at once.  With the gloo backend (CPU tests)
                   the same code path runs on host tensors.
  IpcPairExchange  same pairing, no SM involved: copy-engine pull through a CUDA-IPC mapping,
                   ordered by stream memory operations (for use under OverlappedBoundary).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import boundary


class LocalTransport:
    def __init__(self, device) -> None:
        self.src = self.dst = torch.device(device)

    def send(self, payload_src: torch.Tensor, payload_dst: torch.Tensor) -> torch.Tensor:
        return boundary.transfer(payload_dst, payload_src)


class PeerTransport:
    def __init__(self, src_device, dst_device) -> None:
        self.src, self.dst = torch.device(src_device), torch.device(dst_device)

    def send(self, payload_src: torch.Tensor, payload_dst: torch.Tensor) -> torch.Tensor:
        boundary.transfer(payload_dst, payload_src)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.src))
        torch.cuda.current_stream(self.dst).wait_event(ev)
        return payload_dst


def pair_partner(rank: int, world_size: int) -> int:
    """Ranks pair up as (0,1), (2,3), ...; an unpaired last rank is its own partner."""
    partner = rank ^ 1
    return partner if partner < world_size else rank


class PairExchange:
    def __init__(self, rank: Optional[int] = None, world_size: Optional[int] = None, group=None) -> None:
        import torch.distributed as dist

        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world_size is None else world_size
        self.partner = pair_partner(self.rank, self.world)

    def send(self, payload_src: torch.Tensor, payload_dst: torch.Tensor) -> torch.Tensor:
        """payload_src -> partner's payload_dst; partner's payload_src -> my payload_dst."""
        if self.partner == self.rank:
            if payload_src.is_cuda:
                return boundary.transfer(payload_dst, payload_src)
            payload_dst.copy_(payload_src)
            return payload_dst
        dist = self.dist
        ops = [dist.P2POp(dist.isend, payload_src, self.partner, self.group),
               dist.P2POp(dist.irecv, payload_dst, self.partner, self.group)]
        if self.rank > self.partner:  # same op order on both sides of the pair
            ops.reverse()
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        return payload_dst


class IpcPairExchange:
    """Pair exchange for one process per GPU that needs no SM: rank r's copy engine PULLS the
    partner's packed payload through a CUDA-IPC mapping (NVLink P2P), and the two ranks order
    themselves with stream memory operations on 32-bit step counters (C ABI: splitq_ipc_*,
    splitq_stream_write_u32 / _wait_geq_u32, splitq_copy_async).  NCCL's send/recv kernel has to
    wait for a gap between two launches of the persistent scoring kernel, which fills every SM;
    a DMA pull and memory operations do not.

    Protocol, step n = 1, 2, ... (slot = plan index, `depth` slots in rotation), identical on both
    ranks; `ready` / `free` are counters in MY memory that the PARTNER writes:
      main stream  wait free >= n - depth      partner has pulled what this slot held before
                   ... score, select, quantise + pack into plans[slot].payload_src ...
                   write partner.ready = n     my payload n is complete
      side stream  wait ready >= n             partner's payload n is complete
                   copy  plans[slot].payload_dst <- partner's plans[slot].payload_src   (copy engine)
                   write partner.free = n
                   ... dequantise ...
    `OverlappedBoundary` calls before_pack / after_pack / send at these points.  torch.distributed
    (any backend, gloo is enough) is used once, to swap the IPC handles.  Two ranks may share one
    GPU (how the 1-GPU test runs it)."""

    def __init__(self, plans, rank: Optional[int] = None, world_size: Optional[int] = None, group=None) -> None:
        import ctypes as C

        import torch.distributed as dist

        from . import _native as N

        self._C, self._N, self.dist, self.group = C, N, dist, group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world_size is None else world_size
        self.partner = pair_partner(self.rank, self.world)
        self.plans = list(plans)
        self.depth = len(self.plans)
        self.dev = self.plans[0].src
        self.n_pack = 0
        self.trace = None  # a list: send() appends (before, after) timing events around its wait for the partner
        self.flags = torch.zeros(2, dtype=torch.int32, device=self.dev)  # [ready, free], written by the partner
        self.peer_payload = [None] * self.depth
        self.peer_flags = None
        self._opened = []
        lib = N.load()
        # Every rank goes through the same two collectives whatever fails locally, and all ranks
        # raise together: a rank that bailed out alone would leave its partner waiting.
        err, exported = None, []
        try:
            for t in [p.payload_src for p in self.plans] + [self.flags]:
                handle = (C.c_ubyte * N.IPC_HANDLE_BYTES)()
                off = C.c_size_t(0)
                N.check(lib.splitq_ipc_export(C.c_void_p(t.data_ptr()), handle, C.byref(off)))
                exported.append((bytes(handle), off.value))
            torch.cuda.synchronize(self.dev)  # the zeroed flags are in memory before anyone maps them
        except Exception as e:  # noqa: BLE001
            err = e
        everyone = [None] * self.world
        dist.all_gather_object(everyone, None if err else exported, group=group)
        if err is None and self.partner != self.rank:
            try:
                if everyone[self.partner] is None:
                    raise N.SplitqError(f"rank {self.partner} could not export its buffers")
                ptrs = []
                for handle, off in everyone[self.partner]:
                    out = C.c_void_p()
                    buf = (C.c_ubyte * N.IPC_HANDLE_BYTES).from_buffer_copy(handle)
                    N.check(lib.splitq_ipc_open(buf, C.c_size_t(off), C.byref(out)))
                    ptrs.append(out.value)
                    self._opened.append((out.value, off))
                self.peer_payload, self.peer_flags = ptrs[:self.depth], ptrs[self.depth]
            except Exception as e:  # noqa: BLE001
                err = e
        status = [None] * self.world
        dist.all_gather_object(status, None if err is None else f"rank {self.rank}: {err}", group=group)
        failed = [m for m in status if m]
        if failed:  # nobody starts unless every mapping exists
            for ptr, off in self._opened:
                lib.splitq_ipc_close(C.c_void_p(ptr), C.c_size_t(off))
            self._opened = []
            raise N.SplitqError("IPC exchange unavailable: " + "; ".join(failed))

    # -- hooks called by OverlappedBoundary ---------------------------------------------------
    def _stream(self):
        return self._C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)

    def before_pack(self, plan) -> None:
        self.n_pack += 1
        plan._ipc_step = self.n_pack
        if self.partner != self.rank and self.n_pack > self.depth:
            self._N.check(self._N.load().splitq_stream_wait_geq_u32(
                self._C.c_void_p(self.flags.data_ptr() + 4), self.n_pack - self.depth, self._stream()))

    def after_pack(self, plan) -> None:
        if self.partner != self.rank:
            self._N.check(self._N.load().splitq_stream_write_u32(
                self._C.c_void_p(self.peer_flags), plan._ipc_step, self._stream()))

    def send(self, payload_src: torch.Tensor, payload_dst: torch.Tensor) -> torch.Tensor:
        """Called on the side stream: fetch the partner's payload of the same step into payload_dst."""
        if self.partner == self.rank:
            return boundary.transfer(payload_dst, payload_src)
        lib, C = self._N.load(), self._C
        slot = next(i for i, p in enumerate(self.plans) if p.payload_src is payload_src)
        n = self.plans[slot]._ipc_step
        st = self._stream()
        if self.trace is not None:
            before, after = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            before.record(torch.cuda.current_stream(self.dev))
        self._N.check(lib.splitq_stream_wait_geq_u32(C.c_void_p(self.flags.data_ptr()), n, st))
        if self.trace is not None:
            after.record(torch.cuda.current_stream(self.dev))
            self.trace.append((before, after))
        self._N.check(lib.splitq_copy_async(C.c_void_p(payload_dst.data_ptr()), C.c_void_p(self.peer_payload[slot]),
                                            payload_dst.numel(), st))
        self._N.check(lib.splitq_stream_write_u32(C.c_void_p(self.peer_flags + 4), n, st))
        return payload_dst

    def close(self) -> None:
        """Both ranks must be past their last pull before either unmaps or frees anything."""
        torch.cuda.synchronize(self.dev)
        self.dist.barrier(group=self.group)
        lib = self._N.load()
        for ptr, off in self._opened:
            lib.splitq_ipc_close(self._C.c_void_p(ptr), self._C.c_size_t(off))
        self._opened = []
