"""Payload transports: the device hop between the two halves of a split model.

The reference only simulates the hop (in-place quant->dequant,
Experiments/Pythia-70M/pythia_model.py:176-188); here the packed payload really moves.

  LocalTransport   both halves on one device: device-to-device copy (keeps the hop in the timing)
  PeerTransport    one process driving two devices: cudaMemcpyPeerAsync over NVLink (C ABI)
  PairExchange     one process per GPU (torch.distributed): rank r sends its payload to its pair
                   partner r^1 and receives the partner's, in one grouped NCCL send/recv, so both
                   NVLink directions carry a payload at once.  With the gloo backend (CPU tests)
                   the same code path runs on host tensors.
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
