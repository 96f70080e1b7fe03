"""Ghost-cell exchange between parts placed one per GPU: the device-side stand-in for
`Part::ShareGhostCellCoeffs` / `UpdateGhostCellCoeffs` (include/mini/mesh/part.hpp:1242-1314).

The library packs the coefficients of the send cells into one device buffer (all neighbours back to
back, reference wire format: kFields doubles per cell) and unpacks the receive buffer into the ghost
cells; this module moves the bytes between ranks with `torch.distributed` point-to-point operations
(NCCL over NVLink on GPUs; gloo on CPU tensors for the host-logic tests).  The inner cells are
advanced while the messages are in flight, the inter cells afterwards (fem.hpp:249-256).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


class _DeviceArray:
    """Expose a raw device pointer through __cuda_array_interface__ so torch can alias it."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}


class HaloExchanger:
    """Exchanges `send` -> neighbours' `recv` (1-D float64 tensors, per-neighbour slices in list order)."""

    def __init__(self, neighbours, send_counts, recv_counts, kfields, group=None):
        self.neighbours = [int(q) for q in neighbours]
        self.send_off = np.concatenate([[0], np.cumsum(send_counts)]) * kfields
        self.recv_off = np.concatenate([[0], np.cumsum(recv_counts)]) * kfields
        self.group = group
        self.send = self.recv = None
        self._reqs = []
        self._ops = None
        self.inflight = False     # an exchange of the current state has been posted and not yet unpacked
        self.received = False     # the receive buffer holds the neighbours' values of some earlier state
        self._pack = self._unpack = lambda: None

    def bind(self, send: torch.Tensor, recv: torch.Tensor, pack=None, unpack=None):
        self.send, self.recv = send, recv
        self._ops = None
        if pack is not None:
            self._pack = pack
        if unpack is not None:
            self._unpack = unpack

    def attach(self, spatial):
        """Bind to a Spatial: alias the library's device buffers and run on torch's current stream."""
        ctx = spatial.ctx
        (sp, sn), (rp, rn) = ctx.halo_buffers()
        dev = torch.device("cuda", torch.cuda.current_device())
        send = torch.as_tensor(_DeviceArray(sp, sn), device=dev) if sn else torch.empty(0, dtype=torch.float64, device=dev)
        recv = torch.as_tensor(_DeviceArray(rp, rn), device=dev) if rn else torch.empty(0, dtype=torch.float64, device=dev)
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
        self.bind(send, recv, ctx.halo_pack, ctx.halo_unpack)

    def attach_viscosity(self, spatial):
        """The same for the per-cell viscosities (5 doubles per cell): mcfd_viscosity_halo_*."""
        ctx = spatial.ctx
        (sp, sn), (rp, rn) = ctx.viscosity_halo_buffers()
        dev = torch.device("cuda", torch.cuda.current_device())
        send = torch.as_tensor(_DeviceArray(sp, sn), device=dev) if sn else torch.empty(0, dtype=torch.float64, device=dev)
        recv = torch.as_tensor(_DeviceArray(rp, rn), device=dev) if rn else torch.empty(0, dtype=torch.float64, device=dev)
        self.bind(send, recv, ctx.viscosity_halo_pack, ctx.viscosity_halo_unpack)

    def _build_ops(self):
        ops = []
        for i, q in enumerate(self.neighbours):
            ops.append(dist.P2POp(dist.irecv, self.recv[self.recv_off[i]:self.recv_off[i + 1]], q, self.group))
        for i, q in enumerate(self.neighbours):
            ops.append(dist.P2POp(dist.isend, self.send[self.send_off[i]:self.send_off[i + 1]], q, self.group))
        return ops

    def start(self):
        """ShareGhostCellCoeffs: pack, then post all sends and receives (one NCCL group)."""
        self._pack()
        if self._ops is None:
            self._ops = self._build_ops()      # the buffers never move: build the descriptors once
        self._reqs = dist.batch_isend_irecv(self._ops) if self._ops else []
        self.inflight = True

    def finish(self):
        """UpdateGhostCellCoeffs: wait for the messages, then unpack into the ghost cells."""
        for r in self._reqs:
            r.wait()
        self._reqs = []
        self._unpack()
        self.inflight = False
        self.received = True

    def exchange(self):
        """Make the ghost cells of the current state current (reuses an exchange already in flight)."""
        if not self.inflight:
            self.start()
        self.finish()

    def drain(self):
        if self.inflight:
            self.finish()

    def allreduce_min_tensor(self, t: torch.Tensor):
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)

    def allreduce_min(self, value: float) -> float:
        """MPI_Allreduce(MIN) of the time step (demo/euler/sourceless.cpp:194)."""
        dev = self.send.device if self.send is not None else torch.device("cpu")
        t = torch.tensor([value], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return float(t.item())


def for_part(part, kfields, group=None) -> HaloExchanger:
    return HaloExchanger(part.neighbours, [len(c) for c in part.send_cells],
                         [len(c) for c in part.recv_cells], kfields, group)
