"""Set-up of the device-side halo exchange (mcfd_peer_*, csrc/peer.cuh): every rank exports its epoch words, time-step
slots and receive stage, the handles travel once through `torch.distributed` (any backend; objects, not tensors), and
every rank connects to its neighbours.  After that the exchange of `Part::ShareGhostCellCoeffs` /
`UpdateGhostCellCoeffs` (include/mini/mesh/part.hpp:1242-1314) and the `MPI_Allreduce(MIN)` of the time step
(demo/euler/sourceless.cpp:194) run entirely on the devices: `spatial.ctx.peer_step(order, t, dt_guess)`.

One process per GPU: CUDA IPC handles.  Several contexts in one process (tests on one GPU, a driver owning several
GPUs): `connect_local` wires them with plain pointers."""
from __future__ import annotations

import numpy as np


def _segments(part):
    """(cells in all recv lists, start of each neighbour's segment among them)"""
    counts = [len(c) for c in part.recv_cells]
    return int(sum(counts)), [int(x) for x in np.concatenate([[0], np.cumsum(counts)])[:-1]]


def connect(spatial, part, group=None):
    """Collective over `group`: wire spatial.ctx to the contexts of the neighbouring parts (one process per part)."""
    import torch.distributed as dist
    ctx = spatial.ctx
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    handles, own = ctx.peer_export(world)
    n_recv, starts = _segments(part)
    mine = {"handles": handles, "neighbours": [int(q) for q in part.neighbours], "n_recv": n_recv, "starts": starts,
            "recv_counts": [len(c) for c in part.recv_cells]}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine, group=group)
    opened = {}

    def pointer(r, which):   # which: 0 epoch words, 1 time-step slots, 2 receive stage
        if r == rank:
            return own[which]
        if (r, which) not in opened:
            opened[(r, which)] = ctx.peer_open(everyone[r]["handles"][which])
        return opened[(r, which)]

    flags, slots, recv, items, offs = [], [], [], [], []
    for i, q in enumerate(part.neighbours):
        q = int(q)
        there = everyone[q]["neighbours"].index(rank)
        assert everyone[q]["recv_counts"][there] == len(part.send_cells[i]), "send / recv lists of two parts disagree"
        flags.append(pointer(q, 0))
        slots.append(there)
        recv.append(pointer(q, 2))
        items.append(everyone[q]["n_recv"])
        offs.append(everyone[q]["starts"][there])
    ctx.peer_connect(rank, world, flags, slots, recv, items, offs, [pointer(r, 0) for r in range(world)],
                     [pointer(r, 1) for r in range(world)])
    dist.barrier(group=group)     # nobody delivers before everybody is connected


def connect_local(spatials, parts):
    """The same for contexts that live in this process (plain pointers; the devices must be peers or the same)."""
    world = len(spatials)
    ptrs = [sp.ctx.peer_export(world)[1] for sp in spatials]
    seg = [_segments(p) for p in parts]
    for rank, (sp, part) in enumerate(zip(spatials, parts)):
        flags, slots, recv, items, offs = [], [], [], [], []
        for i, q in enumerate(part.neighbours):
            q = int(q)
            there = [int(x) for x in parts[q].neighbours].index(rank)
            assert len(parts[q].recv_cells[there]) == len(part.send_cells[i])
            flags.append(ptrs[q][0]); slots.append(there); recv.append(ptrs[q][2])
            items.append(seg[q][0]); offs.append(seg[q][1][there])
        sp.ctx.peer_connect(rank, world, flags, slots, recv, items, offs, [p[0] for p in ptrs], [p[1] for p in ptrs])
