"""Set-up of the peer-memory halo (mcfd_peer_*, csrc/peer.cuh): every rank exports its state buffers, epoch words and
time-step slots, the handles travel once through `torch.distributed` (any backend; objects, not tensors), and every rank
connects to its neighbours.  After that the exchange of `Part::ShareGhostCellCoeffs` / `UpdateGhostCellCoeffs`
(include/mini/mesh/part.hpp:1242-1314) and the `MPI_Allreduce(MIN)` of the time step (demo/euler/sourceless.cpp:194) run
entirely on the devices: `spatial.ctx.peer_step(order, t, dt_guess)`.

One process per GPU: CUDA IPC handles.  Several contexts in one process (tests on one GPU, a driver owning several
GPUs): `connect_local` wires them with plain pointers."""
from __future__ import annotations

import numpy as np


def connect(spatial, part, group=None):
    """Collective over `group`: wire spatial.ctx to the contexts of the neighbouring parts (one process per part)."""
    import torch.distributed as dist
    ctx = spatial.ctx
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    handles, own = ctx.peer_export(world)
    mine = {"handles": handles, "neighbours": [int(q) for q in part.neighbours],
            "recv": [np.asarray(c, dtype=np.int64) for c in part.recv_cells]}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine, group=group)
    opened = {}

    def pointer(r, which):   # which: 0..2 state, 3 epoch words, 4 dt slots
        if r == rank:
            return own[which]
        if (r, which) not in opened:
            opened[(r, which)] = ctx.peer_open(everyone[r]["handles"][which])
        return opened[(r, which)]

    state, flags, slots, dest = [], [], [], []
    for i, q in enumerate(part.neighbours):
        q = int(q)
        state.append([pointer(q, b) for b in range(3)])
        flags.append(pointer(q, 3))
        me_there = everyone[q]["neighbours"].index(rank)
        slots.append(me_there)
        ghost_ids = everyone[q]["recv"][me_there]
        assert len(ghost_ids) == len(part.send_cells[i]), "send / recv lists of two parts disagree"
        dest.append(ghost_ids)
    dest = np.concatenate(dest) if dest else np.zeros(0, dtype=np.int64)
    ctx.peer_connect(rank, world, state, flags, slots, dest, [pointer(r, 3) for r in range(world)],
                     [pointer(r, 4) for r in range(world)])
    dist.barrier(group=group)     # nobody pushes before everybody is connected


def connect_local(spatials, parts):
    """The same for contexts that live in this process (plain pointers; the devices must be peers or the same)."""
    world = len(spatials)
    ptrs = [sp.ctx.peer_export(world)[1] for sp in spatials]
    for rank, (sp, part) in enumerate(zip(spatials, parts)):
        state, flags, slots, dest = [], [], [], []
        for i, q in enumerate(part.neighbours):
            q = int(q)
            state.append(ptrs[q][:3])
            flags.append(ptrs[q][3])
            me_there = [int(x) for x in parts[q].neighbours].index(rank)
            slots.append(me_there)
            dest.append(np.asarray(parts[q].recv_cells[me_there], dtype=np.int64))
        dest = np.concatenate(dest) if dest else np.zeros(0, dtype=np.int64)
        sp.ctx.peer_connect(rank, world, state, flags, slots, dest, [p[3] for p in ptrs], [p[4] for p in ptrs])
