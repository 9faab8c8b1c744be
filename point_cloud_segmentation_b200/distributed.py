"""Multi-GPU voxel-map exchange: one process per GPU, frames sharded by rank, voxels owned by spatial hash.

The reference has no multi-process path (single ROS spinner thread).  The scan-level accumulation it
performs -- the octree built from the concatenation of all frames' labelled points
(pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:105-116) -- is a commutative count reduction
keyed by voxel, so it shards with exactly one exchange step:

  1. every rank integrates its own frames into a LOCAL map (pre-aggregation: tens of points per voxel per frame)
  2. local voxels are packed as 16-byte records {key48, count0 | count1 << 32}, grouped by owner = mix64(key48) % G
  3. one all-to-all of the group sizes, one all-to-all of the records (NCCL over NVLink / NVSwitch)
  4. each rank adds what it received into its OWNED map slice (atomic adds; integers, so arrival order is irrelevant)

`PeerMapExchange` is the NVLink-native form of steps 2-4: the owned maps are mapped into every process through CUDA
IPC and ONE kernel (pcs_map_push_to_owners_dev) adds each local voxel straight into its owner's table with
system-scope atomics over NVLink -- no packing, no staging buffers, no data collective; NCCL only provides the two
tiny barriers around the push.

For the mesh mask every rank needs the full map: `allgather_owned` gathers the owned slices and merges them.
torch.distributed is plumbing only (process group, NCCL calls, the exchange buffers); packing and merging are
kernels of libpcs_b200 reached through the C ABI.  The same class runs on the gloo backend with CPU tensors when the
maps are given as objects with the small interface below, which is how the host-side logic is tested without GPUs.
"""
from __future__ import annotations

import numpy as np

RECORD_BYTES = 16


def check_shared_stream(ctx, torch):
    """The exchange interleaves libpcs_b200 kernels with torch.distributed collectives; they are only ordered if both
    are enqueued on the same CUDA stream.  A context created with cuda_stream=None (or with torch's legacy default
    stream, whose handle is 0) runs on a private stream and would race with the collectives."""
    cur = torch.cuda.current_stream().cuda_stream
    if not ctx.cuda_stream or ctx.cuda_stream != cur:
        raise ValueError(
            "pcs context and torch must share one non-default CUDA stream: create `s = torch.cuda.Stream()`, "
            "`torch.cuda.set_stream(s)` and `Context(device, s.cuda_stream)` before building the exchange")


class GpuMapAdapter:
    """Adapts a capi.VoxelMap (device hash) to the exchange: records live in torch CUDA tensors."""

    def __init__(self, ctx, vmap):
        self.ctx = ctx
        self.vmap = vmap

    def size(self) -> int:
        return self.vmap.size()[0]

    def clear(self):
        self.vmap.clear()

    def pack_by_owner_into(self, n_owners: int, buf) -> np.ndarray:
        """buf: torch.int64 tensor with 2 elements per record, on this rank's device."""
        return self.vmap.pack_by_owner_dev(n_owners, buf.data_ptr(), buf.numel() // 2)

    def merge_records_from(self, buf, n: int):
        self.vmap.merge_records_dev(buf.data_ptr(), n)


class MapExchange:
    """Owner-partitioned reduction of per-rank voxel maps."""

    def __init__(self, ctx, vmap, dist, torch, capacity_voxels: int, local=None, owned=None, device=None):
        self.dist = dist
        self.torch = torch
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        if local is None:
            check_shared_stream(ctx, torch)
            local = GpuMapAdapter(ctx, vmap)
            self._owned_map = ctx.create_map(vmap.resolution, capacity_voxels)
            owned = GpuMapAdapter(ctx, self._owned_map)
            device = torch.device("cuda", ctx.device)
        self.local = local
        self.owned = owned
        self.device = device
        self._send = None
        self._recv = None
        self.last_sent_records = 0
        self.last_recv_records = 0

    def _buffer(self, which: str, n_records: int):
        cur = getattr(self, which)
        if cur is None or cur.numel() < 2 * n_records:
            cur = self.torch.empty(2 * max(n_records, 1024) * 5 // 4, dtype=self.torch.int64, device=self.device)
            setattr(self, which, cur)
        return cur

    def exchange(self, clear_owned: bool = True):
        """Route every local voxel record to its owner and merge the received ones into the owned slice."""
        torch, dist = self.torch, self.dist
        n_local = self.local.size()
        send = self._buffer("_send", n_local)
        counts = np.asarray(self.local.pack_by_owner_into(self.world, send), dtype=np.int64)
        assert int(counts.sum()) == n_local
        send_counts = torch.from_numpy(counts.copy()).to(self.device)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts)
        rc = recv_counts.cpu().numpy().astype(np.int64)
        n_recv = int(rc.sum())
        recv = self._buffer("_recv", n_recv)
        dist.all_to_all_single(recv[: 2 * n_recv], send[: 2 * n_local],
                               output_split_sizes=[int(2 * c) for c in rc],
                               input_split_sizes=[int(2 * c) for c in counts])
        if clear_owned:
            self.owned.clear()
        self.owned.merge_records_from(recv, n_recv)
        self.last_sent_records = n_local
        self.last_recv_records = n_recv
        return n_recv

    def allgather_owned(self, into):
        """Merge every rank's owned slice into `into` (same interface) so that each rank holds the whole map."""
        n, self._send, self._recv = allgather_map(self.owned, into, self.dist, self.torch, self.device, self._send, self._recv)
        return n


def _grow(torch, cur, n_records: int, device):
    if cur is None or cur.numel() < 2 * n_records:
        cur = torch.empty(2 * max(n_records, 1024) * 5 // 4, dtype=torch.int64, device=device)
    return cur


def allgather_map(owned, into, dist, torch, device, send=None, recv=None):
    """Every rank's `owned` slice (GpuMapAdapter or an object with its interface) merged into `into` on every rank: the
    gathered map the mesh mask runs on (north_star).  One all-gather of the slice sizes, one all-gather of the
    16-byte records padded to the largest slice.  Returns (total voxels, send buffer, receive buffer) -- the buffers can be
    passed back in to avoid reallocation."""
    world = dist.get_world_size()
    n_own = owned.size()
    send = _grow(torch, send, n_own, device)
    owned.pack_by_owner_into(1, send)
    sizes = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(sizes, torch.tensor([n_own], dtype=torch.int64, device=device))
    sz = sizes.cpu().numpy().astype(np.int64)
    mx = int(sz.max()) if sz.size else 0
    recv = _grow(torch, recv, mx * world, device)
    padded = send[: 2 * mx] if send.numel() >= 2 * mx else torch.cat(
        [send, torch.zeros(2 * mx - send.numel(), dtype=torch.int64, device=device)])
    dist.all_gather_into_tensor(recv[: 2 * mx * world], padded.contiguous())
    for r in range(world):
        n = int(sz[r])
        if n:
            into.merge_records_from(recv[2 * mx * r: 2 * mx * r + 2 * n], n)
    return int(sz.sum()), send, recv


def triangle_range(nt: int, world: int, rank: int):
    """Triangles [lo, hi) of rank `rank`: an even split in input order (the mask's output order is the input order)."""
    return rank * nt // world, (rank + 1) * nt // world


def gather_triangle_flags(dist, torch, device, nt: int, compute_slice):
    """maskMesh sharded over the ranks (octomap_mesh_masking.cpp:118-283 decides every triangle independently): each rank
    decides its own range of triangles against the gathered map, the flags are all-gathered, every rank returns the nt
    flags in input order.  compute_slice(lo, hi, out) must fill out[: hi - lo] (uint8 tensor on `device`) with the contact
    flags of triangles [lo, hi); on a GPU that is pcs_mesh_contacts_dev on a pointer offset by 16 * lo bytes."""
    world, rank = dist.get_world_size(), dist.get_rank()
    width = (nt + world - 1) // world  # longest slice
    mine = torch.zeros(max(width, 1), dtype=torch.uint8, device=device)
    lo, hi = triangle_range(nt, world, rank)
    if hi > lo:
        compute_slice(lo, hi, mine)
    allf = torch.empty(world * max(width, 1), dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allf, mine)
    if nt % world == 0 and nt > 0:
        return allf  # every slice has exactly `width` triangles: the gathered buffer is already in input order
    out = torch.empty(nt, dtype=torch.uint8, device=device)
    for r in range(world):
        a, b = triangle_range(nt, world, r)
        out[a:b] = allf[r * width: r * width + (b - a)]
    return out


def sharded_mesh_flags(ctx, full_map, dist, torch, verts_dev: int, nv: int, faces4_dev: int, nt: int):
    """The GPU form of gather_triangle_flags: returns a uint8 CUDA tensor of nt flags, identical on every rank."""
    device = torch.device("cuda", ctx.device)

    def compute(lo, hi, out):
        ctx.mesh_contacts_dev(full_map, verts_dev, nv, faces4_dev + 16 * lo, hi - lo, out.data_ptr())

    return gather_triangle_flags(dist, torch, device, nt, compute)


def _push(ctx, local, owners, clear_local):
    """pcs_map_push_to_owners_dev on `ctx`'s stream (a map handle is not tied to the context that created it)."""
    import ctypes as C

    from . import capi

    arr = (C.c_void_p * len(owners))(*[o.h for o in owners])
    capi._check(ctx.lib.pcs_map_push_to_owners_dev(ctx.h, local.h, arr, len(owners), int(clear_local)))


class PeerMapExchange:
    """Owner-partitioned reduction through peer memory: local voxels are pushed to their owners by one kernel.

    mode="mailbox" (default): the push kernel stores 16-byte records into the owners' inboxes (plain posted stores over
    NVLink, slots from local cursors) and adds the voxels it owns itself to its own table; after the barrier each owner
    merges its inbox with local atomics.  mode="atomics": the push kernel adds straight into the owners' tables with
    system-scope CAS / RED -- one launch less, but remote atomics cost ~1 us per thousand voxels and collapse when several
    ranks update the same voxels (overlapping scans).

    Set-up (once): every rank creates TWO owned tables, exports their CUDA IPC handles, the handles are all-gathered and
    each rank opens its peers' tables.  exchange() number e: clear the table of the other parity (it held the result of
    exchange e - 1, consumed by now), push kernel into the tables of parity e % 2, one barrier (every push has landed).
    The double buffer is what makes a single barrier enough: a peer can only push into my parity-(e+1) table after it
    has passed barrier e, which I signal after my clear of that table has been enqueued.

    The barrier is made of epoch words in peer memory: after its push rank r writes e into word r of every peer's parity-0
    table and waits for the words of all peers in its own table.  barrier="kernel" (default) does both in one 32-thread
    kernel (pcs_exchange_barrier_dev; one process per GPU, so the waiting kernel cannot starve the one it waits for);
    barrier="flags" uses stream memory operations instead (cuStreamWriteValue32 / cuStreamWaitValue32: nothing runs on an
    SM, but 2 (N-1) stream operations cost more than the one launch: 38 vs 13 us at 8 GPUs).  No collective and no host
    synchronisation either way.  barrier="allreduce" is the first form, a 4-byte NCCL all-reduce on the shared stream; it is also
    the fallback when the driver refuses the stream memory operations on IPC memory.  `owned_map` is the table holding the
    latest result; it stays valid until the next exchange() call."""

    def __init__(self, ctx, vmap, dist, torch, capacity_voxels: int, barrier: str = "kernel", clear_local: bool = False,
                 mode: str = "mailbox"):
        from . import capi

        assert mode in ("mailbox", "atomics")
        self.mode = mode  # mailbox: records stored into the owners' inboxes, merged locally; atomics: remote CAS / RED
        self.capacity = int(capacity_voxels)
        self.clear_local = clear_local  # the push kernel leaves the local map empty (saves the caller's clear)
        check_shared_stream(ctx, torch)
        self.ctx, self.local, self.dist, self.torch = ctx, vmap, dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.device = torch.device("cuda", ctx.device)
        self._mine = [ctx.create_map(vmap.resolution, capacity_voxels) for _ in range(2)]
        ctx.sync()
        mine = torch.tensor(list(self._mine[0].ipc_export() + self._mine[1].ipc_export()), dtype=torch.uint8,
                            device=self.device)
        assert self.world <= 16, "the exchange tracks at most 16 ranks"
        allh = torch.empty(self.world * 256, dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(allh, mine)
        allh = allh.cpu().numpy().tobytes()
        self._owners = [[], []]  # per parity: the table of every rank
        for r in range(self.world):
            for b in range(2):
                if r == self.rank:
                    self._owners[b].append(self._mine[b])
                else:
                    h = allh[256 * r + 128 * b: 256 * r + 128 * (b + 1)]
                    self._owners[b].append(capi.VoxelMap.ipc_open(ctx, h, vmap.resolution, capacity_voxels))
        # mailbox mode: one inbox per parity on every rank, mapped by all peers
        self._inbox, self._peer_inbox = [None, None], [[], []]
        if mode == "mailbox":
            nbytes = int(ctx.lib.pcs_map_inbox_bytes(self.capacity, self.world))
            for b in range(2):
                self._inbox[b] = ctx.dev_alloc(nbytes)
                ctx.memset(self._inbox[b], 0, 256)
            ctx.sync()
            mine_h = torch.tensor(list(ctx.ipc_export(self._inbox[0]) + ctx.ipc_export(self._inbox[1])), dtype=torch.uint8,
                                  device=self.device)
            all_h = torch.empty(self.world * 128, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(all_h, mine_h)
            all_h = all_h.cpu().numpy().tobytes()
            for b in range(2):
                for r in range(self.world):
                    self._peer_inbox[b].append(0 if r == self.rank else ctx.ipc_open(all_h[128 * r + 64 * b: 128 * r + 64 * (b + 1)]))
        self._token = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._epoch = 0
        self.owned_map = self._mine[0]
        # flag barrier: my word in every peer's parity-0 table, and the peers' words in mine
        self.barrier_kind = barrier
        if barrier in ("flags", "kernel"):
            try:
                self._their_flag = [self._owners[0][r].epoch_flag_dev(self.rank) for r in range(self.world) if r != self.rank]
                self._my_flags = [self._mine[0].epoch_flag_dev(r) for r in range(self.world) if r != self.rank]
                if barrier == "flags":  # probe: the driver must accept stream memory operations on IPC memory
                    for f in self._their_flag:
                        ctx.stream_write_u32(f, 0)
                    for f in self._my_flags:
                        ctx.stream_wait_geq_u32(f, 0)
                    ctx.sync()
                ok = 1
            except Exception:
                ok = 0
            flag = torch.tensor([ok], device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0:
                self.barrier_kind = "allreduce"
        dist.barrier()

    @property
    def owned(self):
        return GpuMapAdapter(self.ctx, self.owned_map)

    def _barrier(self):
        if self.barrier_kind == "kernel":
            self.ctx.exchange_barrier(self._their_flag, self._my_flags, self._epoch + 1)
        elif self.barrier_kind == "flags":
            e = self._epoch + 1  # epochs count from 1: the words start at 0
            for f in self._their_flag:
                self.ctx.stream_write_u32(f, e)
            for f in self._my_flags:
                self.ctx.stream_wait_geq_u32(f, e)
        else:
            self.dist.all_reduce(self._token)  # stream-ordered, no host synchronisation

    def exchange(self, mark=None, local=None):
        """`mark`, if given, is called after each phase (clear, push, barrier) -- bench.py records CUDA events there.
        `local` overrides the map that is pushed (a pipeline that alternates two local maps, one being filled while the
        other is exchanged on this exchange's own stream)."""
        b = self._epoch % 2
        self._mine[1 - b].clear()                      # result of the previous exchange: consumed, recycle it
        if mark:
            mark("clear_owned")
        src = self.local if local is None else local
        # the maps are plain device tables: every call below runs on THIS exchange's context (stream), whoever made them
        if self.mode == "mailbox":
            self.ctx.push_mailbox(src, self._mine[b], self.rank, self._peer_inbox[b], self.capacity, self.clear_local)
        else:
            _push(self.ctx, src, self._owners[b], self.clear_local)
        if mark:
            mark("push")
        self._barrier()                                # all pushes into parity b have landed everywhere
        if mark:
            mark("barrier")
        if self.mode == "mailbox":
            self.ctx.merge_inbox(self._mine[b], self._inbox[b], self.rank, self.world, self.capacity)
            if mark:
                mark("merge")
        self.owned_map = self._mine[b]
        self._epoch += 1

    def close(self):
        self.ctx.sync()
        self.dist.barrier()
        for b in range(2):
            for r, m in enumerate(self._owners[b]):
                if r != self.rank:
                    m.destroy()
        self._owners = [[], []]
        for b in range(2):
            for p in self._peer_inbox[b]:
                if p:
                    self.ctx.ipc_close(p)
        self._peer_inbox = [[], []]
        self.dist.barrier()
        for b in range(2):
            if self._inbox[b]:
                self.ctx.dev_free(self._inbox[b])
                self._inbox[b] = None
        for m in self._mine:
            m.destroy()
