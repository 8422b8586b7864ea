"""Multi-GPU slab decomposition of the MPM substep: host side.

The reference is single-GPU (one `ti.init(arch=ti.cuda)`, dense global fields: main.py:28,
model.py:118-121); this module has no counterpart there.  One process per GPU (torchrun).  Rank r owns
the base-cell planes [bounds[r], bounds[r+1]) of the grid's slowest axis (x) and simulates the
particles whose base cell lies there with an ordinary `MPM_Simulator` over its own full-size grid
arrays (touched sparsely).  Per substep the CUDA library exchanges the partial P2G sums of the planes
shared with each neighbour (csrc/mg.cu: ncclSend/ncclRecv on the compute stream, symmetric add, both
sides then run the grid update on them).  Between batches of `migrate_every` substeps this module
moves the particles whose base cell left the slab to the neighbouring rank (torch.distributed P2P:
NCCL over NVLink on GPUs, gloo in the CPU tests) -- a particle may sit up to `ghost` cells outside
its owner's slab in between, which is what the shared region [x_hi-ghost, x_hi+ghost+2) allows for.

Everything above `SlabSimulator` is device-agnostic torch code so that the partitioning and the
migration protocol are exercised on CPU tensors by the world_size-2 gloo tests (tests/test_slabs.py).
"""
from __future__ import annotations

import os
import time

import torch
import torch.distributed as dist

_TIMING = bool(os.environ.get("MPMK_SLAB_TIMING"))
_marks = []


def _mark(name, device):
    """MPMK_SLAB_TIMING=1: wall time of each migration phase (synchronising; tuning aid only)"""
    if _TIMING:
        if device.type == "cuda":
            torch.cuda.synchronize(device)
        _marks.append((name, time.perf_counter()))


def timing_report():
    out, acc = [], {}
    for (a, ta), (b, tb) in zip(_marks, _marks[1:]):
        if b != "begin":
            acc.setdefault(b, []).append(round((tb - ta) * 1e3, 3))
    return acc


RECORD_ARRAYS = ("xm", "vv", "c0", "c1", "cs", "s1", "f0", "f1", "f2", "meta", "cv0", "cv1", "ft0", "ft1", "ft2")


class SlabError(RuntimeError):
    pass


# ---- keys ---------------------------------------------------------------------------------------
def key_base_xyz(keys: torch.Tensor, nt: int):
    """base cell (bx, by, bz) from the cell key of csrc/common.cuh make_key: tile<<6 | lx<<4 | ly<<2 | lz"""
    k = keys.to(torch.int64)
    tile = k >> 6
    tz = tile % nt
    ty = (tile // nt) % nt
    tx = tile // (nt * nt)
    return 4 * tx + ((k >> 4) & 3), 4 * ty + ((k >> 2) & 3), 4 * tz + (k & 3)


def base_cell_x(x: torch.Tensor, inv_dx: float, n_grid: int):
    """utils.py:94-95 base cell of coordinate x in fp32: (int)(x*inv_dx - 0.5), clamped like the kernels"""
    gp = x.to(torch.float32) * torch.tensor(inv_dx, dtype=torch.float32, device=x.device)
    return (gp - 0.5).to(torch.int32).clamp_(0, n_grid - 3).to(torch.int64)


# ---- slab bounds --------------------------------------------------------------------------------
def uniform_bounds(n_grid: int, lo: int, hi: int, world: int, align: int = 4):
    """`world` slabs of equal width over the occupied planes [lo, hi); the outer slabs extend to the grid
    faces.  Interior cuts are multiples of `align` (the 4-cell tile edge) so no tile straddles a cut."""
    cuts = [0]
    for r in range(1, world):
        c = lo + (hi - lo) * r / world
        cuts.append(int(round(c / align)) * align)
    cuts.append(n_grid)
    return cuts


def quantile_bounds(plane_hist: torch.Tensor, world: int, align: int = 4, min_width: int = 8):
    """cuts at particle-count quantiles of the per-plane histogram of base-cell x (SURVEY.md 7.3-7:
    compact scenes starve equal-width slabs), aligned to tiles and at least `min_width` planes wide"""
    n = plane_hist.numel()
    cum = torch.cumsum(plane_hist.to(torch.float64), 0)
    total = float(cum[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        c = int(torch.searchsorted(cum, torch.tensor(target, dtype=torch.float64, device=cum.device)).item()) + 1
        c = int(round(c / align)) * align
        c = max(c, cuts[-1] + min_width)
        cuts.append(c)
    cuts.append(n)
    for r in range(world - 1, 0, -1):  # keep the tail slabs wide enough too
        cuts[r] = min(cuts[r], cuts[r + 1] - min_width)
    if any(cuts[i + 1] - cuts[i] < min_width for i in range(world)):
        raise SlabError(f"grid of {n} planes is too small for {world} slabs of >= {min_width} planes")
    return cuts


def shared_ranges(bounds, rank: int, ghost: int, n_grid: int):
    """plane ranges whose P2G sums rank shares with rank-1 / rank+1 (mirrors mpmk_mg_set_slab)"""
    world = len(bounds) - 1
    x_lo, x_hi = bounds[rank], bounds[rank + 1]
    lo = (max(0, x_lo - ghost), min(n_grid, x_lo + ghost + 2)) if rank > 0 else (0, 0)
    hi = (max(0, x_hi - ghost), min(n_grid, x_hi + ghost + 2)) if rank < world - 1 else (0, 0)
    return lo, hi


def check_bounds(bounds, ghost: int):
    for r in range(len(bounds) - 1):
        if bounds[r + 1] - bounds[r] < 2 * ghost + 2:
            raise SlabError(f"slab {r} = [{bounds[r]},{bounds[r + 1]}) is thinner than 2*ghost+2 = {2 * ghost + 2} planes")


# ---- particle records ---------------------------------------------------------------------------
def pack_records(store: dict, idx: torch.Tensor) -> torch.Tensor:
    """rows idx of every packed array (16-byte records) plus the key, bit-cast to one int32 matrix"""
    cols = [store[a][idx].view(torch.int32) for a in RECORD_ARRAYS]
    cols.append(store["keys"][idx].view(torch.int32).reshape(-1, 1))
    return torch.cat(cols, dim=1).contiguous()


RECORD_WIDTH = 4 * len(RECORD_ARRAYS) + 1


def unpack_records(store: dict, idx: torch.Tensor, rows: torch.Tensor):
    for j, a in enumerate(RECORD_ARRAYS):
        store[a][idx] = rows[:, 4 * j:4 * j + 4].contiguous().view(store[a].dtype)
    store["keys"][idx] = rows[:, -1].contiguous().view(store["keys"].dtype)


# ---- migration ----------------------------------------------------------------------------------
def _base_z_i32(keys: torch.Tensor, nt: int):
    return ((keys >> 6) % nt) * 4 + (keys & 3)


def _base_xy_i32(keys: torch.Tensor, nt: int):
    """(bx, by) of int32 keys without widening (one pass over 4 B per particle)"""
    tile = keys >> 6
    bx = (tile // (nt * nt)) * 4 + ((keys >> 4) & 3)
    by = ((tile // nt) % nt) * 4 + ((keys >> 2) & 3)
    return bx, by


def _payload_exchange(send_lo, send_hi, n_from_lo, n_from_hi, rank, world, group, device, width=RECORD_WIDTH):
    """leavers to rank-1 / rank+1, arrivals from them; the counts are already known on every rank"""
    recv_lo = torch.empty((n_from_lo, width), dtype=torch.int32, device=device)
    recv_hi = torch.empty((n_from_hi, width), dtype=torch.int32, device=device)
    ops = []
    for peer, snd, rcv in ((rank - 1, send_lo, recv_lo), (rank + 1, send_hi, recv_hi)):
        if not 0 <= peer < world:
            continue
        if snd.shape[0] > 0:
            ops.append(dist.P2POp(dist.isend, snd, peer, group=group))
        if rcv.shape[0] > 0:
            ops.append(dist.P2POp(dist.irecv, rcv, peer, group=group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return recv_lo, recv_hi


SCAN_WIDTH = 9   # integers every rank contributes to the migration's one all_gather (TorchRecordOps.scan)


class TorchRecordOps:
    """scan / pack / refill on plain torch tensors of any device (the CPU tests run this over gloo)"""

    def __init__(self, store: dict, nt: int, n_grid: int):
        self.store, self.nt, self.n_grid = store, nt, n_grid
        self.device = store["keys"].device

    def scan(self, n, x_lo, x_hi, ghost):
        """int64[SCAN_WIDTH] on the device: leavers down, leavers up, beyond ghost, -min row, max row, max |v| component
        (float bits), -min z, max z, halo timeout epoch (0 here: no peer-memory halo without the library)"""
        if n == 0:
            self.lo_mask = self.hi_mask = torch.zeros(0, dtype=torch.bool, device=self.device)
            return torch.tensor([0, 0, 0, -self.n_grid, -1, 0, -self.n_grid, -1, 0], dtype=torch.int64, device=self.device)
        bx, by = _base_xy_i32(self.store["keys"][:n], self.nt)
        bz = _base_z_i32(self.store["keys"][:n], self.nt)
        self.lo_mask, self.hi_mask = bx < x_lo, bx >= x_hi
        far = (bx < x_lo - ghost) | (bx >= x_hi + ghost)
        vmax = self.store["vv"][:n, :3].abs().max().reshape(1).to(torch.float32).view(torch.int32)[0]
        zero = torch.zeros((), dtype=torch.int64, device=self.device)
        return torch.stack([self.lo_mask.sum(), self.hi_mask.sum(), far.sum(), -by.min().to(torch.int64),
                            by.max().to(torch.int64), vmax.to(torch.int64), -bz.min().to(torch.int64), bz.max().to(torch.int64), zero])

    def pack_leavers(self, n_lo, n_hi):
        self.idx_lo = torch.nonzero_static(self.lo_mask, size=n_lo).flatten()
        self.idx_hi = torch.nonzero_static(self.hi_mask, size=n_hi).flatten()
        return pack_records(self.store, self.idx_lo), pack_records(self.store, self.idx_hi)

    def pack_idx(self, idx):
        """records at the given slots, in that order; the slots become the holes the next refill() fills"""
        self.idx_lo, self.idx_hi = idx.to(torch.int64), torch.zeros(0, dtype=torch.int64, device=self.device)
        return pack_records(self.store, self.idx_lo)

    def refill(self, n, n_new, arrivals):
        store, device = self.store, self.device
        holes = torch.cat([self.idx_lo, self.idx_hi])
        free = [holes[holes < n_new]]
        rows_in = [arrivals]
        if n_new > n:
            free.append(torch.arange(n, n_new, device=device))
        elif n_new < n:
            tail_is_hole = torch.zeros(n - n_new, dtype=torch.bool, device=device)
            tail_is_hole[holes[holes >= n_new] - n_new] = True
            live_tail = torch.nonzero(~tail_is_hole).flatten() + n_new
            rows_in.append(pack_records(store, live_tail))
        free = torch.cat(free)
        rows_in = torch.cat(rows_in)
        if free.numel() != rows_in.shape[0]:
            raise SlabError(f"migration bookkeeping mismatch ({free.numel()} free slots, {rows_in.shape[0]} records)")
        if free.numel():
            unpack_records(store, free, rows_in)


class CudaRecordOps:
    """the same three steps as kernels of libmpmk (csrc/mg.cu) on the simulator's packed arrays: one pass for the
    scan, one launch each to pack / unpack, two to close the holes of a shrinking slab; no host synchronisation"""

    def __init__(self, backend, stream_fn):
        self.b, self._stream = backend, stream_fn
        dev = backend.device
        self.device = dev
        self.stats = torch.zeros(12, dtype=torch.int32, device=dev)
        self.idx_lo = torch.zeros(backend.capacity, dtype=torch.int32, device=dev)
        self.idx_hi = torch.zeros(backend.capacity, dtype=torch.int32, device=dev)

    def scan(self, n, x_lo, x_hi, ghost):
        self.b.ctx.call("mpmk_mg_scan", self.stats.data_ptr(), self.idx_lo.data_ptr(), self.idx_hi.data_ptr(), self._stream())
        return self.stats[:SCAN_WIDTH].to(torch.int64)

    def to_host(self, t):
        """a small device tensor on the host without a copy engine (model.py: _Backend.read_small)"""
        return self.b.read_small(t)

    # Sigma_0 (export only; an original-order table indexed by global id) travels with the particle: six more words behind
    # the record.  Every rank then only ever needs the rows of the particles it holds -- no replication of the whole table
    # at load time (an all_reduce of 24 B x N_global per reload: 793 MB at 33 M particles).
    EXTRA = 6
    ID_WORD = 4 * RECORD_ARRAYS.index("meta")

    def pack_leavers(self, n_lo, n_hi):
        self.n_lo, self.n_hi = n_lo, n_hi
        out = []
        for idx, cnt in ((self.idx_lo, n_lo), (self.idx_hi, n_hi)):
            rows = torch.empty((cnt, RECORD_WIDTH), dtype=torch.int32, device=self.device)
            self.b.ctx.call("mpmk_mg_pack", idx.data_ptr(), cnt, rows.data_ptr(), self._stream())
            if self.b.ids_init is not None:
                cov0 = self.b.init_cov[rows[:, self.ID_WORD].long()].view(torch.int32)
                rows = torch.cat([rows, cov0], dim=1)
            out.append(rows)
        return out

    def pack_idx(self, idx):
        """records (+ Sigma_0) at the given slots, in that order; the slots become the holes the next refill() fills"""
        cnt = int(idx.numel())
        self.idx_lo[:cnt] = idx.to(torch.int32)
        return self.pack_leavers(cnt, 0)[0]

    msg_records = 0   # > 0: neighbours' exchange blocks are mapped with message slots for that many records
    _tag = 0

    def exchange(self, send_lo, send_hi, n_from_lo, n_from_hi):
        """the payload as flagged 16-byte stores straight into the neighbours' message slots (mpmk_mg_ll_send / _recv):
        both sends, then both receives, all on the simulator's stream; every rank takes this path for a given migration
        or none does (the caller decides from the all-gathered table); the tag is the migration count"""
        call, st = self.b.ctx.call, self._stream()
        recv_lo = torch.empty((n_from_lo, self.width), dtype=torch.int32, device=self.device)
        recv_hi = torch.empty((n_from_hi, self.width), dtype=torch.int32, device=self.device)
        for side, snd in ((0, send_lo), (1, send_hi)):
            if snd.shape[0]:
                snd = snd.contiguous()
                call("mpmk_mg_ll_send", side, snd.data_ptr(), snd.numel(), self._tag, st)
        for side, rcv in ((0, recv_lo), (1, recv_hi)):
            if rcv.shape[0]:
                call("mpmk_mg_ll_recv", side, rcv.data_ptr(), rcv.numel(), self._tag, st)
        return recv_lo, recv_hi

    @property
    def width(self):
        """words per travelling record"""
        return RECORD_WIDTH + (self.EXTRA if self.b.ids_init is not None else 0)

    def refill(self, n, n_new, arrivals):
        call, st = self.b.ctx.call, self._stream()
        if arrivals.shape[1] > RECORD_WIDTH:     # Sigma_0 of the arrivals -> this rank's rows of the table
            self.b.init_cov[arrivals[:, self.ID_WORD].long()] = arrivals[:, RECORD_WIDTH:].contiguous().view(torch.float32)
            arrivals = arrivals[:, :RECORD_WIDTH]
        L, R = self.n_lo + self.n_hi, int(arrivals.shape[0])
        holes = torch.cat([self.idx_lo[: self.n_lo], self.idx_hi[: self.n_hi]])
        arrivals = arrivals.contiguous()
        k = min(L, R)
        if k:
            call("mpmk_mg_unpack", holes.data_ptr(), 0, k, arrivals.data_ptr(), st)
        if R > L:     # the rest is appended at [n, n_new)
            call("mpmk_mg_unpack", None, n, R - L, arrivals[L:].contiguous().data_ptr(), st)
        elif L > R:   # the slab shrinks: live records of the tail move into the unfilled holes below n_new
            rest = holes[R:].contiguous()
            nbytes = 8 + 4 * (L - R) + (n - n_new)
            scratch = torch.empty(nbytes + 16, dtype=torch.uint8, device=self.device)
            call("mpmk_mg_close_holes", rest.data_ptr(), L - R, n_new, scratch.data_ptr(), nbytes, st)


def migrate(store: dict, n: int, capacity: int, bounds, rank: int, ghost: int, nt: int, n_grid: int = None, group=None,
            ops=None):
    """Move the particles whose base-cell plane left [bounds[rank], bounds[rank+1]) to the neighbouring
    rank and take in what the neighbours send.  `store` maps the packed array names (+ "keys") to tensors
    of `capacity` records whose first `n` are live.  Holes left by leavers are refilled by arrivals and,
    if the slab shrank, by live records from the tail, so the live records stay contiguous in [0, n_new)
    (their order is irrelevant: every substep starts with a sort).

    One collective and one host synchronisation: every rank contributes {leavers down, leavers up, particles
    beyond the ghost margin, -min row, max row, max speed} to an all_gather, from which each rank reads its
    neighbours' counts, the global error flag, the global row range and the global speed bound.
    `ops` does the data movement (TorchRecordOps by default, CudaRecordOps on the simulator's arrays).
    Returns (n_new, stats)."""
    world = len(bounds) - 1
    x_lo, x_hi = bounds[rank], bounds[rank + 1]
    device = store["keys"].device
    n_grid = n_grid if n_grid is not None else 4 * nt
    if ops is None:
        ops = TorchRecordOps(store, nt, n_grid)
    _mark("begin", device)
    if hasattr(ops, "exchange"):
        ops._tag += 1          # the message tag = the migration count: in step on every rank, whoever has something to send
    mine = ops.scan(n, x_lo, x_hi, ghost)
    table = torch.empty(world * SCAN_WIDTH, dtype=torch.int64, device=device)
    if world > 1:
        dist.all_gather_into_tensor(table, mine.contiguous(), group=group)
    else:
        table[:SCAN_WIDTH] = mine
    _mark("scan", device)
    # the one synchronisation
    table = ops.to_host(table).view(world, SCAN_WIDTH) if hasattr(ops, "to_host") else table.view(world, SCAN_WIDTH).cpu()
    _mark("allgather+sync", device)
    if int(table[:, 8].max()):     # collective: every rank sees the same table, every rank raises
        bad = [r for r in range(world) if int(table[r, 8])]
        raise SlabError(f"ranks {bad}: a neighbour's halo push did not arrive within the spin budget (substep epochs "
                        f"{[int(table[r, 8]) for r in bad]}): a rank died or skipped a substep; the grid sums since then are invalid")
    if int(table[:, 2].sum()):
        bad = [r for r in range(world) if int(table[r, 2])]
        raise SlabError(f"ranks {bad}: particles drifted more than ghost={ghost} planes outside their slab between "
                        "migrations; migrate more often or raise ghost")
    n_lo, n_hi = int(table[rank, 0]), int(table[rank, 1])
    n_from_lo = int(table[rank - 1, 1]) if rank > 0 else 0
    n_from_hi = int(table[rank + 1, 0]) if rank < world - 1 else 0
    y_min, y_max = -int(table[:, 3].max()), int(table[:, 4].max())
    rows = (0, n_grid) if y_max < y_min else (max(0, y_min - ghost - 1), min(n_grid, y_max + 3 + ghost + 1))
    z_min, z_max = -int(table[:, 6].max()), int(table[:, 7].max())
    zs = (0, n_grid) if z_max < z_min else (max(0, z_min - ghost - 1), min(n_grid, z_max + 3 + ghost + 1))
    vmax = float(table[:, 5].max().to(torch.int32).reshape(1).view(torch.float32)[0])
    stats = {"sent_lo": n_lo, "sent_hi": n_hi, "recv_lo": n_from_lo, "recv_hi": n_from_hi, "rows": rows, "zs": zs, "vmax": vmax}
    if n_lo + n_hi + n_from_lo + n_from_hi == 0:
        return n, stats
    n_new = n - n_lo - n_hi + n_from_lo + n_from_hi
    if n_new > capacity:
        raise SlabError(f"rank {rank}: {n_new} particles after migration exceed the capacity {capacity}")
    send_lo, send_hi = ops.pack_leavers(n_lo, n_hi)
    _mark("pack", device)
    if getattr(ops, "msg_records", 0) and int(table[:, :2].max()) <= ops.msg_records:    # same table, same answer on every rank
        recv_lo, recv_hi = ops.exchange(send_lo, send_hi, n_from_lo, n_from_hi)
    else:
        recv_lo, recv_hi = _payload_exchange(send_lo, send_hi, n_from_lo, n_from_hi, rank, world, group, device,
                                             width=getattr(ops, "width", RECORD_WIDTH))
    _mark("p2p", device)
    ops.refill(n, n_new, torch.cat([recv_lo, recv_hi]))
    _mark("refill", device)
    return n_new, stats


def recut(store: dict, n: int, capacity: int, rank: int, world: int, ghost: int, nt: int, n_grid: int, group=None, ops=None,
          align: int = 4):
    """New slab cuts at particle-count quantiles of the CURRENT cloud (a scene that collapses or
    spreads starves equal-width slabs), and every particle to the rank that owns its base-cell plane under them.
    Unlike migrate() a particle may have to cross several ranks, so the records travel in one all_to_all; between
    re-cuts the cheap neighbour migration keeps ownership.  Collective; two host synchronisations (the global plane
    histogram, the count matrix).  Returns (n_new, new_bounds, stats)."""
    device = store["keys"].device
    if ops is None:
        ops = TorchRecordOps(store, nt, n_grid)
    rd = ops.to_host if hasattr(ops, "to_host") else (lambda t: t.cpu())
    bx = _base_xy_i32(store["keys"][:n], nt)[0].to(torch.int64) if n > 0 else torch.zeros(0, dtype=torch.int64, device=device)
    hist = torch.bincount(bx, minlength=n_grid)[:n_grid].to(torch.int64)
    if world > 1:
        dist.all_reduce(hist, group=group)
    bounds = quantile_bounds(rd(hist), world, align=align, min_width=max(2 * ghost + 2, 2 * align))
    check_bounds(bounds, ghost)
    cuts = torch.tensor(bounds[1:-1], dtype=torch.int64, device=device)
    dest = torch.searchsorted(cuts, bx, right=True)                   # plane b belongs to slab r iff bounds[r] <= b < bounds[r+1]
    idx = torch.nonzero(dest != rank).flatten()
    order = torch.argsort(dest[idx], stable=True)                     # grouped by destination rank, slot order inside a group
    idx = idx[order]
    counts = torch.cat([torch.bincount(dest[idx], minlength=world).to(torch.int64),
                        torch.tensor([n, capacity], dtype=torch.int64, device=device)])
    table = torch.empty(world * (world + 2), dtype=torch.int64, device=device)
    if world > 1:
        dist.all_gather_into_tensor(table, counts.contiguous(), group=group)
    else:
        table[: world + 2] = counts
    table = rd(table).view(world, world + 2)                          # table[src, dst], then the source's count and capacity
    send = [int(c) for c in table[rank, :world]]
    recv = [int(table[r, rank]) for r in range(world)]
    n_new = n - sum(send) + sum(recv)
    stats = {"sent": sum(send), "received": sum(recv), "bounds": bounds}
    # collective: every rank sees the same table, every rank raises (one rank leaving alone would hang the others' all_to_all)
    after = [int(table[r, world]) - int(table[r, :world].sum()) + int(table[:, r].sum()) for r in range(world)]
    over = [r for r in range(world) if after[r] > int(table[r, world + 1])]
    if over:
        raise SlabError(f"ranks {over}: {[after[r] for r in over]} particles after the re-cut exceed the capacities "
                        f"{[int(table[r, world + 1]) for r in over]} (construct the SlabSimulator with a larger capacity)")
    table = table[:, :world]
    if int(table.sum()) == 0:
        return n, bounds, stats
    rows = ops.pack_idx(idx)
    arrivals = torch.empty((sum(recv), rows.shape[1]), dtype=torch.int32, device=device)
    if world > 1:
        dist.all_to_all_single(arrivals, rows.contiguous(), output_split_sizes=recv, input_split_sizes=send, group=group)
    ops.refill(n, n_new, arrivals)
    return n_new, bounds, stats


def global_row_range(store: dict, n: int, nt: int, n_grid: int, margin: int, group=None, axis="y", to_host=None):
    """[y_lo, y_hi): rows of the shared planes any rank's stencils can reach before the next migration (axis="z": the
    same along the fastest axis)"""
    device = store["keys"].device
    if n > 0:
        by = _base_xy_i32(store["keys"][:n], nt)[1] if axis == "y" else _base_z_i32(store["keys"][:n], nt)
        lo, hi = by.min().to(torch.int64), by.max().to(torch.int64)
    else:
        lo = torch.tensor(n_grid, dtype=torch.int64, device=device)
        hi = torch.tensor(-1, dtype=torch.int64, device=device)
    v = torch.stack([-lo, hi]).to(torch.int64)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(v, op=dist.ReduceOp.MAX, group=group)
    v = to_host(v) if to_host is not None else v.cpu()
    lo, hi = -int(v[0]), int(v[1])
    if hi < lo:
        return 0, n_grid
    return max(0, lo - margin), min(n_grid, hi + 3 + margin)


# ---- the per-rank simulator -----------------------------------------------------------------------
class SlabSimulator:
    """One rank's share of a slab-decomposed simulation: an `MPM_Simulator` over the local particles plus
    the halo/migration plumbing.  Construct it on every rank of an initialised torch.distributed group
    (backend nccl) with that rank's particles:

        ids      global particle ids of the local particles (int, unique across ranks)
        n_global total number of particles over all ranks
        bounds   world+1 plane cuts (uniform_bounds / quantile_bounds), identical on every rank
    """

    def __init__(self, xyz, cov, vol, args, ids, n_global, bounds, vel=None, group=None, ghost=1, migrate_every=10,
                 capacity=None, boundary_conditions=None, collider=None, **sim_kw):
        from . import _native
        from .solver import MPM_Simulator
        from .utils import _stream
        self._native, self._stream = _native, _stream
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.bounds = [int(b) for b in bounds]
        if len(self.bounds) != self.world + 1:
            raise SlabError("bounds must hold world+1 cuts")
        self.ghost = int(ghost)
        check_bounds(self.bounds, self.ghost)
        self.migrate_every = int(migrate_every) if migrate_every not in (None, 0, "auto") else 0
        self.n_global = int(n_global)
        n = int(xyz.shape[0])
        cap = int(capacity if capacity is not None else max(int(1.25 * n) + 4096, 8192))
        self.sim = MPM_Simulator(xyz, cov, vol, args, capacity=cap, ids=ids, n_global=n_global, **sim_kw)
        b = self.backend = self.sim._backend
        self.device = b.device
        self.n_grid = int(args.n_grid)
        self.nt = (self.n_grid + 3) // 4
        if vel is not None and n:
            b.cur["vv"][:n, :3] = vel.to(device=self.device, dtype=torch.float32)
        if boundary_conditions:
            self.sim.set_boundary_conditions(boundary_conditions, args)
        if collider:
            self.sim.add_surface_collider(collider["point"], collider["normal"], friction=collider.get("friction", 0.0))
        # communicator of the halo data path: rank 0 makes the NCCL id, torch.distributed carries it
        lib_path = _native.nccl_library_path().encode()
        uid = torch.zeros(_native.NCCL_ID_BYTES, dtype=torch.uint8)
        if self.rank == 0:
            import ctypes as C
            buf = (C.c_char * _native.NCCL_ID_BYTES)()
            rc = _native.lib().mpmk_mg_unique_id(lib_path, C.cast(buf, C.c_void_p), _native.NCCL_ID_BYTES)
            if rc != 0:
                raise SlabError("mpmk_mg_unique_id failed")
            uid = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        uid = uid.to(self.device)
        dist.broadcast(uid, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self._uid = bytes(uid.cpu().numpy().tobytes())
        b.ctx.call("mpmk_mg_init", lib_path, self._uid, _native.NCCL_ID_BYTES, self.rank, self.world)
        planes = 2 * self.ghost + 2
        z = lambda: torch.zeros((planes * self.n_grid * self.n_grid, 4), dtype=torch.float32, device=self.device)
        self._recv_lo = z() if self.rank > 0 else None
        self._recv_hi = z() if self.rank < self.world - 1 else None
        self.halo_path = "nccl"
        if os.environ.get("MPMK_MG_PEER", "1") != "0":
            self._open_peer_halo(planes)
        self._ops = CudaRecordOps(b, _stream)
        if self.halo_path == "peer" and os.environ.get("MPMK_MG_LL_MIGRATE", "1") != "0":
            self._ops.msg_records = self._msg_records     # the payload travels as flagged stores over NVLink (mpmk_mg_ll_send)
        self._auto = migrate_every in (None, 0, "auto")
        self.max_interval = 64
        self._dt = float(args.substep_dt)
        if self._auto:
            self.migrate_every = 1   # until the first migration has measured the speeds
        self.stats = {"migrations": 0, "sent": 0, "received": 0}
        self._since_migration = 0
        self._warm_up_p2p()
        self._set_slab()

    # -- plumbing ---------------------------------------------------------------------------------
    @property
    def n_particles(self):
        return self.backend.n

    def _store(self):
        b = self.backend
        st = dict(b.cur)
        st["ft0"], st["ft1"], st["ft2"] = b.ft   # F_trial of the last substep travels with the particle (export)
        st["keys"] = b.keys
        return st

    def _open_peer_halo(self, planes):
        """Peer-memory halo path (csrc/mg.cu): every rank exports its exchange block as a CUDA IPC handle, the handles
        travel once over torch.distributed, the neighbours map them; from then on the per-substep halo sum is stores
        over NVLink into the neighbour's memory plus a flag -- no NCCL call on the substep path.  One node only."""
        import ctypes as C
        b = self.backend
        hbuf = (C.c_char * 64)()
        ok = 1
        try:
            # message slots for the migration payload: 1/32 of a rank's share per neighbour and migration (a migration
            # that moves more goes through NCCL instead), LL-encoded = 8 bytes per 4-byte word
            self._msg_records = max(32768, (self.n_global // self.world) // 32)
            msg_bytes = ((self._msg_records * (RECORD_WIDTH + CudaRecordOps.EXTRA) + 1) // 2) * 16
            b.ctx.call("mpmk_mg_peer_alloc_ex", planes * self.n_grid * self.n_grid * 16, msg_bytes, C.cast(hbuf, C.c_void_p), 64)
        except _native.MpmkError:
            ok = 0
        mine = torch.frombuffer(bytearray(hbuf.raw), dtype=torch.uint8).clone().to(self.device)
        handles = [torch.empty(64, dtype=torch.uint8, device=self.device) for _ in range(self.world)]
        dist.all_gather(handles, mine, group=self.group)
        raw = lambda r: handles[r].cpu().numpy().tobytes()
        lo = raw(self.rank - 1) if self.rank > 0 else None
        hi = raw(self.rank + 1) if self.rank < self.world - 1 else None
        if ok:
            try:
                b.ctx.call("mpmk_mg_peer_open", lo, hi)
            except _native.MpmkError:      # no peer access between these two devices
                ok = 0
        # every rank or none: the two sides of a halo must use the same path
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 1:
            self.halo_path = "peer"
        else:
            b.ctx.call("mpmk_mg_peer_close")
        dist.barrier(group=self.group)   # nobody pushes before every mapping exists

    def _warm_up_p2p(self):
        """one dummy record each way now, so that the first real migration does not pay for setting up the
        point-to-point connections (~0.1 s with NCCL)"""
        dummy = torch.zeros((1, RECORD_WIDTH), dtype=torch.int32, device=self.device)
        n_lo = 1 if self.rank > 0 else 0
        n_hi = 1 if self.rank < self.world - 1 else 0
        _payload_exchange(dummy[:n_lo], dummy[:n_hi], n_lo, n_hi, self.rank, self.world, self.group, self.device)

    def _choose_interval(self, vmax):
        """migrate_every="auto": as many substeps as the fastest particle needs to cross half the ghost margin"""
        if not self._auto:
            return
        dx = self.sim.mpm_model.dx
        per_substep = max(vmax * self._dt / dx, 1e-9)
        self.migrate_every = int(max(1, min(self.max_interval, 0.5 * self.ghost / per_substep)))

    def _set_slab(self, rows=None, zs=None):
        b = self.backend
        rd = self._ops.to_host
        y_lo, y_hi = rows if rows is not None else global_row_range(self._store(), b.n, self.nt, self.n_grid, self.ghost + 1, self.group, to_host=rd)
        z_lo, z_hi = zs if zs is not None else global_row_range(self._store(), b.n, self.nt, self.n_grid, self.ghost + 1, self.group, axis="z", to_host=rd)
        self.rows, self.zs = (y_lo, y_hi), (z_lo, z_hi)
        b.ctx.call("mpmk_mg_set_slab", self.bounds[self.rank], self.bounds[self.rank + 1], self.ghost, y_lo, y_hi,
                   self._recv_lo.data_ptr() if self._recv_lo is not None else None,
                   self._recv_hi.data_ptr() if self._recv_hi is not None else None)
        b.ctx.call("mpmk_mg_set_window", z_lo, z_hi)

    def migrate(self):
        b = self.backend
        _mark("begin", self.device)
        n_new, st = migrate(self._store(), b.n, b.capacity, self.bounds, self.rank, self.ghost, self.nt, self.n_grid, self.group,
                            ops=self._ops)
        b.n = n_new
        self.sim.n_particles = n_new
        b.ctx.call("mpmk_mg_set_count", n_new)
        b.invalidate()
        self.stats["migrations"] += 1
        self.stats["sent"] += st["sent_lo"] + st["sent_hi"]
        self.stats["received"] += st["recv_lo"] + st["recv_hi"]
        self._since_migration = 0
        self._choose_interval(st["vmax"])
        self._set_slab(st["rows"], st["zs"])
        _mark("set_slab", self.device)
        return st

    def recut(self):
        """re-balance: new cuts at particle-count quantiles of the current cloud, every particle to its new owner (one
        all_to_all; module-level recut()).  Call it between frames when the counts have drifted apart -- `imbalance()` says
        by how much; the per-substep work of a rank is proportional to its count, and the halo makes every rank wait for
        the slowest.  Returns the new bounds."""
        b = self.backend
        n_new, bounds, st = recut(self._store(), b.n, b.capacity, self.rank, self.world, self.ghost, self.nt, self.n_grid, self.group,
                                  ops=self._ops)
        self.bounds = list(bounds)
        b.n = n_new
        self.sim.n_particles = n_new
        b.ctx.call("mpmk_mg_set_count", n_new)
        b.invalidate()
        self.stats["sent"] += st["sent"]
        self.stats["received"] += st["received"]
        self.stats["recuts"] = self.stats.get("recuts", 0) + 1
        self._since_migration = 0
        self._set_slab()
        return self.bounds

    def imbalance(self):
        """max over ranks of the local particle count / the mean count (1.0 = perfectly balanced); collective"""
        t = torch.tensor([float(self.backend.n)], dtype=torch.float64, device=self.device)
        counts = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(counts, t, group=self.group)
        c = self._ops.to_host(torch.cat(counts))
        return float(c.max() / max(float(c.mean()), 1.0))

    def reload(self, xyz, cov, vol, vel, ids):
        """re-initialise this rank's particles in place from new data (the constructor's state
        initialisation, model.py:101-115, without reallocating); registered grid operations stay"""
        b = self.backend
        n = int(xyz.shape[0])
        if n > b.capacity:
            raise SlabError("reload: more particles than capacity")
        b.n = n
        self.sim.n_particles = n
        b.ids_init = ids.to(device=self.device, dtype=torch.int32)
        self.sim.mpm_state._init_particles(xyz, cov, vol)
        self.sim.mpm_model._reinit_packed()
        b.cur["vv"][:n, :3] = vel.to(device=self.device, dtype=torch.float32)
        self.sim.time = 0.0
        self.sim._sync_flags()
        self._since_migration = 0
        self._set_slab()

    # -- stepping ---------------------------------------------------------------------------------
    def step(self, n_substeps, dt=None):
        """n_substeps substeps (MPM_Simulator.step, whose substeps now include the halo sum), with a
        migration after every `migrate_every` of them"""
        done = 0
        while done < n_substeps:
            k = min(n_substeps - done, self.migrate_every - self._since_migration)
            if dt is not None:
                self._dt = float(dt)
            self.sim.step(k, dt)
            done += k
            self._since_migration += k
            if self._since_migration >= self.migrate_every:
                self.migrate()

    def p2g2p(self, dt):
        self.step(1, dt)

    # -- results ----------------------------------------------------------------------------------
    def gather(self):
        """(ids, xyz, vel, F) of the local particles in physical order"""
        b = self.backend
        n, cur = b.n, b.cur
        F = torch.cat([cur["f0"][:n], cur["f1"][:n], cur["f2"][:n, :1]], dim=1).reshape(n, 3, 3)
        return cur["meta"][:n, 0].long(), cur["xm"][:n, :3].clone(), cur["vv"][:n, :3].clone(), F

    def export(self):
        """postprocess() on every rank, summed into complete (xyz [N,3], cov [N,6], R [N,3,3]) tensors in
        global id order on every rank (each id is written by exactly one rank; the others hold zeros)"""
        b = self.backend
        for t in (b.out_xyz, b.out_cov, b.out_R):
            t.zero_()
        b.export(want_R=True)
        for t in (b.out_xyz, b.out_cov, b.out_R):
            dist.all_reduce(t, group=self.group)
        return b.out_xyz, b.out_cov, b.out_R

    def global_count(self):
        t = torch.tensor([self.backend.n], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, group=self.group)
        return int(t.item())

    def close(self):
        self.backend.ctx.call("mpmk_mg_finalize")


def partition_cloud(cloud: dict, bounds, rank: int, inv_dx: float, n_grid: int):
    """the rows of a host-side cloud (numpy dict xyz/vel/cov/vol) whose base-cell plane lies in this rank's
    slab, plus their global ids"""
    import numpy as np
    bx = base_cell_x(torch.from_numpy(cloud["xyz"][:, 0].copy()), inv_dx, n_grid).numpy()
    sel = np.nonzero((bx >= bounds[rank]) & (bx < bounds[rank + 1]))[0]
    return {k: v[sel] for k, v in cloud.items()}, sel.astype(np.int64)
