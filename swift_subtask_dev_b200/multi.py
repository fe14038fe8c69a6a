"""Multi-GPU driver of the SPH neighbour-loop path: spatial decomposition of the
top-level cell grid with halo refresh between the phases.

Role in the reference: partition_initial_partition (src/partition.c:1914-1950,
grid of ranks over the top-level cells), engine_makeproxies
(src/engine_proxy.c:41), and the send/recv tasks of
engine_addtasks_send_hydro / _recv_hydro (src/engine_maketasks.c:141-324,
567-): `xv` before the density loop, `rho` after the ghost.

One process per GPU.  Every rank holds the particles of the top-level cells it
owns (ACTIVE) plus the particles of the neighbouring ranks' adjacent top-level
cells (its halo, INACTIVE on this rank: they are read by the loops, never
updated; reference rule, src/scheduler.c:1275, tests/testHydroMPIrules.c).  Only
the halo particles that an interaction of an owned leaf actually reads
(swift_cuda_get_touched) are refreshed: packed on the owner's GPU
(swift_cuda_halo_pack: 48 B `xv`, 24 B `rho` per particle instead of the
reference's 128-byte struct part), moved with NCCL send/recv over NVLink, and
unpacked on the receiver's GPU.  There is no collective on the data path.
The results of the owned particles are bit-identical to a single-GPU run.
"""
import numpy as np

from . import ics
from .backend import HydroBackend, make_params


# --------------------------------------------------------------------------- partition
class Partition:
    """gx x gy x gz blocks of top-level cells; nodeID = ix + gx (iy + gy iz) as
    partition_initial_partition's INITPART_GRID."""

    def __init__(self, cdim, grid, periodic=True, weights=None):
        """weights=None: equal blocks (the grid must divide cdim).  weights = cost of every top-level
        cell (flat, x-major): the cuts of the gx x gy x gz blocks are moved so that every block
        carries the same cost (recursive bisection: x, then y inside each x slab, then z inside
        each column) -- the role of the reference's weighted repartition (src/partition.c:419-530,
        1382: METIS vertex weights from particle counts / task costs) for the clustered config."""
        self.cdim = tuple(int(c) for c in cdim)
        self.grid = tuple(int(g) for g in grid)
        self.periodic = periodic
        self.nranks = self.grid[0] * self.grid[1] * self.grid[2]
        ijk = np.stack(np.meshgrid(*[np.arange(c) for c in self.cdim], indexing="ij"), axis=-1).reshape(-1, 3)
        self.ijk = ijk
        if weights is None:
            for c, g in zip(self.cdim, self.grid):
                if c % g != 0:
                    raise ValueError("grid %s does not divide the top-level cells %s" % (grid, cdim))
            blk = [ijk[:, d] // (self.cdim[d] // self.grid[d]) for d in range(3)]
        else:
            w = np.asarray(weights, dtype=np.float64).reshape(self.cdim)
            blk = self._weighted_blocks(w)
        self.owner = (blk[0] + self.grid[0] * (blk[1] + self.grid[1] * blk[2])).astype(np.int32)

    @staticmethod
    def _cuts(w1d, parts):
        """Block index of every slice of a 1-D cost profile: `parts` contiguous blocks of equal cost,
        each at least one slice wide."""
        n = len(w1d)
        if parts > n:
            raise ValueError("more blocks than top-level cells along an axis")
        cum = np.cumsum(w1d)
        total = cum[-1] if cum[-1] > 0 else 1.0
        idx = np.zeros(n, dtype=np.int64)
        start = 0
        for b in range(parts - 1):
            target = total * (b + 1) / parts
            end = int(np.searchsorted(cum, target, side="left")) + 1   # first `end` slices reach the target
            if end - 1 > start and target - cum[end - 2] < cum[end - 1] - target:
                end -= 1                                               # the nearer of the two cuts
            end = min(max(end, start + 1), n - (parts - 1 - b))
            idx[end:] = b + 1
            start = end
        return idx

    def _weighted_blocks(self, w):
        gx, gy, gz = self.grid
        cx, cy, cz = self.cdim
        bi = np.zeros(self.cdim, dtype=np.int64)
        bj = np.zeros(self.cdim, dtype=np.int64)
        bk = np.zeros(self.cdim, dtype=np.int64)
        ix = self._cuts(w.sum(axis=(1, 2)), gx)
        for a in range(gx):
            sx = ix == a
            iy = self._cuts(w[sx].sum(axis=(0, 2)), gy)
            for b in range(gy):
                sy = iy == b
                sub = w[np.ix_(sx, sy, np.ones(cz, dtype=bool))]
                iz = self._cuts(sub.sum(axis=(0, 1)), gz)
                m = np.ix_(sx, sy, np.ones(cz, dtype=bool))
                bi[m] = a
                bj[m] = b
                bk[m] = iz[None, None, :]
        return [bi.reshape(-1), bj.reshape(-1), bk.reshape(-1)]

    def flat(self, i, j, k):
        return (i * self.cdim[1] + j) * self.cdim[2] + k

    def owned_cells(self, rank):
        return np.nonzero(self.owner == rank)[0]

    def halo_cells(self, rank):
        """Foreign top-level cells adjacent (26 neighbours, periodic wrap) to a cell of `rank`."""
        own = self.owned_cells(rank)
        out = set()
        for c in own:
            i, j, k = self.ijk[c]
            for di in (-1, 0, 1):
                for dj in (-1, 0, 1):
                    for dk in (-1, 0, 1):
                        n = [i + di, j + dj, k + dk]
                        ok = True
                        for d in range(3):
                            if n[d] < 0 or n[d] >= self.cdim[d]:
                                if self.periodic:
                                    n[d] %= self.cdim[d]
                                else:
                                    ok = False
                        if ok:
                            f = self.flat(*n)
                            if self.owner[f] != rank:
                                out.add(int(f))
        return np.array(sorted(out), dtype=np.int64)


FOREIGN_BIN = 64   # SWIFT_CUDA_FOREIGN_BIN of the library (csrc/pack.cuh: kBinMask)


def grid_for(nranks):
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[nranks]


def local_particles(part, rank, dim, per_cell, seed=42):
    """Owned particles first, then the halo ones; inside each group cell by cell in the order of
    the global cell index (the particles of a cell come in the same relative order on every rank
    and in a single-GPU run; owned-first lets the host move only its own particles over PCIe:
    swift_cuda_set_transfer_count).  A halo copy carries its owner's time bin + FOREIGN_BIN: never
    active here, while the owner's real bin still enters the neighbour-bin limit of the particles that
    see it (runner_iact_nonsym_timebin)."""
    own = part.owned_cells(rank)
    halo = part.halo_cells(rank)
    cells = np.concatenate([np.sort(own), np.sort(halo)])
    ic = ics.cell_block_random(cells, part.cdim, dim, per_cell, seed=seed)
    is_own = part.owner[ic["cell"]] == rank
    ic["time_bin"] = np.where(is_own, 1, 1 + FOREIGN_BIN).astype(np.int8)
    ic["owner"] = part.owner[ic["cell"]]
    return ic


def top_cell_of(x, cdim, dim):
    """Flat top-level cell of every position, with the library's own arithmetic
    (host_tree.cpp: idx = (int)(x * (1 / width)), clamped)."""
    idx = []
    for d in range(3):
        iw = 1.0 / (dim[d] / cdim[d])
        idx.append(np.clip((x[:, d] * iw).astype(np.int64), 0, cdim[d] - 1))
    return (idx[0] * cdim[1] + idx[1]) * cdim[2] + idx[2]


def local_particles_from(ic, part, rank, dim):
    """The same for an arbitrary particle set `ic` (every rank holds or generates the whole set):
    the particles of the rank's own top-level cells first, then those of its halo cells, each group
    in the order of the global arrays."""
    cell = top_cell_of(ic["x"], part.cdim, dim)
    own_mask = part.owner[cell] == rank
    halo_cells = part.halo_cells(rank)
    is_halo = np.zeros(int(np.prod(part.cdim)), dtype=bool)
    is_halo[halo_cells] = True
    sel = np.concatenate([np.nonzero(own_mask)[0], np.nonzero(is_halo[cell] & ~own_mask)[0]])
    out = {k: (np.ascontiguousarray(a[sel]) if isinstance(a, np.ndarray) and len(a) == len(cell) else a)
           for k, a in ic.items()}
    out["cell"] = cell[sel]
    out["owner"] = part.owner[out["cell"]]
    base = out["time_bin"].astype(np.int64) if out.get("time_bin") is not None else np.ones(len(sel), dtype=np.int64)
    out["time_bin"] = (base + FOREIGN_BIN * (out["owner"] != rank)).astype(np.int8)
    return out


# --------------------------------------------------------------------------- transports
def plan_recv(ic, need):
    """{peer: (ids, caller indices)} of the halo particles flagged in `need`, ids ascending."""
    out = {}
    for peer in np.unique(ic["owner"][need]):
        sel = np.nonzero(need & (ic["owner"] == peer))[0]
        sel = sel[np.argsort(ic["id"][sel])]
        out[int(peer)] = (ic["id"][sel], sel)
    return out


def match_send(ic, is_own, wanted):
    """{peer: caller indices} of my owned particles in the order of the ids each peer asked for."""
    my_ids = ic["id"]
    srt = np.argsort(my_ids)
    out = {}
    for peer, ids in wanted.items():
        pos = srt[np.searchsorted(my_ids[srt], ids)]
        if not (np.array_equal(my_ids[pos], ids) and is_own[pos].all()):
            raise ValueError("rank %d asked for particles this rank does not own" % peer)
        out[int(peer)] = pos
    return out


class NcclTransport:
    """torch.distributed: NCCL send/recv for the halo buffers, a gloo group for the set-up lists."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.rank = dist.get_rank()
        self.nranks = dist.get_world_size()
        self.setup_group = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else None

    def all_to_all_objects(self, objs):
        """objs: {peer: object}; returns {src: object} (one-off, host side)."""
        out = [None] * self.nranks
        send = [objs.get(p) for p in range(self.nranks)]
        gathered = [None] * self.nranks
        self.dist.all_gather_object(gathered, send, group=self.setup_group)
        for src in range(self.nranks):
            out[src] = gathered[src][self.rank]
        return {src: o for src, o in enumerate(out) if o is not None and src != self.rank}

    def exchange(self, send, recv):
        """send / recv: {peer: cuda uint8 tensor}.  Grouped point-to-point, no collective.
        The NCCL kernels order themselves with torch's CURRENT stream: call this under the stream the
        pack / unpack kernels are launched on (RankState.stream)."""
        dist = self.dist
        ops = []
        for peer in sorted(set(send) | set(recv)):
            if peer in recv and recv[peer].numel():
                ops.append(dist.P2POp(dist.irecv, recv[peer], peer))
            if peer in send and send[peer].numel():
                ops.append(dist.P2POp(dist.isend, send[peer], peer))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()


# --------------------------------------------------------------------------- one rank
class RankState:
    """The backend of one rank plus its halo plan."""

    def __init__(self, rank, part, dim, per_cell, seed=42, device=0, params_kw=None, ic=None):
        """ic=None: the uniform per-cell generator (config 4); else the rank's particles as
        local_particles_from gives them."""
        import torch
        self.torch = torch
        self.rank = rank
        self.part = part
        self.device = device
        self.ic = local_particles(part, rank, dim, per_cell, seed=seed) if ic is None else ic
        kw = dict(dim=dim, periodic=part.periodic, max_active_bin=56)
        kw.update(params_kw or {})
        # SPHENIX: three exchanges per step (halo phases 0 xv, 3 rho + u + alpha, 2 the alphas after the extra ghost)
        self.sphenix = kw.get("scheme", 0) == 2
        self.rho_phase = 3 if self.sphenix else 1
        self.be = HydroBackend(make_params(**kw), device=device)
        self.be.build(self.ic)
        tree = self.be.tree()
        assert tuple(tree["cdim"]) == tuple(part.cdim), (tree["cdim"], part.cdim)
        order = tree["order"]
        self.slot_of = np.empty(len(order), dtype=np.int64)   # caller index -> cell-order slot
        self.slot_of[order] = np.arange(len(order))
        touched = self.be.touched_mask()
        self.is_own = self.ic["time_bin"] < FOREIGN_BIN
        need = touched & ~self.is_own
        # what I need from each peer: ids, sorted, and where they live on my device
        self.recv_ids, self.recv_idx = {}, {}
        for peer, (ids, sel) in plan_recv(self.ic, need).items():
            self.recv_ids[peer] = ids
            self.recv_idx[peer] = self._dev_index(self.slot_of[sel])
        self.send_idx = {}
        self.bufs = {}
        # The halo stream: pack -> NCCL send/recv -> unpack are ordered on it.  It must not be torch's
        # default stream: its handle is 0, which swift_cuda_halo_pack/unpack read as "the library's own
        # (non-blocking) stream", and the kernels would then race with the NCCL transfers.
        with torch.cuda.device(device):
            self.stream = torch.cuda.Stream(device=device)
        assert self.stream.cuda_stream != 0

    def _dev_index(self, slots):
        return self.torch.as_tensor(np.ascontiguousarray(slots, dtype=np.int32), device="cuda:%d" % self.device)

    def set_send_lists(self, wanted):
        """wanted: {peer: sorted ids the peer needs from me}."""
        for peer, pos in match_send(self.ic, self.is_own, wanted).items():
            self.send_idx[peer] = self._dev_index(self.slot_of[pos])

    def buffers(self, phase):
        if phase not in self.bufs:
            nb = self.be.halo_bytes(phase)
            dev = "cuda:%d" % self.device
            send = {p: self.torch.empty(len(ix) * nb, dtype=self.torch.uint8, device=dev)
                    for p, ix in self.send_idx.items()}
            recv = {p: self.torch.empty(len(ix) * nb, dtype=self.torch.uint8, device=dev)
                    for p, ix in self.recv_idx.items()}
            self.bufs[phase] = (send, recv)
        return self.bufs[phase]

    def pack(self, phase):
        send, _ = self.buffers(phase)
        st = self.stream.cuda_stream
        for p, ix in self.send_idx.items():
            self.be.halo_pack(phase, ix.data_ptr(), len(ix), send[p].data_ptr(), st)
        return send

    def unpack(self, phase):
        _, recv = self.buffers(phase)
        st = self.stream.cuda_stream
        for p, ix in self.recv_idx.items():
            self.be.halo_unpack(phase, ix.data_ptr(), len(ix), recv[p].data_ptr(), st)
        self.stream.synchronize()   # the library's stream reads the halo copies next
        if phase != 2:              # (the gradient halo carries no smoothing lengths)
            self.be.refresh_cell_hmax(after_ghost=(phase != 0))

    def halo_bytes_per_step(self):
        per = self.be.halo_bytes(0) + self.be.halo_bytes(self.rho_phase) + (self.be.halo_bytes(2) if self.sphenix else 0)
        return sum(len(ix) for ix in self.send_idx.values()) * per

    def owned_results(self):
        out = self.be.results()
        if self.sphenix:
            out.update({k: v for k, v in self.be.sphenix_results().items() if k != "n_gradient"})
        sel = self.is_own
        res = {k: v[sel] for k, v in out.items()}
        res["id"] = self.ic["id"][sel]
        return res


def step_ranks_local(states):
    """One density -> ghost -> force step of all ranks living in this process (LocalTransport)."""
    def move(phase):
        for s in states:
            s.pack(phase)
        for s in states:
            s.stream.synchronize()
        for s in states:
            _, recv = s.buffers(phase)
            with s.torch.cuda.stream(s.stream):
                for peer, buf in recv.items():
                    buf.copy_(states[peer].buffers(phase)[0][s.rank])
        for s in states:
            s.unpack(phase)
    for s in states:
        s.be.restore_particles()
    move(0)
    for s in states:
        s.be.runner_do_hydro_sort()
        s.be.runner_do_density()
        s.be.runner_do_ghost()
    move(states[0].rho_phase)
    if states[0].sphenix:
        for s in states:
            s.be.runner_do_gradient()
            s.be.runner_do_extra_ghost()
        move(2)
    for s in states:
        s.be.runner_do_force()


def setup_local(nranks, dim, cdim, per_cell, seed=42, periodic=True, params_kw=None, ic=None, weighted=False):
    """All ranks in this process on device 0 (tests).  ic: a whole particle set to be split (config 5
    shape), weighted: balance the particle counts of the blocks."""
    weights = None
    if ic is not None and weighted:
        weights = np.bincount(top_cell_of(ic["x"], cdim, dim), minlength=int(np.prod(cdim)))
    part = Partition(cdim, grid_for(nranks), periodic, weights=weights)
    states = [RankState(r, part, dim, per_cell, seed=seed, device=0, params_kw=params_kw,
                        ic=None if ic is None else local_particles_from(ic, part, r, dim)) for r in range(nranks)]
    for s in states:
        s.set_send_lists({r.rank: r.recv_ids[s.rank] for r in states if s.rank in r.recv_ids})
    return part, states


# --------------------------------------------------------------------------- torchrun entry
PHASES = ("halo_xv", "sort", "density", "ghost", "halo_rho", "force")


def step_rank_nccl(state, transport, phase_ms=None):
    """One step of this process's rank: xv halo, sort + density + ghost, rho halo, force.
    phase_ms: dict to add the wall time of every phase to (each phase ends synchronised)."""
    import time
    be = state.be
    t = [time.perf_counter()]

    def lap(name):
        if phase_ms is not None:
            state.torch.cuda.synchronize()
            t.append(time.perf_counter())
            phase_ms[name] = phase_ms.get(name, 0.0) + 1e3 * (t[-1] - t[-2])
    be.restore_particles()
    send = state.pack(0)
    _, recv = state.buffers(0)
    with state.torch.cuda.stream(state.stream):
        transport.exchange(send, recv)
    state.unpack(0)
    lap("halo_xv")
    be.runner_do_hydro_sort()
    lap("sort")
    be.runner_do_density()
    lap("density")
    be.runner_do_ghost()
    lap("ghost")
    send = state.pack(state.rho_phase)
    _, recv = state.buffers(state.rho_phase)
    with state.torch.cuda.stream(state.stream):
        transport.exchange(send, recv)
    state.unpack(state.rho_phase)
    lap("halo_rho")
    if state.sphenix:
        be.runner_do_gradient()
        be.runner_do_extra_ghost()
        lap("gradient")
        send = state.pack(2)
        _, recv = state.buffers(2)
        with state.torch.cuda.stream(state.stream):
            transport.exchange(send, recv)
        state.unpack(2)
        lap("halo_gradient")
    be.runner_do_force()
    lap("force")


def nccl_parity_check(tr, world, rank, local, per_cell=150, seed=11, clustered=False, sphenix=False):
    """A small box split over the ranks and stepped through the SAME code the timed loop runs
    (step_rank_nccl: halo pack, NCCL send/recv, unpack), TWICE (the second step starts from what the
    first left in the halo copies), against one context holding the whole box on rank 0: every owned
    particle's results and the summed counters must be bit-identical.  clustered: a clumpy box in
    cost-weighted blocks (the shape of BASELINE configs[4]) instead of the uniform one.
    Role of the reference's tests/testHydroMPIrules.c + engine_maketasks.c:141-324 (send/recv xv, rho).
    sphenix: the three-loop scheme with its three exchanges (xv; rho + u + alpha; the alphas after the extra
    ghost), ONE step (the scheme's alphas evolve from step to step), its own fields compared as well.
    Returns "ok ..." or raises on every rank."""
    import torch.distributed as dist
    grid = grid_for(world)
    if clustered:
        dim = (1.0, 1.0, 1.0)
        ic = ics.clustered(400000, nclumps=40, seed=seed)
        ic["id"] = np.arange(1, len(ic["h"]) + 1, dtype=np.int64)
        cdim = ics.top_grid(dim, float(ic["h"].max()))
        weights = np.bincount(top_cell_of(ic["x"], cdim, dim), minlength=int(np.prod(cdim)))
        part = Partition(cdim, grid, periodic=True, weights=weights)
        st = RankState(rank, part, dim, 0, device=local, ic=local_particles_from(ic, part, rank, dim),
                       params_kw=dict(scheme=2, alpha=0.1) if sphenix else None)
    else:
        dim = tuple(float(g) for g in grid)
        cdim = tuple(int(12 * d / max(dim)) for d in dim)
        h0 = ics.ETA / (per_cell * np.prod(cdim) / np.prod(dim)) ** (1.0 / 3.0)
        assert ics.top_grid(dim, h0) == cdim
        part = Partition(cdim, grid, periodic=True)
        st = RankState(rank, part, dim, per_cell, seed=seed, device=local,
                       params_kw=dict(scheme=2, alpha=0.1) if sphenix else None)
        ic = None
    st.set_send_lists(tr.all_to_all_objects({p: ids for p, ids in st.recv_ids.items()}))
    for _ in range(1 if sphenix else 2):
        # poison the receive buffers: a pack / transfer / unpack ordering bug then shows as NaNs
        # instead of hiding behind what the previous step left there
        for phase in ((0, 3, 2) if sphenix else (0, 1)):
            with st.torch.cuda.stream(st.stream):
                for buf in st.buffers(phase)[1].values():
                    buf.fill_(255)
        step_rank_nccl(st, tr)
    mine = st.owned_results()
    cnt = st.be.counters()
    mine["_counts"] = (cnt["n_density"], cnt["n_density_redo"], cnt["n_force"])
    if sphenix:
        mine["_counts"] += (st.be.sphenix_results()["n_gradient"],)
    st.be.close()
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0, group=tr.setup_group)
    verdict = ["ok"]
    npart = 0
    if rank == 0:
        if ic is None:
            ic = ics.cell_block_random(np.arange(int(np.prod(cdim))), cdim, dim, per_cell, seed=seed)
        npart = len(ic["h"])
        be = HydroBackend(make_params(dim=dim, periodic=True, **(dict(scheme=2, alpha=0.1) if sphenix else {})),
                          device=local)
        be.build(ic)
        be.step()
        ref, cref = be.results(), be.counters()
        want = (cref["n_density"], cref["n_density_redo"], cref["n_force"])
        if sphenix:
            sr = be.sphenix_results()
            want += (sr.pop("n_gradient"),)
            ref.update(sr)
        be.close()
        pos = {int(v): k for k, v in enumerate(ic["id"])}
        tot = np.zeros(len(want), dtype=np.int64)
        nchecked = 0
        for r, res in enumerate(gathered):
            idx = np.array([pos[int(v)] for v in res["id"]], dtype=np.int64)
            nchecked += len(idx)
            for k in ref:
                if len(idx) and not np.array_equal(res[k], ref[k][idx]):
                    verdict[0] = "rank %d: field %s differs from the single-GPU run" % (r, k)
            tot += np.array(res["_counts"], dtype=np.int64)
        if nchecked != npart:
            verdict[0] = "owned particles of all ranks: %d, box: %d" % (nchecked, npart)
        if tuple(tot) != want:
            verdict[0] = "interaction counters differ: %s vs %s" % (tuple(tot), want)
    dist.broadcast_object_list(verdict, src=0, group=tr.setup_group)
    if verdict[0] != "ok":
        raise RuntimeError("NCCL path parity check failed: " + verdict[0])
    return "ok (%d ranks, %s box, %s, every owned field and the counters bit-identical to one GPU)" % (
        world, "clustered" if clustered else "uniform",
        "SPHENIX with its three exchanges, one step" if sphenix else "two steps")


def bench_multi(args, cfg, metric, unit, sampler_cls=None):
    """bench.py under torchrun: weak scaling, `side`^3 particles per GPU (BASELINE config 4 at 8 GPUs:
    512^3 over 2 x 2 x 2)."""
    import json
    import os
    import time
    import torch
    import torch.distributed as dist
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tr = NcclTransport()
    # correctness of the timed path first: the same halo code on a small box, against one GPU
    nccl_parity = nccl_parity_check(tr, world, rank, local)
    if getattr(args, "workload", "uniform") == "clustered":
        nccl_parity += "; " + nccl_parity_check(tr, world, rank, local, clustered=True)
    grid = grid_for(world)
    side = args.side
    clustered = getattr(args, "workload", "uniform") == "clustered"
    fast = getattr(args, "fields", "fast") == "fast"
    pkw = dict(fields_mode=1 if fast else 0)
    t0 = time.perf_counter()
    if clustered:
        # BASELINE configs[4]: ONE clustered box of side^3 particles split over the ranks (strong
        # scaling), blocks weighted by particle count; every rank generates the same seeded set
        dim = (1.0, 1.0, 1.0)
        gic = ics.clustered(side ** 3, nclumps=1000, seed=7)
        gic["id"] = np.arange(1, len(gic["h"]) + 1, dtype=np.int64)
        cdim = ics.top_grid(dim, float(gic["h"].max()))
        weights = np.bincount(top_cell_of(gic["x"], cdim, dim), minlength=int(np.prod(cdim)))
        part = Partition(cdim, grid, periodic=True, weights=weights)
        per_cell = 0
        st = RankState(rank, part, dim, per_cell, device=local, ic=local_particles_from(gic, part, rank, dim),
                       params_kw=pkw)
        del gic
    else:
        dim = tuple(float(g) for g in grid)
        h0 = ics.ETA / side
        cdim = ics.top_grid(dim, h0)
        part = Partition(cdim, grid, periodic=True)
        per_cell = int(round(side ** 3 * world / float(np.prod(cdim))))
        st = RankState(rank, part, dim, per_cell, seed=cfg["seed"], device=local, params_kw=pkw)
    wanted = tr.all_to_all_objects({p: ids for p, ids in st.recv_ids.items()})
    st.set_send_lists(wanted)
    build_s = time.perf_counter() - t0
    n_own = int(st.is_own.sum())

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_rank_nccl(st, tr)
    sampler = None
    if rank == 0 and sampler_cls is not None:
        sampler = sampler_cls(local)
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_rank_nccl(st, tr)
    e1.record()
    torch.cuda.synchronize()
    ms_dev = e0.elapsed_time(e1) / args.steps
    ms_wall = 1e3 * (time.perf_counter() - t0) / args.steps
    clocks = sampler.stop() if sampler is not None else None
    # per-phase split (one more step, every phase synchronised: outside the timed region)
    phase_ms = {}
    barrier()
    step_rank_nccl(st, tr, phase_ms)
    pt = torch.tensor([phase_ms[k] for k in PHASES], dtype=torch.float64, device="cuda")
    dist.all_reduce(pt, op=dist.ReduceOp.MAX)
    phase_max = {k: float(v) for k, v in zip(PHASES, pt.tolist())}
    cnt = st.be.counters()
    inter = cnt["n_density"] + cnt["n_density_redo"] + cnt["n_force"]
    # end to end: H2D of this rank's particle state from pinned host memory, the step, D2H of the results
    ic_pin, keep = {}, []
    for k, a in st.ic.items():
        if isinstance(a, np.ndarray) and k not in ("cell", "owner"):
            tt = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
            keep.append(tt)
            ic_pin[k] = tt.numpy()
        elif a is None:
            ic_pin[k] = None
    out_pin = st.be.alloc_results(pinned=True)
    nloc = len(st.ic["h"])
    e2e_s = []
    # a rank's own particles are a prefix of its arrays: only they cross PCIe, the halo copies are
    # refreshed over NVLink inside the step
    assert st.is_own[:n_own].all()
    st.be.set_transfer_count(n_own)
    for k in range(3):
        barrier()
        t1 = time.perf_counter()
        st.be.update_particles(ic_pin)
        t2 = time.perf_counter()
        step_rank_nccl(st, tr)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        st.be.results(out=out_pin)
        torch.cuda.synchronize()
        if k > 0:
            e2e_s.append(time.perf_counter() - t1)
            e2e_split = {"h2d_ms": 1e3 * (t2 - t1), "step_ms": 1e3 * (t3 - t2), "d2h_ms": 1e3 * (time.perf_counter() - t3)}
    e2e_ms = 1e3 * float(np.mean(e2e_s))
    t = torch.tensor([max(ms_dev, ms_wall), float(inter), float(n_own), float(st.halo_bytes_per_step()), e2e_ms,
                      float(nloc)], dtype=torch.float64, device="cuda")
    tmax = t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = t.clone()
    dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    ms = float(tmax[0])
    if rank == 0:
        cfg = dict(cfg)
        if clustered:
            cfg["workload"] = ("clustered periodic box (1000 Plummer clumps + 30 %% uniform background), %d^3 "
                               "particles in TOTAL, %s top-level cells in %dx%dx%d blocks of equal particle count, "
                               "density+ghost+force, Minimal SPH + ideal gas"
                               % (side, "x".join(map(str, cdim)), grid[0], grid[1], grid[2]))
        else:
            cfg["workload"] = ("periodic uniform-random box, %d particles per top-level cell, %s top-level cells "
                               "over a %dx%dx%d GPU grid (%d^3 particles per GPU), density+ghost+force, Minimal SPH "
                               "+ ideal gas" % (per_cell, "x".join(map(str, cdim)), grid[0], grid[1], grid[2], side))
        cfg["parallelism"] = "spatial %dx%dx%d, NCCL send/recv halo (xv 48 B, rho 24 B per particle)" % grid
        cfg["fields_mode"] = ("FAST: neighbour lists, counts and h bit-exact; force-loop fields with FMA / MUFU "
                              "arithmetic, within 1e-4 of the reference" if fast else
                              "EXACT: every field bit-identical to the reference")
        nloc_tot = int(tsum[5])
        line = {"metric": metric, "value": float(tsum[1]) / (ms * 1e-3), "unit": unit, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong" if clustered else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
                "clocks": clocks,
                "gpu_launches": (2 * st.be.params.max_smoothing_iterations + 5 + 4 * len(st.send_idx))
                * args.steps * world,
                "e2e": {"value": float(tsum[1]) / (float(tmax[4]) * 1e-3), "unit": unit,
                        "h2d_bytes_per_step": int(tsum[2]) * 49, "d2h_bytes_per_step": int(tsum[2]) * 49},
                "detail": {"n_particles_owned_total": int(tsum[2]), "interactions_per_step": int(tsum[1]),
                           "halo_bytes_per_step_total": int(tsum[3]), "build_s": build_s,
                           "load_imbalance_interactions_max_over_mean": float(tmax[1]) * world / float(tsum[1]),
                           "ms_device_events_rank0": ms_dev, "ms_wall_rank0": ms_wall,
                           "nccl_parity": nccl_parity, "phase_ms": phase_max, "e2e_ms_max_over_ranks": float(tmax[4]), "e2e_split_rank0": e2e_split}}
        print(json.dumps(line))
    dist.barrier()
    dist.destroy_process_group()
    return 0
