"""Directions::update (sources/map/paths.cpp:24-78) of ONE large grid split into row stripes, one rank (GPU) per stripe.

The protocol (include/mazetd_b200.h, "row stripes"): every rank relaxes its stripe to local convergence, then the ranks
swap their boundary rows with the stripes above / below (torch.distributed point-to-point: NCCL over NVLink on the GPU
box, gloo in the CPU tests), merge what arrived into their halo copies and relax again; one all-reduce of the number of
improved halo cells per sweep ends the loop. The distance fixpoint is unique and next hops are a pure function of final
distances, so the result is bit-identical to a single Directions::update over the whole grid whatever the cut.

A "stripe solver" is anything with begin / boundary_rows / exchange / finalize; the product's is GpuStripe (libmazetd_b200
through the C ABI). Tests inject a CPU solver to exercise the protocol itself on gloo without a GPU.
"""
from . import shard


class GpuStripe:
    """rows [y0, y1) of a W x H grid on one CUDA device, through the C ABI. torch supplies the buffers and the stream."""

    def __init__(self, device, W, H, y0, y1, elev_rows, flags_rows, sources, stream=None, round_budget=0):
        import torch

        from . import capi

        self.torch = torch
        self.W, self.H, self.y0, self.y1 = W, H, y0, y1
        self.dev = torch.device("cuda", device)
        self.ctx = capi.Context(device, stream=stream)
        if isinstance(elev_rows, torch.Tensor):
            assert elev_rows.is_cuda and flags_rows.is_cuda and elev_rows.is_contiguous() and flags_rows.is_contiguous()
            self._keep = (elev_rows, flags_rows)
            self.ctx.stripe_upload_dev(W, H, y0, y1, elev_rows.data_ptr(), flags_rows.data_ptr())
        else:
            self.ctx.stripe_upload(W, H, y0, y1, elev_rows, flags_rows)
        self.sources = sources
        self.ctx.stripe_round_budget(round_budget)  # > 0: swap rows every that many relaxation rounds instead of at local convergence
        self.dist = torch.empty((y1 - y0, W), dtype=torch.int32, device=self.dev)  # uint32 bit patterns
        self.tiles = torch.empty((y1 - y0, W), dtype=torch.int32, device=self.dev)

    def begin(self):
        self.ctx.stripe_begin_dev(self.sources, self.dist.data_ptr())

    def boundary_rows(self):
        """(first owned row, last owned row) as they stand now; contiguous views, no copy"""
        return self.dist[0], self.dist[-1]

    def new_row(self):
        return self.torch.empty(self.W, dtype=self.torch.int32, device=self.dev)

    def exchange(self, row_above, row_below):
        """merge the neighbours' rows and relax again. Returns how much this call had to do: halo cells improved + tiles
        that were still queued from the previous (budgeted) call; 0 on every rank = nothing moved anywhere = converged."""
        self.torch.cuda.current_stream(self.dev).synchronize()  # the received rows were written on torch's stream
        queued = self.ctx.stripe_pending()
        return queued + self.ctx.stripe_exchange_dev(row_above.data_ptr() if row_above is not None else None, row_below.data_ptr() if row_below is not None else None)

    def finalize(self):
        self.ctx.stripe_finalize_dev(self.tiles.data_ptr())
        self.ctx.synchronize()
        return self.tiles, self.dist

    # ---- peer-coupled mode: one kernel run per rank, rows and keys go straight into the neighbours' memory ----
    def peer_launch(self):
        """queue the whole update (every rank must); returns at once"""
        self.ctx.stripe_peer_launch_dev(self.sources, self.dist.data_ptr(), self.tiles.data_ptr())

    def peer_finish(self):
        """wait for this rank; returns (tiles, dist, rounds)"""
        rounds = self.ctx.stripe_peer_finish()
        return self.tiles, self.dist, rounds

    def close(self):
        self.ctx.close()


def solve_distributed(solver, rank, world, group=None):
    """Run the stripe protocol on this rank (every rank of `group` calls it). Returns (tiles, dist, sweeps)."""
    import torch
    import torch.distributed as dist

    solver.begin()
    sweeps = 0
    while world > 1:
        first, last = solver.boundary_rows()
        above = solver.new_row() if rank > 0 else None
        below = solver.new_row() if rank < world - 1 else None
        ops = []
        if rank > 0:
            ops += [dist.P2POp(dist.isend, first, rank - 1, group), dist.P2POp(dist.irecv, above, rank - 1, group)]
        if rank < world - 1:
            ops += [dist.P2POp(dist.isend, last, rank + 1, group), dist.P2POp(dist.irecv, below, rank + 1, group)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        improved = solver.exchange(above, below)
        t = torch.tensor([improved], dtype=torch.int64, device=first.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        sweeps += 1
        if int(t.item()) == 0:
            break
    tiles, d = solver.finalize()
    return tiles, d, sweeps


def connect_peers(stripe, rank, world, group=None):
    """once per stripe: every rank publishes the CUDA IPC handle of its peer block and maps its neighbours' and rank 0's"""
    import torch.distributed as dist

    mine = stripe.ctx.stripe_peer_export()
    handles = [None] * world
    dist.all_gather_object(handles, mine, group=group)
    stripe.ctx.stripe_peer_connect(rank, world, handles)
    dist.barrier(group=group)  # nobody launches before every block is mapped everywhere


def solve_peer(stripe):
    """Directions::update over all ranks in one kernel run per rank (connect_peers first). Returns (tiles, dist, rounds).
    No host-side collective: the ranks meet at the round barriers on the device."""
    stripe.peer_launch()
    return stripe.peer_finish()


def solve_peer_local(stripes, ctas=64):
    """the peer-coupled protocol with every stripe a context of this process on ONE GPU (tests): each stripe's kernel is
    limited to `ctas` CTAs so that all of them are resident together. Every kernel involved must have been launched once
    before (CUDA loads kernels lazily, and a load while the first stripe's kernel waits holds the others back).
    Returns ([(tiles, dist) per stripe], rounds)."""
    n = len(stripes)
    for s in stripes:
        s.ctx.stripe_peer_export()
    for i, s in enumerate(stripes):
        s.ctx.stripe_peer_connect_local(i, n, [t.ctx for t in stripes], ctas)
    for s in stripes:
        s.peer_launch()
    out, err = [], None
    for s in stripes:  # wait for every stripe even if one reports a barrier timeout
        try:
            out.append(s.peer_finish())
        except Exception as e:  # noqa: BLE001
            err = err or e
    if err is not None:
        raise err
    return [(t, d) for t, d, _ in out], out[0][2]


def solve_local(solvers):
    """The same protocol with every stripe in this process (stripes top to bottom; e.g. several contexts on one GPU).
    Returns ([(tiles, dist) per stripe], sweeps)."""
    for s in solvers:
        s.begin()
    n = len(solvers)
    sweeps = 0
    while n > 1:
        rows = [tuple(r.clone() for r in s.boundary_rows()) for s in solvers]  # a sweep swaps the rows as they stood at its start
        improved = 0
        for i, s in enumerate(solvers):
            improved += s.exchange(rows[i - 1][1] if i > 0 else None, rows[i + 1][0] if i < n - 1 else None)
        sweeps += 1
        if improved == 0:
            break
    return [s.finalize() for s in solvers], sweeps


def make_gpu_stripe(device, rank, world, W, H, elev, flags, sources, stream=None, round_budget=0):
    """cut rank's stripe (with its halo rows) out of full host arrays elev / flags ([H][W]) and upload it"""
    y0, y1 = shard.stripe_rows(rank, world, H)
    ya, yb = shard.stripe_halo_rows(y0, y1, H)
    e = elev.reshape(H, W)[ya:yb]
    f = flags.reshape(H, W)[ya:yb]
    return GpuStripe(device, W, H, y0, y1, e, f, sources, stream=stream, round_budget=round_budget)
