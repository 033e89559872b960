"""y-slab multi-GPU driver: one process per GPU, ``torch.distributed`` (NCCL over NVLink).

Decomposition = the reference's MPI layout with npx = 1, npy = N (src/grid.cxx:142-150):
  * stencil phases: north/south halo rows exchanged with the two ring neighbours
    (Boundary_cyclic MPI path, src/boundary_cyclic.cxx:139-174); east/west wrap stays local;
  * Poisson solve: x FFT local -> all-to-all (Transpose::exec_xy) -> y FFT + tridiagonal solve local
    -> inverse y FFT -> all-to-all (exec_yx) -> inverse x FFT  (src/fft.cxx:452-583 with npx = 1);
  * CFL / dnmul / divergence maxima and means: all-reduce (Master::max/sum, src/master_parallel.cxx:233-256).

Two data paths:
  * peer (default on GPUs): every rank's exchange arena is mapped into the other ranks (CUDA IPC over NVLink / NVSwitch,
    ``host.PeerArena``); halos are ONE kernel per exchange that stores my edge rows into the neighbours' ghost rows
    (``mhh_peer_halo_push``), and the two all-to-alls of the solver happen INSIDE the forward x / inverse y transforms, which
    write their chunk-layout output straight into the peers' receive buffers (``mhh_fft_x/y_chunks_to_peers``), followed by a
    device-side barrier (``mhh_peer_barrier``).  No pack / unpack passes, no host waits; NCCL only carries the reductions.
  * nccl (MHH_SLAB_PATH=nccl, and the gloo CPU tests): ``SlabComm`` send/recv + all_to_all_single with the pack/unpack kernels
    ``mhh_halo_ns`` / chunk-layout transforms around them.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import lib as L
from .host import Boundary_cyclic, Pres
from .model import Model


def nccl_env_defaults():
    """Call before dist.init_process_group: the all-to-all of the pressure solver is point-to-point traffic, and NCCL's
    default P2P channel count leaves NVLink half idle (measured at 2 x B200, 0.54 GB sent per GPU: 418 GB/s with the
    defaults, 577 GB/s with 64 channels; tools/bench_a2a.py)."""
    import os
    os.environ.setdefault("NCCL_MIN_P2P_NCHANNELS", "64")
    os.environ.setdefault("NCCL_MAX_P2P_NCHANNELS", "64")


class SlabComm:
    """Ring neighbours, halo send/recv, all-to-all and reductions for ``world`` y-slabs."""

    def __init__(self, rank, world, group=None):
        self.rank, self.world, self.group = rank, world, group
        self.north = (rank + 1) % world      # holds rows j >= my jend   (MPI nnorth)
        self.south = (rank - 1) % world      # holds rows j <  my jstart (MPI nsouth)

    def halo_exchange(self, to_north, to_south, from_south, from_north):
        """Send my top interior rows north and bottom interior rows south; receive the matching ghosts."""
        if self.world == 1:
            from_south.copy_(to_north)
            from_north.copy_(to_south)
            return
        ops = [dist.P2POp(dist.isend, to_north, self.north, self.group),
               dist.P2POp(dist.irecv, from_south, self.south, self.group),
               dist.P2POp(dist.isend, to_south, self.south, self.group),
               dist.P2POp(dist.irecv, from_north, self.north, self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()

    def all_to_all(self, out, inp):
        """Equal chunks, chunk r of ``inp`` goes to rank r (Transpose::exec_xy / exec_yx)."""
        if self.world == 1:
            out.copy_(inp)
        else:
            dist.all_to_all_single(out, inp, group=self.group)

    def all_to_all_async(self, out, inp):
        """all_to_all on NCCL's own stream; returns a handle whose wait() makes the current stream wait for it (or None)."""
        if self.world == 1:
            out.copy_(inp)
            return None
        return dist.all_to_all_single(out, inp, group=self.group, async_op=True)

    def allreduce_max(self, value, device):
        t = torch.tensor([value], dtype=torch.float64, device=device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def allreduce_max_many(self, values, device):
        """Several maxima in one collective.  NaN must win (the reference's CFL guard): NCCL's MAX is not specified for NaN,
        so a NaN flag rides along as one more value."""
        vals = [float(v) for v in values]
        nan = any(v != v for v in vals)
        t = torch.tensor([0. if v != v else v for v in vals] + [1. if nan else 0.], dtype=torch.float64, device=device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        out = t.tolist()
        return [float("nan")] * len(vals) if out[-1] > 0. else out[:-1]

    def allreduce_sum(self, value, device):
        t = torch.as_tensor(value, dtype=torch.float64, device=device).reshape(-1).clone()
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t


class SlabBoundaryCyclic(Boundary_cyclic):
    """Boundary_cyclic on a y-slab rank: east/west local, north/south through the ring neighbours."""

    def __init__(self, grid, comm):
        super().__init__(grid)
        self.comm = comm
        self.peer = False            # SlabModel switches it on once the arenas are connected
        n3 = grid.kcells * grid.jgc * grid.icells
        mk = lambda n: torch.empty(n, dtype=grid.dtype, device=grid.device)
        if grid.arena is None:
            self.to_n, self.to_s, self.from_s, self.from_n = mk(n3), mk(n3), mk(n3), mk(n3)
        else:
            self.to_n = self.to_s = self.from_s = self.from_n = mk(0)

    def _push(self, flds, is2d):
        """North/south rows of up to 8 fields per launch straight into the neighbours' ghost rows (peer path)."""
        g = self.grid
        for q in range(0, len(flds), 8):
            part = flds[q:q + 8]
            L.check(g.ctx, g.lib.mhh_peer_halo_push(g.ctx, len(part), L.ptr_array(part), int(is2d)), "mhh_peer_halo_push")

    def _ns(self, fld, is2d):
        g = self.grid
        if self.peer:
            self._push([fld], is2d)
            return
        n = (1 if is2d else g.kcells) * g.jgc * g.icells
        if self.to_n.numel() < n:
            mk = lambda: torch.empty(g.kcells * g.jgc * g.icells, dtype=g.dtype, device=g.device)
            self.to_n, self.to_s, self.from_s, self.from_n = mk(), mk(), mk(), mk()
        L.check(g.ctx, g.lib.mhh_halo_ns(g.ctx, L.ptr(fld), L.ptr(self.to_n), L.ptr(self.to_s), 0, int(is2d)), "mhh_halo_ns pack")
        self.comm.halo_exchange(self.to_n[:n], self.to_s[:n], self.from_s[:n], self.from_n[:n])
        L.check(g.ctx, g.lib.mhh_halo_ns(g.ctx, L.ptr(fld), L.ptr(self.from_n), L.ptr(self.from_s), 1, int(is2d)), "mhh_halo_ns unpack")

    def exec_many(self, flds):
        """Boundary_cyclic::exec of several 3-D fields with ONE ring exchange: the north / south rows of all fields travel in
        one message per neighbour (the prognostic fields at the start of a substep: 4 exchanges -> 1)."""
        g = self.grid
        flds = list(flds)
        for fld in flds:
            L.check(g.ctx, g.lib.mhh_cyclic(g.ctx, L.ptr(fld), L.EDGE_EAST_WEST), "mhh_cyclic")
        if g.jtot <= 1 or not flds:
            return
        if self.peer:
            self._push(flds, False)
            return
        n = g.kcells * g.jgc * g.icells
        if self.to_n.numel() < n * len(flds):
            mk = lambda: torch.empty(n * len(flds), dtype=g.dtype, device=g.device)
            self.to_n, self.to_s, self.from_s, self.from_n = mk(), mk(), mk(), mk()
        seg = lambda buf, q: L.ptr(buf[q * n:(q + 1) * n])
        for q, fld in enumerate(flds):
            L.check(g.ctx, g.lib.mhh_halo_ns(g.ctx, L.ptr(fld), seg(self.to_n, q), seg(self.to_s, q), 0, 0), "mhh_halo_ns pack")
        m = n * len(flds)
        self.comm.halo_exchange(self.to_n[:m], self.to_s[:m], self.from_s[:m], self.from_n[:m])
        for q, fld in enumerate(flds):
            L.check(g.ctx, g.lib.mhh_halo_ns(g.ctx, L.ptr(fld), seg(self.from_n, q), seg(self.from_s, q), 1, 0), "mhh_halo_ns unpack")

    def exec(self, fld, edge=L.EDGE_BOTH):
        g = self.grid
        if edge in (L.EDGE_EAST_WEST, L.EDGE_BOTH):
            L.check(g.ctx, g.lib.mhh_cyclic(g.ctx, L.ptr(fld), L.EDGE_EAST_WEST), "mhh_cyclic")
        if edge in (L.EDGE_NORTH_SOUTH, L.EDGE_BOTH) and g.jtot > 1:
            self._ns(fld, False)

    def exec_2d(self, fld2d):
        g = self.grid
        L.check(g.ctx, g.lib.mhh_cyclic_2d(g.ctx, L.ptr(fld2d)), "mhh_cyclic_2d")
        if g.jtot > 1:
            self._ns(fld2d, True)

    def exec_2d_many(self, planes):
        """Boundary_cyclic::exec_2d of several planes with one ring exchange on the peer path (the two momentum flux planes
        of Boundary_surface::exec, src/boundary_surface.cxx:894-895)."""
        g = self.grid
        planes = list(planes)
        if not (self.peer and g.jtot > 1):
            for pl in planes:
                self.exec_2d(pl)
            return
        for pl in planes:
            L.check(g.ctx, g.lib.mhh_cyclic_2d(g.ctx, L.ptr(pl)), "mhh_cyclic_2d")
        self._push(planes, True)


def sub_block_count(iblock, world, wanted=4):
    """kx sub-blocks per rank: the largest power-of-two fraction of `wanted` that divides the rank's block, keeps at least 8
    columns per sub-block and fits ranks x sub-blocks into the transforms' 32-entry segment table."""
    B = max(1, wanted)
    while B > 1 and (iblock % B or world * B > 32 or iblock // B < 8):
        B //= 2
    return B


def sub_block_plan(world, me, nsub):
    """Where the segments of the kx sub-block pipeline live (pure index arithmetic; segment = [k][j < jmax][ii < ib], slots
    counted in segments).  Segment v = r * nsub + b of my x lines is rank r's sub-block b.
      x_fwd[v]    : (buffer, slot) the forward x pass writes segment v to — the staging array S, or for my own block my
                    receive buffer R1, where slot (source rank) * nsub + b is what the y pass of sub-block b reads;
      y_bwd[b][r] : (buffer, slot) the inverse y pass of sub-block b writes the rows of rank r to — S, or my R2;
      copies[b]   : (peer r, destination slot in r's R1 / R2, source slot in my S) of the transfers of sub-block b, starting
                    with the next rank round the ring (the same list serves the forward and the return transfer)."""
    B = nsub
    x_fwd = [("R1" if r == me else "S", r * B + b) for r in range(world) for b in range(B)]
    y_bwd = [[("R2", me * B + b) if r == me else ("S", r * B + b) for r in range(world)] for b in range(B)]
    copies = [[((me + d) % world, me * B + b, ((me + d) % world) * B + b) for d in range(1, world)] for b in range(B)]
    return dict(x_fwd=x_fwd, y_bwd=y_bwd, copies=copies)


class SlabPres(Pres):
    """Pres_2 / Pres_4 on y-slabs (src/pres_2.cxx:267-361, src/pres_4.cxx:320-529 with src/fft.cxx:452-583, npx = 1): the stage
    calls of the C ABI with the two all-to-alls in between; the vertical solve (Thomas / 7-band LU) is column-local."""

    def __init__(self, grid, fields, swpres, comm, cyclic):
        super().__init__(grid, fields, swpres)
        if swpres not in ("2", "4"):
            raise L.MhhError("the y-slab pressure pipeline needs swpres=2 or 4")
        self.comm, self.cyclic = comm, cyclic
        n = grid.imax * grid.jmax * grid.kmax
        self.peer = False          # SlabModel switches it on once the arenas are connected
        if grid.arena is not None:
            # X: this rank's work array; R1 / R2: receive buffers the peers' forward x / inverse y transforms write into
            self.A, self.R1, self.R2 = (grid.arena.alloc((n,), grid.dtype) for _ in range(3))
            if self.R2 is None:
                raise L.MhhError("peer arena too small for the pressure solver's buffers")
            self.B = self.R1
        else:
            self.A = torch.empty(n, dtype=grid.dtype, device=grid.device)
            self.B = torch.empty(n, dtype=grid.dtype, device=grid.device)
        self.chunk_fft = None      # decided at the first exec (needs the solver's plans): mhh_fft_chunks_supported
        self.kranges = 4           # k ranges whose all-to-alls overlap the transforms of their neighbours (nccl path)
        # how the peer path moves the all-to-all chunks (MHH_SLAB_A2A).  Measured at 512^3 per GPU (r02y, r02v):
        #   2 x B200: ce_overlap 44.6 ms/step, fused 45.0, fused_pipe 44.6, ce 45.3;   8 x B200: ce_overlap 49.6, nccl 50.1, fused 50.4.
        # ce_overlap: own chunk written in place, one copy-engine transfer per peer on parallel streams while the SMs run the
        # scalar tendency pass (which does not depend on the pressure); fused: the transforms store straight into the peers'
        # buffers (no staging, but the NVLink-bound stores hold the SMs: fft_fwd_x 1.01 instead of 0.51 ms at N = 2).
        self.a2a = os.environ.get("MHH_SLAB_A2A", "ce_sub" if comm.world <= 4 else "ce_overlap")
        self.timing = [] if os.environ.get("MHH_SLAB_TIMING") else None     # ce_overlap: CUDA events of every exchange

    def timing_report(self):
        """ms, averaged over the recorded exchanges.  ce_overlap: copies, barrier wait after the copies, hook kernel, main stream
        stalled; ce_sub: the main stream's waits for a sub-block (arrival or return), per wait and per solve."""
        torch.cuda.synchronize(self.grid.device)
        if self.a2a == "ce_sub":
            n = max(len(self.timing), 1)
            names = list(self.timing[0]) if self.timing else []
            out = dict(mode="ce_sub", nsub=self.nsub, streams=len(self._comp), solves=len(self.timing),
                       timeline_ms={k: round(sum(t["start"].elapsed_time(t[k]) for t in self.timing) / n, 3) for k in names if k != "start"},
                       stalled_ms_per_solve=round(sum(t["hooks"].elapsed_time(t["resumed"]) for t in self.timing) / n, 3))
            self.timing.clear()
            return out
        rows = [(a.elapsed_time(b), b.elapsed_time(c), a.elapsed_time(d), d.elapsed_time(e)) for a, b, c, d, e in self.timing]
        self.timing.clear()
        n = max(len(rows), 1)
        return dict(exchanges=len(rows), copies_ms=sum(r[0] for r in rows) / n, barrier_ms=sum(r[1] for r in rows) / n,
                    hook_ms=sum(r[2] for r in rows) / n, main_stalled_ms=sum(r[3] for r in rows) / n)

    def _exec_ce_overlap(self, p, hooks):
        """MHH_SLAB_A2A=ce_overlap: the sending transforms write their chunks locally, the copy engines move them into the
        peers' receive buffers (one cudaMemcpyAsync per peer on a side stream, SMs not involved), and in the meantime the
        main stream runs `hooks[0]` / `hooks[1]` — the two k halves of the scalar tendency + RK pass, which do not depend on
        the pressure — before it waits for the device barrier and runs the receiving transforms."""
        g, lib = self.grid, self.grid.lib
        ck = lambda err, what: L.check(g.ctx, err, what)
        X, R1, R2, S = L.ptr(self.A), L.ptr(self.R1), L.ptr(self.R2), L.ptr(self.S)
        world, me = self.comm.world, self.comm.rank
        R = g.kmax * g.jmax * (g.itot // world)
        main, xfer = torch.cuda.current_stream(g.device), self._side[0]
        on = lambda st: lib.mhh_set_stream(g.ctx, C.c_void_p(st.cuda_stream))

        T = self.timing is not None
        mk = lambda: torch.cuda.Event(enable_timing=T)

        def exchange(peers_t, hook):
            ev = mk(); ev.record(main)
            for n, d in enumerate(range(1, world)):     # my own chunk is already in place (_ce_tables)
                st = self._copy_streams[n % len(self._copy_streams)]
                if n < len(self._copy_streams):
                    st.wait_event(ev)
                r = (me + d) % world
                with torch.cuda.stream(st):
                    peers_t[r][me * R:(me + 1) * R].copy_(self.S[r * R:(r + 1) * R], non_blocking=True)
            for st in self._copy_streams[:max(world - 1, 0)]:
                if st is not xfer:
                    e = torch.cuda.Event(); e.record(st)
                    xfer.wait_event(e)
            if world == 1:
                xfer.wait_event(ev)
            ev_c = mk(); ev_c.record(xfer)
            on(xfer)
            ck(lib.mhh_peer_barrier(g.ctx), "peer barrier")
            ev2 = mk(); ev2.record(xfer)
            on(main)
            if hook is not None:
                hook()
            ev_h = mk(); ev_h.record(main)
            main.wait_event(ev2)
            ev_w = mk(); ev_w.record(main)
            if T:       # (send transform done, copies done, barrier done, hook done, main resumes)
                self.timing.append((ev, ev_c, ev2, ev_h, ev_w))

        t1, t2 = self._ce_tabs
        ck(lib.mhh_fft_x_chunks_to_peers(g.ctx, X, t1, 0, g.kmax), "fft_x fwd -> chunks")
        exchange(self._r1_t, hooks[0] if hooks else None)
        ck(lib.mhh_fft_y_chunks(g.ctx, -1, R1, X, 0, g.kmax), "chunks -> fft_y fwd")
        ck(lib.mhh_pres_solve_spectral_ws(g.ctx, X, R1), "tdma")
        ck(lib.mhh_fft_y_chunks_to_peers(g.ctx, X, t2, 0, g.kmax), "fft_y bwd -> chunks")
        exchange(self._r2_t, hooks[1] if hooks else None)
        ck(lib.mhh_fft_x_chunks(g.ctx, +1, R2, None, L.ptr(p), 0, g.kmax), "chunks -> fft_x bwd")

    def _sub_setup(self):
        """MHH_SLAB_A2A=ce_sub: tables of the kx sub-block pipeline (mhh_slab_*_sub, include/mhh_b200.h)."""
        g = self.grid
        world, me = self.comm.world, self.comm.rank
        self.nsub = B = sub_block_count(g.itot // world, world, int(os.environ.get("MHH_SLAB_NSUB", "4")))
        ib = g.itot // world // B
        self._Rb = Rb = g.kmax * g.jmax * ib
        es = self.S.element_size()
        plan = sub_block_plan(world, me, B)
        buf = dict(S=self.S, R1=self.R1, R2=self.R2)
        at = lambda where: buf[where[0]].data_ptr() + where[1] * Rb * es
        self._sub_xtab = (C.c_void_p * (world * B))(*[at(w) for w in plan["x_fwd"]])
        self._sub_ytab = [(C.c_void_p * world)(*[at(w) for w in plan["y_bwd"][b]]) for b in range(B)]
        self._sub_copies = plan["copies"]
        ncomp = max(1, int(os.environ.get("MHH_SLAB_SUB_STREAMS", str(B))))
        self._comp = [torch.cuda.Stream(device=g.device, priority=-1) for _ in range(ncomp)]
        self._W = [torch.empty((g.kmax // 16 + 2) * g.jtot * ib, dtype=g.dtype, device=g.device) for _ in range(ncomp)]   # pivot checkpoints
        self._bar = torch.cuda.Stream(device=g.device, priority=-1)
        self._hooks_first = os.environ.get("MHH_SLAB_SUB_HOOKS", "first" if world >= 8 else "overlap") == "first"

    def _exec_ce_sub(self, p, hooks):
        """MHH_SLAB_A2A=ce_sub: the solve pipelined over kx sub-blocks.  The forward x pass writes its segments into a staging
        array; the copy engines send sub-block b of every peer's block (one cudaMemcpyAsync per peer), and as soon as sub-block
        b has arrived from everyone (device barrier on its own stream) a high-priority stream of its own runs y transform ->
        Thomas sweep -> inverse y transform on it, while sub-block b+1 is still in flight; its rows then travel back while
        b+1 is computed.  The vertical solve is no longer a join between the two all-to-alls.  The scalar tendency pass
        (`hooks`) runs on the main stream beside all of this (or first, see `_hooks_first`); the main stream waits only for
        the last sub-block's return before the inverse x pass.  Measured: 43.5 against 45.4 ms/step (2 GPUs), 46.9 against
        48.0 (4), 52.0 against 50.8 (8) next to ce_overlap on the same boxes (profiles/r02_n8_phases.txt)."""
        g, lib = self.grid, self.grid.lib
        ck = lambda err, what: L.check(g.ctx, err, what)
        X, R1, R2 = L.ptr(self.A), L.ptr(self.R1), L.ptr(self.R2)
        world, me, B, Rb = self.comm.world, self.comm.rank, self.nsub, self._Rb
        main, bar = torch.cuda.current_stream(g.device), self._bar
        cs = self._copy_streams[:max(1, min(len(self._copy_streams), world - 1))]
        on = lambda st: lib.mhh_set_stream(g.ctx, C.c_void_p(st.cuda_stream))
        T = self.timing is not None

        def transfer(peers_t, b, after):
            """Sub-block b to every peer's buffer, then the barrier: returns the event 'sub-block b is complete everywhere'."""
            for st in cs:
                st.wait_event(after)
            for n, (r, dst, src) in enumerate(self._sub_copies[b]):
                with torch.cuda.stream(cs[n % len(cs)]):
                    peers_t[r][dst * Rb:(dst + 1) * Rb].copy_(self.S[src * Rb:(src + 1) * Rb], non_blocking=True)
            for st in cs:
                e = torch.cuda.Event(); e.record(st)
                bar.wait_event(e)
            on(bar)
            ck(lib.mhh_peer_barrier(g.ctx), "peer barrier")
            on(main)
            done = torch.cuda.Event(); done.record(bar)
            return done

        # MHH_SLAB_TIMING: the solve's timeline, events on the streams they belong to (offsets from the start of the x pass)
        tl = {} if T else None

        def mark(name, st):
            if T:
                e = torch.cuda.Event(enable_timing=True); e.record(st)
                tl[name] = e

        mark("start", main)
        # The scalar tendency pass saturates HBM, and copy-engine traffic next to it all but stops (r02n8b timeline on 8 GPUs:
        # the first quarter of the forward transfer arrived 0.57 ms after the pass ended, 2.3 ms after it was issued).  With 7
        # peers the transfers are the critical path, so there the pass runs first and the transfers overlap only the
        # sub-blocks' transforms and sweeps; with 1-3 peers the pass runs beside them (nothing stalls: r02z3).
        first = self._hooks_first
        if first:
            for h in hooks or ():
                h()
        ck(lib.mhh_slab_fft_x_fwd_sub(g.ctx, X, self._sub_xtab, B), "fft_x fwd -> sub-block segments")
        ev = torch.cuda.Event(); ev.record(main)
        mark("x_fwd", main)
        arrived = []
        for b in range(B):
            arrived.append(transfer(self._r1_t, b, ev))
            mark(f"arrived{b}", bar)
        # Every sub-block has its own high-priority stream: a sub-block's Thomas sweep has too few columns to fill the GPU
        # (k-sequential: it runs at the latency of its chain, tools/bench_tdma_sub.py), so the sweeps of the sub-blocks share
        # the SMs with each other, with the transforms and with the scalar tendency pass (`hooks`, HBM bound) on the main stream.
        back = []
        for b in range(B):
            st = self._comp[b % len(self._comp)]
            st.wait_event(arrived[b])
            on(st)
            ck(lib.mhh_slab_fft_y_fwd_sub(g.ctx, R1, X, b, B), "segments -> fft_y fwd")
            mark(f"y_fwd{b}", st)
            ck(lib.mhh_slab_solve_sub(g.ctx, X, L.ptr(self._W[b % len(self._comp)]), b, B), "tdma")
            mark(f"tdma{b}", st)
            ck(lib.mhh_slab_fft_y_bwd_sub(g.ctx, X, self._sub_ytab[b], b, B), "fft_y bwd -> segments")
            ev = torch.cuda.Event(); ev.record(st)
            mark(f"y_bwd{b}", st)
            on(main)
            back.append(transfer(self._r2_t, b, ev))
            mark(f"back{b}", bar)
        if not first:
            for h in hooks or ():
                h()
        mark("hooks", main)
        main.wait_event(back[-1])          # the barrier stream is ordered: the last return implies the earlier ones
        mark("resumed", main)
        ck(lib.mhh_slab_fft_x_bwd_sub(g.ctx, R2, L.ptr(p), B), "segments -> fft_x bwd")
        mark("x_bwd", main)
        if T:
            self.timing.append(tl)

    def _ce_tables(self):
        """Destination tables for the sending transforms of ce_overlap (mhh_fft_*_chunks_to_peers writes segment r at
        table[r] + rank * R): the segments for the other ranks land at S[r * R] in the local staging array, from where the
        copy engines take them; my own segment goes straight into my receive buffer, so no local device-to-device copy
        competes with the kernels for HBM."""
        g = self.grid
        world, me = self.comm.world, self.comm.rank
        R = g.kmax * g.jmax * (g.itot // world)
        es = self.S.element_size()

        def table(recv):
            ptrs = [recv.data_ptr() if r == me else self.S.data_ptr() + (r - me) * R * es for r in range(world)]
            return (C.c_void_p * 8)(*(ptrs + [0] * (8 - world)))
        return table(self.R1), table(self.R2)

    def _exec_peer_transforms(self, p, hooks=None):
        """Forward transforms, tridiagonal solve, inverse transforms with the all-to-alls fused into the transforms' stores
        (MHH_SLAB_A2A=fused, one k range), the same cut into k ranges whose NVLink-bound sending transform overlaps the
        receiving transform of the previous range on a second stream (fused_pipe, default), or the chunks written locally and
        moved by the copy engines, one cudaMemcpyAsync per peer and k range, while the SMs transform the next range (ce)."""
        g, lib = self.grid, self.grid.lib
        ck = lambda err, what: L.check(g.ctx, err, what)
        X, R1, R2 = L.ptr(self.A), L.ptr(self.R1), L.ptr(self.R2)
        if self._r1_peers is None:
            self._r1_peers, self._r2_peers = g.arena.peer_ptrs(self.R1), g.arena.peer_ptrs(self.R2)
            if self.a2a not in ("fused", "fused_pipe", "ce", "ce_overlap", "ce_sub"):
                raise L.MhhError("MHH_SLAB_A2A must be fused, fused_pipe, ce, ce_overlap or ce_sub")
            if self.a2a == "ce_sub" and self.swpres != "2":
                self.a2a = "ce_overlap"          # the sub-block sweep exists for the tridiagonal solve only
            if self.a2a != "fused":
                self._side = [torch.cuda.Stream(device=g.device), torch.cuda.Stream(device=g.device)]
            if self.a2a in ("ce", "ce_overlap", "ce_sub"):
                from .host import _RawCuda
                nb = self.R1.numel() * self.R1.element_size()
                wrap = lambda ptr: torch.as_tensor(_RawCuda(ptr, nb), device=g.device).view(g.dtype)
                self._r1_t = [wrap(q) for q in self._r1_peers]
                self._r2_t = [wrap(q) for q in self._r2_peers]
                self.S = torch.empty_like(self.A)
            if self.a2a in ("ce_overlap", "ce_sub"):
                # one cudaMemcpyAsync per peer, spread over a few streams so that several copy engines / NVLink ports are busy
                ncs = max(1, int(os.environ.get("MHH_SLAB_COPY_STREAMS", "4")))
                self._copy_streams = [self._side[0]] + [torch.cuda.Stream(device=g.device) for _ in range(ncs - 1)]
                self._ce_tabs = self._ce_tables()
                if self.a2a == "ce_sub":
                    self._sub_setup()
        if self.a2a == "ce_overlap":
            return self._exec_ce_overlap(p, hooks)
        if self.a2a == "ce_sub":
            return self._exec_ce_sub(p, hooks)
        for h in hooks or ():
            h()
        if self.a2a == "fused":
            ck(lib.mhh_fft_x_chunks_to_peers(g.ctx, X, self._r1_peers, 0, g.kmax), "fft_x fwd -> peers")
            ck(lib.mhh_peer_barrier(g.ctx), "peer barrier")
            ck(lib.mhh_fft_y_chunks(g.ctx, -1, R1, X, 0, g.kmax), "chunks -> fft_y fwd")
            ck(lib.mhh_pres_solve_spectral_ws(g.ctx, X, R1), "tdma")      # R1 is free until the peers' next forward pass
            ck(lib.mhh_fft_y_chunks_to_peers(g.ctx, X, self._r2_peers, 0, g.kmax), "fft_y bwd -> peers")
            ck(lib.mhh_peer_barrier(g.ctx), "peer barrier")
            ck(lib.mhh_fft_x_chunks(g.ctx, +1, R2, None, L.ptr(p), 0, g.kmax), "chunks -> fft_x bwd")
            return
        nc = self.kranges if g.kmax % self.kranges == 0 else 1
        nk = g.kmax // nc
        world, me = self.comm.world, self.comm.rank
        R = nk * g.jmax * (g.itot // world)                  # one peer's chunk of one k range (elements)
        main = torch.cuda.current_stream(g.device)
        xfer, recv = self._side
        ce = self.a2a == "ce"
        S = L.ptr(self.S) if ce else None
        on = lambda st: lib.mhh_set_stream(g.ctx, C.c_void_p(st.cuda_stream))

        def half(send, recv_fn, peers_t):
            """send(c): the sending transform of range c on the main stream; then on `xfer`: (ce: the copies,) the barrier;
            then on `recv`: the receiving transform of range c."""
            for c in range(nc):
                on(main)
                send(c)
                ev = torch.cuda.Event(); ev.record(main)
                xfer.wait_event(ev)
                if ce:
                    with torch.cuda.stream(xfer):
                        base = c * world * R
                        for d in range(world):
                            r = (me + d) % world          # start with my own chunk, then round the ring
                            peers_t[r][base + me * R: base + (me + 1) * R].copy_(self.S[base + r * R: base + (r + 1) * R], non_blocking=True)
                on(xfer)
                ck(lib.mhh_peer_barrier(g.ctx), "peer barrier")
                ev2 = torch.cuda.Event(); ev2.record(xfer)
                recv.wait_event(ev2)
                on(recv)
                recv_fn(c)
            ev3 = torch.cuda.Event(); ev3.record(recv)
            main.wait_event(ev3)
            on(main)

        if ce:
            half(lambda c: ck(lib.mhh_fft_x_chunks(g.ctx, -1, X, S, None, c * nk, nk), "fft_x fwd -> chunks"),
                 lambda c: ck(lib.mhh_fft_y_chunks(g.ctx, -1, R1, X, c * nk, nk), "chunks -> fft_y fwd"), self._r1_t)
        else:
            half(lambda c: ck(lib.mhh_fft_x_chunks_to_peers(g.ctx, X, self._r1_peers, c * nk, nk), "fft_x fwd -> peers"),
                 lambda c: ck(lib.mhh_fft_y_chunks(g.ctx, -1, R1, X, c * nk, nk), "chunks -> fft_y fwd"), None)
        ck(lib.mhh_pres_solve_spectral_ws(g.ctx, X, R1), "tdma")
        if ce:
            half(lambda c: ck(lib.mhh_fft_y_chunks(g.ctx, +1, X, S, c * nk, nk), "fft_y bwd -> chunks"),
                 lambda c: ck(lib.mhh_fft_x_chunks(g.ctx, +1, R2, None, L.ptr(p), c * nk, nk), "chunks -> fft_x bwd"), self._r2_t)
        else:
            half(lambda c: ck(lib.mhh_fft_y_chunks_to_peers(g.ctx, X, self._r2_peers, c * nk, nk), "fft_y bwd -> peers"),
                 lambda c: ck(lib.mhh_fft_x_chunks(g.ctx, +1, R2, None, L.ptr(p), c * nk, nk), "chunks -> fft_x bwd"), None)
    _r1_peers = None

    def exec(self, dt, stats=None, timeloop=None, hooks=None):
        g, f, lib = self.grid, self.fields, self.grid.lib
        if self.chunk_fft is None:
            self.chunk_fft = bool(lib.mhh_fft_chunks_supported(g.ctx))
        if hooks and not (self.peer and self.chunk_fft):
            for h in hooks:
                h()
            hooks = None
        u, v, w = (f.mp[n].fld for n in "uvw")
        ut, vt, wt = (f.mt[n].fld for n in "uvw")
        p = f.sd["p"].fld
        A, B = L.ptr(self.A), L.ptr(self.B)
        ck = lambda err, what: L.check(g.ctx, err, what)
        # Pres_2::input: halos of the tendencies first (src/pres_2.cxx:180-181)
        self.cyclic.exec(ut, L.EDGE_EAST_WEST)
        self.cyclic.exec(vt, L.EDGE_NORTH_SOUTH)
        ck(lib.mhh_pres_input_to(g.ctx, L.ptr(u), L.ptr(v), L.ptr(w), L.ptr(ut), L.ptr(vt), L.ptr(wt), C.c_double(dt), A), "pres_input")
        if self.peer and self.chunk_fft:
            self._exec_peer_transforms(p, hooks)
        elif self.chunk_fft:
            # The transforms read / write the all-to-all chunk layout themselves (no pack / unpack passes), and the slab is
            # cut into k ranges: the collective of one range runs on NCCL's stream while the transforms of the next are
            # computed (the tridiagonal solve in the middle needs all levels and is the only join).
            nc = self.kranges if g.kmax % self.kranges == 0 else 1
            nk, seg = g.kmax // nc, (g.kmax // nc) * g.jmax * g.itot
            part = lambda t, c: t[c * seg:(c + 1) * seg]
            works = []
            for c in range(nc):
                ck(lib.mhh_fft_x_chunks(g.ctx, -1, A, B, None, c * nk, nk), "fft_x fwd -> chunks")
                works.append(self.comm.all_to_all_async(part(self.A, c), part(self.B, c)))
            for c in range(nc):
                if works[c] is not None:
                    works[c].wait()
                ck(lib.mhh_fft_y_chunks(g.ctx, -1, A, B, c * nk, nk), "chunks -> fft_y fwd")
            ck(lib.mhh_pres_solve_spectral_ws(g.ctx, B, A), "tdma")
            works = []
            for c in range(nc):
                ck(lib.mhh_fft_y_chunks(g.ctx, +1, B, A, c * nk, nk), "fft_y bwd -> chunks")
                works.append(self.comm.all_to_all_async(part(self.B, c), part(self.A, c)))
            for c in range(nc):
                if works[c] is not None:
                    works[c].wait()
                ck(lib.mhh_fft_x_chunks(g.ctx, +1, B, None, L.ptr(p), c * nk, nk), "chunks -> fft_x bwd")
        else:
            # forward: x FFT, exec_xy, y FFT (src/fft.cxx:452-516 with npx = 1)
            ck(lib.mhh_fft_x(g.ctx, -1, A, None), "fft_x fwd")
            ck(lib.mhh_transpose_copy(g.ctx, 0, B, A), "pack xy")
            self.comm.all_to_all(self.A, self.B)
            ck(lib.mhh_transpose_copy(g.ctx, 1, B, A), "unpack xy")
            ck(lib.mhh_fft_y(g.ctx, -1, B), "fft_y fwd")
            ck(lib.mhh_pres_solve_spectral_ws(g.ctx, B, A), "tdma")
            # backward: y FFT, exec_yx, x FFT scattering into the ghost-celled p (src/fft.cxx:519-583)
            ck(lib.mhh_fft_y(g.ctx, +1, B), "fft_y bwd")
            ck(lib.mhh_transpose_copy(g.ctx, 2, A, B), "pack yx")
            self.comm.all_to_all(self.B, self.A)
            ck(lib.mhh_transpose_copy(g.ctx, 3, A, B), "unpack yx")
            ck(lib.mhh_fft_x(g.ctx, +1, A, L.ptr(p)), "fft_x bwd")
        # Pres_2::solve tail + output (src/pres_2.cxx:352-387)
        ck(lib.mhh_pres_output_stage(g.ctx, 0, L.ptr(ut), L.ptr(vt), L.ptr(wt), L.ptr(p)), "pres_output 0")
        self.cyclic.exec(p, L.EDGE_NORTH_SOUTH)
        if timeloop is not None:       # gradient stage + Timeloop::exec for u,v,w in one pass
            ck(lib.mhh_pres_output_rk(g.ctx, L.ptr(ut), L.ptr(vt), L.ptr(wt), L.ptr(p), L.ptr(u), L.ptr(v), L.ptr(w),
                                      timeloop.rkorder, timeloop.substep, C.c_double(timeloop.dt)), "pres_output_rk")
            return True
        ck(lib.mhh_pres_output_stage(g.ctx, 1, L.ptr(ut), L.ptr(vt), L.ptr(wt), L.ptr(p)), "pres_output 1")
        return False

    def check_divergence(self):
        return self.comm.allreduce_max(super().check_divergence(), self.grid.device)


class SlabModel(Model):
    """Model on one y-slab of an npy = world decomposition: the same Model::exec sequence (fields.exec, exec_viscosity, the
    surface layer, tendencies, Buffer, Force, pressure, RK — Model.exec_substep), with the halo fills going through the ring
    neighbours and every maximum / mean reduced over the ranks (Master::max / Master::sum)."""

    def __init__(self, ini, z, device=None, rank=0, world=1, surface=None, group=None, profs=None, mo=False, path=None,
                 precision="f64"):
        path = path or os.environ.get("MHH_SLAB_PATH", "peer")
        if path not in ("peer", "nccl"):
            raise L.MhhError("MHH_SLAB_PATH / path must be peer or nccl")
        if world > 8:
            path = "nccl"            # the peer tables hold 8 ranks (one NVSwitch box)
        super().__init__(ini, z, device=device, surface=surface, npy=world, rank_y=rank, profs=profs, mo=mo,
                         peer_arena=(path == "peer"), precision=precision)
        self.comm = SlabComm(rank, world, group)
        self.fields.comm = self.comm
        self.cyclic = SlabBoundaryCyclic(self.grid, self.comm)
        self.boundary.boundary_cyclic = self.cyclic
        self.pres = SlabPres(self.grid, self.fields, self.pres.swpres, self.comm, self.cyclic)
        self.path = path
        if path == "peer":
            self.grid.arena.connect(self.comm)       # CUDA IPC handles over torch.distributed; host barrier
            self.cyclic.peer = self.pres.peer = True

    def check_peers(self):
        """Raises when a device-side flag wait of the peer path timed out (a rank was lost)."""
        err = C.c_int()
        L.check(self.grid.ctx, self.grid.lib.mhh_peer_error(self.grid.ctx, C.byref(err)), "mhh_peer_error")
        if err.value:
            raise L.MhhError("peer-memory exchange timed out waiting for a neighbour's flag")

    def _cyclic_prognostic(self):
        self.cyclic.exec_many(fld.fld for fld in self.fields.ap.values())      # set_prognostic_cyclic_bcs, one exchange

    def _after_viscosity(self):
        if self.diff.swdiff == "smag2":
            self.cyclic.exec(self.fields.sd["evisc"].fld, L.EDGE_NORTH_SOUTH)   # east/west was done inside the call

    def _reduce_max(self, values):
        """Master::max over the ranks; several maxima travel in one collective (a NaN flag rides along)."""
        return self.comm.allreduce_max_many(values, self.grid.device)

    def _scalar_halves(self):
        return self.path == "peer" and self.pres.a2a in ("ce_overlap", "ce_sub")

    def print_status(self):
        g, f, tl, dev = self.grid, self.fields, self.timeloop, self.grid.device
        if self.path == "peer":
            self.check_peers()
        for fld in f.mp.values():
            self.cyclic.exec(fld.fld)
        # per-rank sums are already divided by the GLOBAL itot*jtot*zsize inside the library
        sums = self.comm.allreduce_sum([f.check_momentum(), f.check_tke(), f.check_mass()], dev).tolist()
        return dict(iter=tl.iteration, time=tl.time, dt=tl.dt,
                    cfl=self.comm.allreduce_max(self.advec.get_cfl(tl.dt), dev),
                    dn=self.comm.allreduce_max(self.diff.get_dn(tl.dt), dev),
                    div=self.pres.check_divergence(), mom=sums[0], tke=sums[1], mass=sums[2])
