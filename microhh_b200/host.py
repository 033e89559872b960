"""Host-side mirror of the reference's operator interfaces for the dynamical core.

Same class names, factory keys and method names as the reference (``include/advec.h:45-70``,
``include/diff.h:37-71``, ``include/pres.h:39-85``, ``include/boundary.h:74-185``,
``include/timeloop.h``), each method a thin call into the C ABI (``include/mhh_b200.h``).
PyTorch is used only to own device memory (``torch.empty(..., device="cuda")``) and streams.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np
import torch

from . import lib as L


class _RawCuda:
    """A raw device allocation seen through the CUDA array interface (torch.as_tensor wraps it without a copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerArena:
    """The exchange arena of a y-slab rank (include/mhh_b200.h, "peer-memory data path"): one cudaMalloc'ed block that holds
    every array other ranks write into — the halo-exchanged fields and planes, the pressure solver's receive buffers and the
    flag words — at the SAME offsets on all ranks (all ranks allocate in the same order), so that the address of any of my
    arrays in rank r's arena is  bases[r] + (ptr - bases[me])."""
    ALIGN = 512

    def __init__(self, grid, nbytes):
        self.grid, self.nbytes = grid, int(nbytes)
        p = C.c_void_p()
        L.check(grid.ctx, grid.lib.mhh_peer_alloc(grid.ctx, C.c_ulonglong(self.nbytes), C.byref(p)), "mhh_peer_alloc")
        self.base = p.value
        self.raw = torch.as_tensor(_RawCuda(self.base, self.nbytes), device=grid.device)
        assert self.raw.data_ptr() == self.base
        self.flags_offset = 0
        self.top = 4096                      # the first page: 32 flag words
        self.bases = None                    # filled by connect()
        self.mapped = []

    def alloc(self, shape, dtype=torch.float64):
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        off = (self.top + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        if off + n > self.nbytes:
            return None
        self.top = off + n
        return self.raw[off:off + n].view(dtype).view(*shape)      # zero-filled by mhh_peer_alloc

    def contains(self, t):
        return self.base <= t.data_ptr() and t.data_ptr() + t.numel() * t.element_size() <= self.base + self.nbytes

    def connect(self, comm):
        """Exchange the CUDA IPC handles over torch.distributed, map the other ranks' arenas, register the table."""
        g = self.grid
        world, me = comm.world, comm.rank
        bases = [None] * world
        bases[me] = self.base
        if world > 1:
            import torch.distributed as dist
            h = (C.c_ubyte * 64)()
            L.check(g.ctx, g.lib.mhh_peer_export(g.ctx, C.c_void_p(self.base), h), "mhh_peer_export")
            handles = [None] * world
            dist.all_gather_object(handles, (bytes(h), self.nbytes, self.top), group=comm.group)
            for r, (hb, nb, top) in enumerate(handles):
                if nb != self.nbytes or top != self.top:
                    raise L.MhhError("peer arenas differ between ranks (all ranks must allocate the same arrays in the same order)")
                if r == me:
                    continue
                q = C.c_void_p()
                hr = (C.c_ubyte * 64).from_buffer_copy(hb)
                L.check(g.ctx, g.lib.mhh_peer_import(g.ctx, hr, C.byref(q)), "mhh_peer_import")
                bases[r] = q.value
                self.mapped.append(q.value)
        arr = (C.c_void_p * world)(*bases)
        L.check(g.ctx, g.lib.mhh_peer_register(g.ctx, world, me, C.c_void_p(self.base), C.c_ulonglong(self.nbytes), arr,
                                               C.c_ulonglong(self.flags_offset)), "mhh_peer_register")
        self.bases = bases
        if world > 1:
            import torch.distributed as dist
            dist.barrier(group=comm.group)

    def peer_ptrs(self, t):
        """C array: the address of my tensor `t` in every rank's arena."""
        off = t.data_ptr() - self.base
        return (C.c_void_p * len(self.bases))(*[b + off for b in self.bases])

    def close(self):
        g = self.grid
        if getattr(self, "base", None) and getattr(g, "ctx", None):
            self.raw = None
            for q in self.mapped:
                g.lib.mhh_peer_unmap(g.ctx, C.c_void_p(q))
            g.lib.mhh_peer_free(g.ctx, C.c_void_p(self.base))
            self.base = None


class Grid:
    """Grid / Grid_data (include/grid.h:50-135). ``settings`` holds the [grid] keys of the .ini."""

    def __init__(self, itot, jtot, ktot, xsize, ysize, zsize, z, swspatialorder="2", gc=(0, 0, 0),
                 npy=1, rank_y=0, device=None, peer_arena=False, precision="f64"):
        """precision: "f64" (the reference's default build) or "f32" (its -DUSESP build: the `_f32` symbol set of the library,
        every field and profile in float; time steps stay double)."""
        if not torch.cuda.is_available():
            raise L.MhhError("microhh_b200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.precision = precision
        self.lib = L.Api(precision)
        self.dtype = torch.float64 if precision == "f64" else torch.float32
        self.npdtype = np.float64 if precision == "f64" else np.float32
        order = int(swspatialorder)
        if order not in (2, 4):
            raise L.MhhError(f"{swspatialorder} is an illegal value for swspatialorder")
        d = self.lib.GridDesc(itot, jtot, ktot, gc[0], gc[1], gc[2], xsize, ysize, zsize, order, npy, rank_y)
        z = np.ascontiguousarray(z, dtype=self.npdtype)
        assert z.shape == (ktot,)
        self.ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            err = self.lib.mhh_create(C.byref(d), z.ctypes.data_as(C.c_void_p), C.byref(self.ctx))
        if err:
            raise L.MhhError(f"mhh_create failed with code {err}")
        if os.environ.get("MHH_TMA") == "0":       # experiments: per-thread cp.async plane staging instead of TMA
            self.lib.mhh_set_tma(self.ctx, 0)
        dims = (C.c_int * 20)()
        self.lib.mhh_get_dims(self.ctx, dims)
        (self.itot, self.jtot, self.ktot, self.imax, self.jmax, self.kmax, self.igc, self.jgc, self.kgc,
         self.icells, self.jcells, self.kcells, self.ijcells, self.ncells,
         self.istart, self.iend, self.jstart, self.jend, self.kstart, self.kend) = (int(x) for x in dims)
        self.order = order
        self.npy, self.rank_y = npy, rank_y
        self.xsize, self.ysize, self.zsize = xsize, ysize, zsize
        self.dx, self.dy = xsize / itot, ysize / jtot
        self.shape = (self.kcells, self.jcells, self.icells)
        self.shape2d = (self.jcells, self.icells)
        for name in ("z", "zh", "dz", "dzh", "dzi", "dzhi", "dzi4", "dzhi4"):
            setattr(self, name, self.metric(name))
        # y-slab rank with the peer-memory data path: every array made by zeros() / zeros2d() lives in the exchange arena
        self.arena = None
        if peer_arena:
            nfree = self.imax * self.jmax * self.kmax
            esize = 8 if precision == "f64" else 4
            self.arena = PeerArena(self, esize * (16 * self.ncells + 3 * nfree + 96 * self.ijcells) + (8 << 20))

    def metric(self, name):
        a = np.zeros(self.kcells, dtype=self.npdtype)
        L.check(self.ctx, self.lib.mhh_get_metric(self.ctx, name.encode(), a.ctypes.data_as(C.c_void_p)), "mhh_get_metric")
        return a

    def set_rhoref(self, rhoref, rhorefh):
        rhoref = np.ascontiguousarray(rhoref, dtype=self.npdtype)
        rhorefh = np.ascontiguousarray(rhorefh, dtype=self.npdtype)
        assert rhoref.shape == (self.kcells,) and rhorefh.shape == (self.kcells,)
        L.check(self.ctx, self.lib.mhh_set_rhoref(self.ctx, rhoref.ctypes.data_as(C.c_void_p),
                                                  rhorefh.ctypes.data_as(C.c_void_p)), "mhh_set_rhoref")

    # launch accounting / live per-kernel CUDA-event timing (bench.py)
    def launch_count(self):
        return int(self.lib.mhh_launch_count(self.ctx))

    def profile_enable(self, on=True):
        L.check(self.ctx, self.lib.mhh_profile_enable(self.ctx, 1 if on else 0), "mhh_profile_enable")

    def profile_report(self):
        """{kernel name: (total ms, launches)} since the last report."""
        names = C.create_string_buffer(8192)
        ms = (C.c_double * 256)()
        cnt = (C.c_long * 256)()
        n = self.lib.mhh_profile_report(self.ctx, names, 8192, ms, cnt, 256)
        keys = [k for k in names.value.decode().split(";") if k]
        return {keys[i]: (ms[i], int(cnt[i])) for i in range(min(n, len(keys)))}

    def zeros(self):
        t = self.arena.alloc(self.shape, self.dtype) if self.arena is not None else None
        return t if t is not None else torch.zeros(self.shape, dtype=self.dtype, device=self.device)

    def zeros2d(self):
        t = self.arena.alloc(self.shape2d, self.dtype) if self.arena is not None else None
        return t if t is not None else torch.zeros(self.shape2d, dtype=self.dtype, device=self.device)

    def close(self):
        if getattr(self, "ctx", None):
            if getattr(self, "arena", None) is not None:
                torch.cuda.synchronize(self.device)
                self.arena.close()
                self.arena = None
            self.lib.mhh_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Field3d:
    """Field3d (include/field3d.h:47-76): a 3-D array plus six 2-D boundary planes."""

    def __init__(self, grid: Grid, name: str, visc: float = 0.0):
        self.name, self.visc = name, visc
        self.fld = grid.zeros()
        self.fld_mean = torch.zeros(grid.kcells, dtype=grid.dtype, device=self.fld.device)   # Field3d::fld_mean
        self.fld_bot, self.fld_top = grid.zeros2d(), grid.zeros2d()
        self.grad_bot, self.grad_top = grid.zeros2d(), grid.zeros2d()
        self.flux_bot, self.flux_top = grid.zeros2d(), grid.zeros2d()


class Fields:
    """Fields (include/fields.h:132-182): the named maps mp/mt/sp/st/sd/ap/at."""

    def __init__(self, grid: Grid, slist=(), visc=0.0, svisc=0.0):
        self.grid, self.visc = grid, visc
        self.mp, self.mt, self.sp, self.st, self.sd = {}, {}, {}, {}, {}
        for n in ("u", "v", "w"):
            self.mp[n] = Field3d(grid, n, visc)
            self.mt[n] = Field3d(grid, n + "t")
        for n in slist:       # svisc: one value for all scalars, or {name: value}
            self.init_prognostic_field(n, svisc.get(n, 0.) if isinstance(svisc, dict) else svisc)
        self.sd["p"] = Field3d(grid, "p")
        self.rhoref = np.ones(grid.kcells)
        self.rhorefh = np.ones(grid.kcells)
        self.calc_mean_profs = False
        self.comm = None          # set by a decomposed model (multigpu.SlabComm): sums over ranks as Master::sum does

    def set_calc_mean_profs(self, sw):
        self.calc_mean_profs = bool(sw)

    def exec(self):
        """Fields::exec, src/fields.cxx: xy-mean profile of every prognostic field into fld_mean (when a consumer asked)."""
        if not self.calc_mean_profs:
            return
        g = self.grid
        prof = np.zeros(g.kcells, dtype=g.npdtype)
        if self.comm is not None and self.comm.world > 1:
            # calc_mean_profile on MPI: per-rank level sums, master.sum, then the division (src/field3d_operators.cxx:46-66)
            flds = list(self.ap.values())
            sums = np.zeros((len(flds), g.kcells), dtype=g.npdtype)
            for q, fld in enumerate(flds):
                L.check(g.ctx, g.lib.mhh_level_sums(g.ctx, L.ptr(fld.fld), sums[q].ctypes.data_as(C.c_void_p)), "mhh_level_sums")
            tot = self.comm.allreduce_sum(sums, g.device).reshape(len(flds), g.kcells) / (g.itot * g.jtot)
            for q, fld in enumerate(flds):
                fld.fld_mean.copy_(tot[q].to(g.dtype))
            return
        for fld in self.ap.values():
            L.check(g.ctx, g.lib.mhh_mean_profile(g.ctx, L.ptr(fld.fld), prof.ctypes.data_as(C.c_void_p)), "mhh_mean_profile")
            fld.fld_mean.copy_(torch.from_numpy(prof))       # prof has the working type

    def init_prognostic_field(self, name, visc):
        self.sp[name] = Field3d(self.grid, name, visc)
        self.st[name] = Field3d(self.grid, name + "t")

    def init_diagnostic_field(self, name):
        if name not in self.sd:
            self.sd[name] = Field3d(self.grid, name)

    @property
    def ap(self):
        return {**self.mp, **self.sp}

    @property
    def at(self):
        return {**self.mt, **self.st}

    def set_rhoref(self, rhoref, rhorefh):
        self.rhoref, self.rhorefh = np.asarray(rhoref, dtype=self.grid.npdtype), np.asarray(rhorefh, dtype=self.grid.npdtype)
        self.grid.set_rhoref(self.rhoref, self.rhorefh)

    # Fields::check_momentum/tke/mass, src/fields.cxx:1175-1217
    def _scalar(self, fn, *flds):
        out = C.c_double()
        L.check(self.grid.ctx, fn(self.grid.ctx, *[L.ptr(f) for f in flds], C.byref(out)), fn.__name__)
        return out.value

    def check_momentum(self):
        return self._scalar(self.grid.lib.mhh_check_momentum, self.mp["u"].fld, self.mp["v"].fld, self.mp["w"].fld)

    def check_tke(self):
        return self._scalar(self.grid.lib.mhh_check_tke, self.mp["u"].fld, self.mp["v"].fld, self.mp["w"].fld)

    def check_mass(self):
        if not self.sp:
            return 0.0
        first = sorted(self.sp)[0]          # std::map iteration order
        return self._scalar(self.grid.lib.mhh_check_mass, self.sp[first].fld)


class Field3d_operators:
    """Field3d_operators (include/field3d_operators.h:40-51)."""

    def __init__(self, grid: Grid):
        self.grid = grid

    def calc_mean_profile(self, fld):
        prof = np.zeros(self.grid.kcells, dtype=self.grid.npdtype)
        L.check(self.grid.ctx, self.grid.lib.mhh_mean_profile(self.grid.ctx, L.ptr(fld), prof.ctypes.data_as(C.c_void_p)),
                "mhh_mean_profile")
        return prof

    def _scalar(self, fn, fld):
        out = C.c_double()
        L.check(self.grid.ctx, fn(self.grid.ctx, L.ptr(fld), C.byref(out)), fn.__name__)
        return out.value

    def calc_mean(self, fld):
        return self._scalar(self.grid.lib.mhh_mean, fld)

    def calc_mean_2d(self, fld2d):
        return self._scalar(self.grid.lib.mhh_mean_2d, fld2d)

    def calc_max(self, fld):
        return self._scalar(self.grid.lib.mhh_max, fld)


class Stats:
    """The profile statistics of Stats (include/stats.h, src/stats.cxx:264-398) over the default mask: xy means, central
    moments and covariances per level, computed on the device.  Profiles are numpy arrays of kcells values (levels
    kstart..kend filled).  On a y-slab rank pass ``comm`` (multigpu.SlabComm): the ranks' profiles are summed (master.sum)."""

    def __init__(self, grid: Grid, comm=None):
        self.grid, self.comm = grid, comm

    def _reduce(self, prof):
        if self.comm is not None and self.comm.world > 1:
            return self.comm.allreduce_sum(prof, self.grid.device).cpu().numpy()
        return prof

    def calc_mean(self, fld):
        g = self.grid
        prof = np.zeros(g.kcells, dtype=g.npdtype)
        L.check(g.ctx, g.lib.mhh_stats_mean(g.ctx, L.ptr(fld), prof.ctypes.data_as(C.c_void_p)), "mhh_stats_mean")
        return self._reduce(prof)

    def calc_moment(self, fld, fld_mean, power, offset=0.):
        g = self.grid
        prof, mean = np.zeros(g.kcells, dtype=g.npdtype), np.ascontiguousarray(fld_mean, dtype=g.npdtype)
        L.check(g.ctx, g.lib.mhh_stats_moment(g.ctx, L.ptr(fld), mean.ctypes.data_as(C.c_void_p), C.c_double(offset), int(power),
                                              prof.ctypes.data_as(C.c_void_p)), "mhh_stats_moment")
        return self._reduce(prof)

    def calc_cov(self, fld1, mean1, fld2, mean2, pow1=1, pow2=1, offset1=0., offset2=0.):
        g = self.grid
        prof = np.zeros(g.kcells, dtype=g.npdtype)
        m1, m2 = np.ascontiguousarray(mean1, dtype=g.npdtype), np.ascontiguousarray(mean2, dtype=g.npdtype)
        L.check(g.ctx, g.lib.mhh_stats_cov(g.ctx, L.ptr(fld1), m1.ctypes.data_as(C.c_void_p), C.c_double(offset1), int(pow1),
                                           L.ptr(fld2), m2.ctypes.data_as(C.c_void_p), C.c_double(offset2), int(pow2),
                                           prof.ctypes.data_as(C.c_void_p)), "mhh_stats_cov")
        return self._reduce(prof)

    def profiles(self, fields):
        """Means and variances of every prognostic field, and the resolved flux w'th' when th exists."""
        out = {}
        for n, f in fields.ap.items():
            out[n] = self.calc_mean(f.fld)
            out[n + "_2"] = self.calc_moment(f.fld, out[n], 2)
        if "th" in fields.sp:
            out["wth"] = self.calc_cov(fields.mp["w"].fld, np.zeros(self.grid.kcells), fields.sp["th"].fld, out["th"])
        return out


class Boundary_cyclic:
    """Boundary_cyclic (include/boundary_cyclic.h:36-49)."""

    def __init__(self, grid: Grid):
        self.grid = grid

    def exec(self, fld, edge=L.EDGE_BOTH):
        L.check(self.grid.ctx, self.grid.lib.mhh_cyclic(self.grid.ctx, L.ptr(fld), edge), "mhh_cyclic")

    def exec_2d(self, fld2d):
        L.check(self.grid.ctx, self.grid.lib.mhh_cyclic_2d(self.grid.ctx, L.ptr(fld2d)), "mhh_cyclic_2d")


_BC = {"dirichlet": L.BC_DIRICHLET, "neumann": L.BC_NEUMANN, "flux": L.BC_FLUX, "ustar": L.BC_USTAR, "off": L.BC_OFF,
       "noslip": L.BC_DIRICHLET, "freeslip": L.BC_NEUMANN}


@dataclass
class Field3dBc:
    """Field3dBc (include/boundary.h:60-67)."""
    bot: float = 0.0
    top: float = 0.0
    bcbot: int = L.BC_NEUMANN
    bctop: int = L.BC_NEUMANN


class Boundary:
    """Boundary base class = swboundary=default (include/boundary.h:74-185, src/boundary.cxx)."""

    def __init__(self, grid: Grid, fields: Fields, mbcbot="noslip", mbctop="noslip", ubot=0., utop=0., vbot=0., vtop=0.,
                 sbc=None, swboundary="default"):
        self.grid, self.fields = grid, fields
        self.swboundary = swboundary
        self.mbcbot, self.mbctop = _BC[mbcbot], _BC[mbctop]
        self.ubot, self.utop, self.vbot, self.vtop = ubot, utop, vbot, vtop
        self.sbc = {}
        for name in fields.sp:
            spec = (sbc or {}).get(name, {})
            self.sbc[name] = Field3dBc(spec.get("sbot", 0.), spec.get("stop", 0.),
                                       _BC[spec.get("sbcbot", "neumann")], _BC[spec.get("sbctop", "neumann")])
        self.boundary_cyclic = Boundary_cyclic(grid)

    def get_switch(self):
        return self.swboundary

    def _set_bc(self, a, agrad, aflux, sw, aval, visc, offset):
        L.check(self.grid.ctx, self.grid.lib.mhh_set_bc(self.grid.ctx, L.ptr(a), L.ptr(agrad), L.ptr(aflux), sw,
                                                        C.c_double(aval), C.c_double(visc), C.c_double(offset)), "mhh_set_bc")

    def set_values(self):
        """Boundary::set_values, src/boundary.cxx:598-682 (no Galilean offset: utrans=vtrans=0)."""
        f = self.fields
        for name, bot, top in (("u", self.ubot, self.utop), ("v", self.vbot, self.vtop)):
            fld = f.mp[name]
            self._set_bc(fld.fld_bot, fld.grad_bot, fld.flux_bot, self.mbcbot, bot, f.visc, 0.)
            self._set_bc(fld.fld_top, fld.grad_top, fld.flux_top, self.mbctop, top, f.visc, 0.)
        for name, fld in f.sp.items():
            bc = self.sbc[name]
            self._set_bc(fld.fld_bot, fld.grad_bot, fld.flux_bot, bc.bcbot, bc.bot, fld.visc, 0.)
            self._set_bc(fld.fld_top, fld.grad_top, fld.flux_top, bc.bctop, bc.top, fld.visc, 0.)

    def set_prognostic_cyclic_bcs(self):
        """src/boundary.cxx:447-458"""
        for fld in self.fields.ap.values():
            self.boundary_cyclic.exec(fld.fld)

    def _ghost(self, fld, bcbot, bctop):
        L.check(self.grid.ctx, self.grid.lib.mhh_ghost_cells(self.grid.ctx, L.ptr(fld.fld), bcbot, bctop,
                                                             L.ptr(fld.fld_bot), L.ptr(fld.fld_top),
                                                             L.ptr(fld.grad_bot), L.ptr(fld.grad_top)), "mhh_ghost_cells")

    def set_ghost_cells(self):
        """Boundary::set_ghost_cells, src/boundary.cxx:919-983 (update_slave_bcs is host bookkeeping, not needed here)."""
        f = self.fields
        self._ghost(f.mp["u"], self.mbcbot, self.mbctop)
        self._ghost(f.mp["v"], self.mbcbot, self.mbctop)
        if self.grid.order == 4:
            self.set_ghost_cells_w(L.BCW_NORMAL)
        for name, fld in f.sp.items():
            self._ghost(fld, self.sbc[name].bcbot, self.sbc[name].bctop)

    def set_ghost_cells_w(self, boundary_w_type):
        """src/boundary.cxx:986-1007"""
        L.check(self.grid.ctx, self.grid.lib.mhh_ghost_cells_w(self.grid.ctx, L.ptr(self.fields.mp["w"].fld), boundary_w_type),
                "mhh_ghost_cells_w")


class Advec:
    """Advec (include/advec.h:45-70); factory keyed on [advec] swadvec (src/advec.cxx:62-83)."""
    _schemes = {"2": L.ADVEC_2, "2i5": L.ADVEC_2I5, "4": L.ADVEC_4}

    def __init__(self, grid, fields, swadvec, cflmax=1.0, fluxlimit_list=()):
        self.grid, self.fields = grid, fields
        self.swadvec, self.scheme = swadvec, self._schemes[swadvec]
        self.cflmax, self.cflmin = cflmax, 1.e-5
        self.fluxlimit_list = tuple(fluxlimit_list)

    @staticmethod
    def factory(grid, fields, swadvec=None, **kw):
        if swadvec is None:
            swadvec = str(grid.order)
        if swadvec not in Advec._schemes:
            raise L.MhhError(f"{swadvec} is an illegal value for swadvec")
        return Advec(grid, fields, swadvec, **kw)

    def get_switch(self):
        return self.swadvec

    def exec(self, stats=None):
        f, g = self.fields, self.grid
        names = list(f.sp)
        s = L.ptr_array([f.sp[n].fld for n in names])
        st = L.ptr_array([f.st[n].fld for n in names])
        flags = (C.c_int * max(len(names), 1))(*[1 if n in self.fluxlimit_list else 0 for n in names])
        L.check(g.ctx, g.lib.mhh_advec_exec(g.ctx, self.scheme, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld),
                                            L.ptr(f.mt["u"].fld), L.ptr(f.mt["v"].fld), L.ptr(f.mt["w"].fld),
                                            len(names), s, st, flags), "mhh_advec_exec")

    def get_cfl(self, dt):
        f, g = self.fields, self.grid
        out = C.c_double()
        L.check(g.ctx, g.lib.mhh_advec_cfl(g.ctx, self.scheme, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld),
                                           C.c_double(dt), C.byref(out)), "mhh_advec_cfl")
        return out.value

    def get_time_limit(self, idt, dt):
        return self.time_limit_from(self.get_cfl(dt) , idt)

    def time_limit_from(self, cfl, idt):
        """Advec::get_time_limit (src/advec_2.cxx:281-294) from an already reduced cfl = dt * max(...)."""
        return int(idt * self.cflmax / max(self.cflmin, cfl))


class Diff:
    """Diff (include/diff.h:37-71); factory keyed on [diff] swdiff (src/diff.cxx:57-85)."""

    def __init__(self, grid, fields, boundary, swdiff, dnmax=0.4, cs=0.23, tPr=1. / 3.):
        self.grid, self.fields, self.boundary = grid, fields, boundary
        self.swdiff, self.dnmax, self.cs, self.tPr = swdiff, dnmax, cs, tPr
        if swdiff == "smag2":
            fields.init_diagnostic_field("evisc")
        self.dnmul = 0.0

    @staticmethod
    def factory(grid, fields, boundary, swdiff=None, **kw):
        if swdiff is None:
            swdiff = str(grid.order)
        if swdiff not in ("0", "2", "4", "smag2"):
            raise L.MhhError(f"{swdiff} is an illegal value for swdiff")
        return Diff(grid, fields, boundary, swdiff, **kw)

    def get_switch(self):
        return self.swdiff

    def create(self):
        """static dnmul, src/diff_2.cxx:139-151 / src/diff_smag2.cxx:921-936"""
        g, f = self.grid, self.fields
        viscmax = max([f.visc] + [s.visc for s in f.sp.values()])
        out = C.c_double()
        L.check(g.ctx, g.lib.mhh_diff_const_dnmul(g.ctx, C.c_double(viscmax), C.byref(out)), "mhh_diff_const_dnmul")
        self.dnmul = out.value

    def _surface_model(self):
        return 1 if self.boundary.get_switch() != "default" else 0

    def exec_viscosity(self, thermo=None, surface=None, limits=False):
        """Diff_smag2::exec_viscosity, src/diff_smag2.cxx:1046. ``thermo`` supplies N2 (get_thermo_field("N2"));
        ``surface`` supplies z0m/dudz/dvdz/dbdz planes when a surface model is on.
        ``limits``: also return (max of calc_cfl before the * dt, dnmul) taken in the same pass (the two maxima
        Model::set_time_step needs right after this call, src/model.cxx:354-357); None when the fused pass does not apply."""
        if self.swdiff != "smag2":
            return None
        g, f = self.grid, self.fields
        sm = self._surface_model()
        planes = surface or {}
        if limits and thermo is not None and getattr(thermo, "fuse_N2", False):
            cfl_dt, dnmul = C.c_double(), C.c_double()
            L.check(g.ctx, g.lib.mhh_smag2_exec_viscosity_dry_limits(
                g.ctx, sm, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld), L.ptr(f.sp["th"].fld),
                L.ptr(planes.get("dudz")), L.ptr(planes.get("dvdz")), L.ptr(planes.get("dbdz")), L.ptr(planes.get("z0m")),
                C.c_double(self.cs), C.c_double(self.tPr), C.c_double(f.visc), L.ptr(f.sd["evisc"].fld),
                C.byref(cfl_dt), C.byref(dnmul)), "mhh_smag2_exec_viscosity_dry_limits")
            return cfl_dt.value, dnmul.value
        if thermo is not None and getattr(thermo, "fuse_N2", False):
            # Thermo_dry: N2 is evaluated inside the evisc kernel from th (no N2 field round trip through HBM)
            L.check(g.ctx, g.lib.mhh_smag2_exec_viscosity_dry(
                g.ctx, sm, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld), L.ptr(f.sp["th"].fld),
                L.ptr(planes.get("dudz")), L.ptr(planes.get("dvdz")), L.ptr(planes.get("dbdz")), L.ptr(planes.get("z0m")),
                C.c_double(self.cs), C.c_double(self.tPr), C.c_double(f.visc), L.ptr(f.sd["evisc"].fld)), "mhh_smag2_exec_viscosity_dry")
            return
        N2 = thermo.get_thermo_field("N2") if thermo is not None and thermo.get_switch() != "0" else None
        L.check(g.ctx, g.lib.mhh_smag2_exec_viscosity(
            g.ctx, sm, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld), L.ptr(N2),
            L.ptr(planes.get("dudz")), L.ptr(planes.get("dvdz")), L.ptr(planes.get("dbdz")), L.ptr(planes.get("z0m")),
            C.c_double(self.cs), C.c_double(self.tPr), C.c_double(f.visc), L.ptr(f.sd["evisc"].fld)), "mhh_smag2_exec_viscosity")

    def exec(self, stats=None):
        g, f = self.grid, self.fields
        if self.swdiff == "0":
            return
        names = list(f.sp)
        s = L.ptr_array([f.sp[n].fld for n in names])
        st = L.ptr_array([f.st[n].fld for n in names])
        svisc = (g.lib.real * max(len(names), 1))(*[f.sp[n].visc for n in names])
        u, v, w = (L.ptr(f.mp[n].fld) for n in "uvw")
        ut, vt, wt = (L.ptr(f.mt[n].fld) for n in "uvw")
        if self.swdiff == "2":
            L.check(g.ctx, g.lib.mhh_diff2_exec(g.ctx, u, v, w, ut, vt, wt, C.c_double(f.visc), len(names), s, st, svisc), "mhh_diff2_exec")
        elif self.swdiff == "4":
            L.check(g.ctx, g.lib.mhh_diff4_exec(g.ctx, u, v, w, ut, vt, wt, C.c_double(f.visc), len(names), s, st, svisc), "mhh_diff4_exec")
        else:
            sfb = L.ptr_array([f.sp[n].flux_bot for n in names])
            sft = L.ptr_array([f.sp[n].flux_top for n in names])
            L.check(g.ctx, g.lib.mhh_smag2_exec(
                g.ctx, self._surface_model(), u, v, w, L.ptr(f.sd["evisc"].fld), ut, vt, wt,
                L.ptr(f.mp["u"].flux_bot), L.ptr(f.mp["u"].flux_top), L.ptr(f.mp["v"].flux_bot), L.ptr(f.mp["v"].flux_top),
                C.c_double(f.visc), len(names), s, st, sfb, sft, svisc, C.c_double(self.tPr)), "mhh_smag2_exec")

    def exec_with_advec2(self, th_buoy=None, buffer=None, timeloop=None, parts=None):
        """Thermo_dry::exec (when th_buoy is given) + Advec_2::exec + Diff_smag2::exec in one pass per prognostic field
        (mhh_advec2_smag2_exec); bit-identical to the three calls in the time loop's order.
        With ``timeloop`` (and optionally ``buffer``) the scalars' Buffer term and RK update ride in their pass
        (mhh_advec2_smag2_exec_rk): the updated scalar lands in Field3d.fld_alt and the two tensors are swapped here, so on
        return ``fields.sp[name].fld`` is the advanced field (ghost cells stale until the next set_*_bcs, as after
        Timeloop::exec).  Only valid when nothing else adds to the scalar tendencies afterwards.
        ``parts`` = (do_uvw, do_scalars, kc0, nkc): only the momentum pass and / or the scalar passes on the 64-level chunks
        [kc0, kc0 + nkc) (mhh_advec2_smag2_exec_rk_parts); the caller swaps fld / fld_alt (swap_scalars) after the last part."""
        g, f = self.grid, self.fields
        if self.swdiff != "smag2" or g.order != 2:
            raise L.MhhError("the fused tendency pass needs swdiff=smag2 on the 2nd-order grid")
        names = list(f.sp)
        s = L.ptr_array([f.sp[n].fld for n in names])
        st = L.ptr_array([f.st[n].fld for n in names])
        svisc = (g.lib.real * max(len(names), 1))(*[f.sp[n].visc for n in names])
        sfb = L.ptr_array([f.sp[n].flux_bot for n in names])
        sft = L.ptr_array([f.sp[n].flux_top for n in names])
        args = (g.ctx, self._surface_model(), L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld), L.ptr(f.sd["evisc"].fld),
                L.ptr(f.mt["u"].fld), L.ptr(f.mt["v"].fld), L.ptr(f.mt["w"].fld),
                L.ptr(f.mp["u"].flux_bot), L.ptr(f.mp["u"].flux_top), L.ptr(f.mp["v"].flux_bot), L.ptr(f.mp["v"].flux_top),
                C.c_double(f.visc), L.ptr(th_buoy), len(names), s, st, sfb, sft, svisc, C.c_double(self.tPr))
        if timeloop is None:
            L.check(g.ctx, g.lib.mhh_advec2_smag2_exec(*args), "mhh_advec2_smag2_exec")
            return
        for n in names:
            if getattr(f.sp[n], "fld_alt", None) is None:
                f.sp[n].fld_alt = g.zeros()
        s_new = L.ptr_array([f.sp[n].fld_alt for n in names])
        sbuf = None
        if buffer is not None and buffer.swbuffer:
            sbuf = L.ptr_array([f.sp[n].fld_mean if buffer.swupdate else buffer.bufferprofs[n] for n in names])
        if parts is not None:
            L.check(g.ctx, g.lib.mhh_advec2_smag2_exec_rk_parts(*args, sbuf, s_new, timeloop.rkorder, timeloop.substep,
                                                                C.c_double(timeloop.dt), *[int(x) for x in parts]),
                    "mhh_advec2_smag2_exec_rk_parts")
            return
        L.check(g.ctx, g.lib.mhh_advec2_smag2_exec_rk(*args, sbuf, s_new, timeloop.rkorder, timeloop.substep, C.c_double(timeloop.dt)),
                "mhh_advec2_smag2_exec_rk")
        self.swap_scalars()

    def swap_scalars(self):
        """After the scalars' fused RK update: Field3d.fld_alt holds the advanced scalar."""
        for fld in self.fields.sp.values():
            fld.fld, fld.fld_alt = fld.fld_alt, fld.fld

    def exec_with_advec4(self):
        """Advec_4::exec + set_ghost_cells_w(Normal) + Diff_4::exec (src/model.cxx:387-392) through mhh_advec4_diff4_exec:
        diff_c / diff_w of u, v, w rides in the advection pass; bit-identical to the separate calls.  w must hold Conservation-type
        ghost rows on entry."""
        g, f = self.grid, self.fields
        if self.swdiff != "4" or g.order != 4:
            raise L.MhhError("the fused 4th-order tendency pass needs swdiff=4 on the 4th-order grid")
        names = list(f.sp)
        s = L.ptr_array([f.sp[n].fld for n in names])
        st = L.ptr_array([f.st[n].fld for n in names])
        svisc = (g.lib.real * max(len(names), 1))(*[f.sp[n].visc for n in names])
        u, v, w = (L.ptr(f.mp[n].fld) for n in "uvw")
        ut, vt, wt = (L.ptr(f.mt[n].fld) for n in "uvw")
        L.check(g.ctx, g.lib.mhh_advec4_diff4_exec(g.ctx, u, v, w, ut, vt, wt, C.c_double(f.visc), len(names), s, st, svisc),
                "mhh_advec4_diff4_exec")

    def _dnmul(self):
        if self.swdiff != "smag2":
            return self.dnmul
        g, f = self.grid, self.fields
        out = C.c_double()
        L.check(g.ctx, g.lib.mhh_smag2_dnmul(g.ctx, L.ptr(f.sd["evisc"].fld), C.c_double(self.tPr), C.byref(out)), "mhh_smag2_dnmul")
        return out.value

    def get_dn(self, dt):
        return self._dnmul() * dt

    def get_time_limit(self, idt, dt):
        if self.swdiff == "0":
            return int(1e18)
        return self.time_limit_from(self._dnmul(), idt, dt)

    def time_limit_from(self, dnmul, idt, dt):
        """Diff::get_time_limit (src/diff_smag2.cxx:884-902, src/diff_2.cxx) from an already reduced dnmul."""
        dnmul = max(1.e-9, dnmul) if self.swdiff == "smag2" else dnmul
        return int(idt * self.dnmax / (dt * dnmul))


class Pres:
    """Pres (include/pres.h:39-85); factory keyed on [pres] swpres (src/pres.cxx:66-84)."""

    def __init__(self, grid, fields, swpres):
        self.grid, self.fields, self.swpres = grid, fields, swpres
        self._side, self._side_done = None, None
        self.overlap_hooks = os.environ.get("MHH_SCALAR_OVERLAP", "0") != "0"

    @staticmethod
    def factory(grid, fields, swpres=None):
        if swpres is None:
            swpres = str(grid.order)
        if swpres not in ("0", "2", "4"):
            raise L.MhhError(f"{swpres} is an illegal value for swpres")
        if swpres != "0" and int(swpres) != grid.order:
            raise L.MhhError("swpres must match swspatialorder (ghost-cell depth)")
        return Pres(grid, fields, swpres)

    def set_values(self):
        g = self.grid
        L.check(g.ctx, g.lib.mhh_pres_set_values(g.ctx), "mhh_pres_set_values")

    def exec(self, dt, stats=None, timeloop=None, hooks=None):
        """Pres::exec.  With `timeloop` the gradient stage is fused with Timeloop::exec for u,v,w (mhh_pres_exec_rk):
        returns True, and the caller must run timeloop.exec(skip=("u", "v", "w")) and must NOT call
        set_ghost_cells_w(Normal) in between (4th order: the fused call has done it on the not yet updated w)."""
        g, f = self.grid, self.fields
        fused = timeloop is not None and self.swpres in ("2", "4") and g.npy == 1
        if hooks and fused and self.overlap_hooks:
            # `hooks` (the scalars' tendency + RK passes) read u,v,w but nothing the solver writes before its last kernel: they
            # run on a second stream beside the solver's input, transforms and vertical solve (HBM-bound passes next to
            # latency-bound ones), and the kernel that updates u,v,w waits for them (mhh_pres_set_output_wait)
            if self._side is None:
                self._side = torch.cuda.Stream(device=g.device)
            main, side = torch.cuda.current_stream(g.device), self._side
            ev = torch.cuda.Event(); ev.record(main)
            side.wait_event(ev)
            g.lib.mhh_set_stream(g.ctx, C.c_void_p(side.cuda_stream))
            for h in hooks:
                h()
            g.lib.mhh_set_stream(g.ctx, C.c_void_p(main.cuda_stream))
            self._side_done = torch.cuda.Event(); self._side_done.record(side)
            L.check(g.ctx, g.lib.mhh_pres_set_output_wait(g.ctx, C.c_void_p(self._side_done.cuda_event)), "mhh_pres_set_output_wait")
        else:
            for h in hooks or ():       # independent work a decomposed solver overlaps with its exchanges: here simply run first
                h()
        if self.swpres == "0":
            return False
        if fused:
            L.check(g.ctx, g.lib.mhh_pres_exec_rk(g.ctx, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld),
                                                  L.ptr(f.mt["u"].fld), L.ptr(f.mt["v"].fld), L.ptr(f.mt["w"].fld),
                                                  L.ptr(f.sd["p"].fld), C.c_double(dt), timeloop.rkorder, timeloop.substep,
                                                  C.c_double(timeloop.dt)), "mhh_pres_exec_rk")
            return True
        L.check(g.ctx, g.lib.mhh_pres_exec(g.ctx, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld),
                                           L.ptr(f.mt["u"].fld), L.ptr(f.mt["v"].fld), L.ptr(f.mt["w"].fld),
                                           L.ptr(f.sd["p"].fld), C.c_double(dt)), "mhh_pres_exec")
        return False

    def check_divergence(self):
        g, f = self.grid, self.fields
        out = C.c_double()
        L.check(g.ctx, g.lib.mhh_pres_check_divergence(g.ctx, L.ptr(f.mp["u"].fld), L.ptr(f.mp["v"].fld), L.ptr(f.mp["w"].fld),
                                                       C.byref(out)), "mhh_pres_check_divergence")
        return out.value


class Timeloop:
    """Timeloop (include/timeloop.h; src/timeloop.cxx): integer-nanosecond clock + low-storage RK."""
    ifactor = 1e9

    def __init__(self, grid, fields, rkorder=3, dt=0.1, dtmax=1e9, adaptivestep=True, starttime=0., endtime=1e9, savetime=None):
        if rkorder not in (3, 4):
            raise L.MhhError("rkorder must be 3 or 4")
        self.grid, self.fields = grid, fields
        self.rkorder, self.adaptivestep = rkorder, adaptivestep
        self.dt, self.dtmax = dt, dtmax
        self.substep = 0
        self.itime = int(self.ifactor * starttime + 0.5)
        self.iendtime = int(self.ifactor * endtime + 0.5)
        self.idt = int(self.ifactor * dt + 0.5)
        self.idtmax = int(self.ifactor * dtmax + 0.5)
        self.isavetime = None if savetime is None else int(self.ifactor * savetime + 0.5)
        self.time = starttime
        self.iteration = 0

    def in_substep(self):
        return self.substep > 0

    def get_sub_time_step(self):
        return self.grid.lib.mhh_rk_sub_dt(self.rkorder, self.substep, C.c_double(self.dt))

    def set_time_step_limit(self):
        """src/timeloop.cxx:138-152 (no wall-clock limit; the savetime term only when [time] savetime was given)"""
        self.idtlim = self.idtmax
        if self.isavetime:
            self.idtlim = min(self.idtlim, self.isavetime - self.itime % self.isavetime)
        if self.itime < self.iendtime:
            self.idtlim = min(self.idtlim, self.iendtime - self.itime)

    def set_time_step(self, *limits):
        """src/timeloop.cxx:229-245 + src/model.cxx:731-751"""
        if self.in_substep() or not self.adaptivestep:
            return
        self.set_time_step_limit()
        idtlim = min([self.idtlim, *limits])
        if self.idt == 0:           # src/timeloop.cxx:237-241: the previous step's limit fell below the clock's precision
            raise L.MhhError(f"Required time step less than precision {1. / self.ifactor:f} of the time stepping")
        self.idt = max(int(idtlim), 0)
        self.dt = self.idt / self.ifactor

    def step_time(self):
        """src/timeloop.cxx:160-180: advance the clock after the last substep."""
        if self.in_substep():
            return
        self.itime += self.idt
        self.time = self.itime / self.ifactor
        self.iteration += 1

    def exec(self, skip=()):
        """Timeloop::exec, src/timeloop.cxx:358-411.  `skip`: fields already advanced by the fused pressure tail."""
        g, f = self.grid, self.fields
        names = [n for n in f.at if n not in skip]
        a = L.ptr_array([f.ap[n].fld for n in names])
        at = L.ptr_array([f.at[n].fld for n in names])
        L.check(g.ctx, g.lib.mhh_rk_substep(g.ctx, self.rkorder, self.substep, C.c_double(self.dt), len(names), a, at),
                "mhh_rk_substep")
        self.substep = (self.substep + 1) % (3 if self.rkorder == 3 else 5)


def _level_profile(grid, prof, name):
    """ktot values -> device profile of kcells in the reference layout (std::rotate by kstart, src/buffer.cxx:139-141)."""
    a = np.zeros(grid.kcells, dtype=grid.npdtype)
    p = np.asarray(prof, dtype=grid.npdtype)
    if p.shape != (grid.ktot,):
        raise L.MhhError(f"profile {name} must have ktot={grid.ktot} values")
    a[grid.kstart:grid.kend] = p
    return torch.from_numpy(a).to(grid.device)


class Buffer:
    """Buffer (include/buffer.h, src/buffer.cxx): damping towards a profile above zstart.  [buffer] swbuffer, swupdate,
    zstart, sigma (default 2), beta (default 2); the profiles are the `init` profiles of the case input (ktot values)."""

    def __init__(self, grid, fields, swbuffer=False, swupdate=False, zstart=None, sigma=2., beta=2., profs=None):
        self.grid, self.fields = grid, fields
        self.swbuffer, self.swupdate = bool(swbuffer), bool(swupdate)
        if not self.swbuffer:
            return
        if zstart is None:
            raise L.MhhError("[buffer] zstart is required when swbuffer=1")
        self.zstart, self.sigma, self.beta = float(zstart), float(sigma), float(beta)
        ks, ksh = C.c_int(), C.c_int()
        L.check(grid.ctx, grid.lib.mhh_buffer_plan(grid.ctx, C.c_double(self.zstart), C.c_double(self.sigma),
                                                   C.c_double(self.beta), C.byref(ks), C.byref(ksh)), "mhh_buffer_plan")
        self.bufferkstart, self.bufferkstarth = ks.value, ksh.value
        self.bufferprofs = {}
        if self.swupdate:
            fields.set_calc_mean_profs(True)
        else:
            profs = profs or {}
            for name in fields.ap:      # w (and anything not given) relaxes to zero, src/buffer.cxx:94-96
                self.bufferprofs[name] = (_level_profile(grid, profs[name], name) if name in profs
                                          else torch.zeros(grid.kcells, dtype=grid.dtype, device=grid.device))

    def exec(self, stats=None, skip=()):
        """`skip`: fields whose Buffer term was applied inside the fused scalar pass."""
        if not self.swbuffer:
            return
        g, f = self.grid, self.fields
        for name, fld in f.ap.items():
            if name in skip:
                continue
            abuf = fld.fld_mean if self.swupdate else self.bufferprofs[name]
            L.check(g.ctx, g.lib.mhh_buffer_exec(g.ctx, L.ptr(f.at[name].fld), L.ptr(fld.fld), L.ptr(abuf),
                                                 1 if name == "w" else 0), "mhh_buffer_exec")


class Force:
    """Force (include/force.h, src/force.cxx:586-720): swlspres in {0, uflux, dpdx, geo}, swls, swwls in {0, mean, local}
    (+ swwls_mom), swnudge.  Time-dependent profiles (Timedep) are out of scope: profiles are constant in time."""

    def __init__(self, grid, fields, swlspres="0", uflux=0., dpdx=0., fc=-1., lat=0., swls="0", lslist=(), swwls="0",
                 swwls_mom=False, swnudge="0", nudgelist=(), profs=None, utrans=0., vtrans=0.):
        self.grid, self.fields = grid, fields
        self.swlspres, self.swls, self.swwls, self.swnudge = str(swlspres), str(swls), str(swwls), str(swnudge)
        if self.swwls == "1":
            self.swwls = "local"                       # src/force.cxx:327-331
        if self.swlspres not in ("0", "uflux", "dpdx", "geo"):
            raise L.MhhError('Invalid option for "swlspres"')
        if self.swls not in ("0", "1"):
            raise L.MhhError('Invalid option for "swls"')
        if self.swwls not in ("0", "mean", "local"):
            raise L.MhhError('Invalid option for "swwls"')
        if self.swnudge not in ("0", "1"):
            raise L.MhhError('Invalid option for "swnduge"')
        self.uflux, self.dpdx, self.utrans, self.vtrans = float(uflux), float(dpdx), float(utrans), float(vtrans)
        self.swwls_mom = bool(swwls_mom)
        profs = profs or {}
        if self.swlspres == "geo":
            self.fc = float(fc)
            if self.fc < 0:                            # src/force.cxx:615-617 (Constants::e_rot = 7.2921e-5)
                self.fc = 2. * 7.2921e-5 * np.sin(lat * np.pi / 180.)
            self.ug, self.vg = _level_profile(grid, profs["u_geo"], "u_geo"), _level_profile(grid, profs["v_geo"], "v_geo")
        self.lslist = list(lslist) if self.swls == "1" else []
        for name in self.lslist:
            if name not in fields.ap:
                raise L.MhhError(f"field {name} in [force][lslist] is illegal")
        self.lsprofs = {n: _level_profile(grid, profs[n + "_ls"], n + "_ls") for n in self.lslist}
        if self.swwls != "0":
            self.wls = _level_profile(grid, profs["w_ls"], "w_ls")
            fields.set_calc_mean_profs(True)
        self.nudgelist = list(nudgelist) if self.swnudge == "1" else []
        if self.swnudge == "1":
            for name in self.nudgelist:
                if name not in fields.ap:
                    raise L.MhhError(f"field {name} in [force][nudgelist] is illegal")
            self.nudge_factor = _level_profile(grid, profs["nudgefac"], "nudgefac")
            self.nudgeprofs = {}
            for n in self.nudgelist:
                p = np.asarray(profs[n + "_nudge"], dtype=np.float64)
                p = p - (self.utrans if n == "u" else self.vtrans if n == "v" else 0.)
                self.nudgeprofs[n] = _level_profile(grid, p, n + "_nudge")
            fields.set_calc_mean_profs(True)

    def exec(self, dt, thermo=None, stats=None, skip_coriolis=False):
        """`skip_coriolis`: the swlspres=geo term was applied inside the fused momentum pass (mhh_advec2_smag2_set_force_tail)."""
        g, f, lib = self.grid, self.fields, self.grid.lib
        if self.swlspres == "uflux" and f.comm is not None and f.comm.world > 1:
            # enforce_fixed_flux with the two volume means summed over the ranks (Field3d_operators::calc_mean -> master.sum):
            # mhh_mean divides this rank's sum by the GLOBAL itot*jtot*zsize, so the rank values add up to the mean
            um, utm = C.c_double(), C.c_double()
            L.check(g.ctx, lib.mhh_mean(g.ctx, L.ptr(f.mp["u"].fld), C.byref(um)), "mhh_mean")
            L.check(g.ctx, lib.mhh_mean(g.ctx, L.ptr(f.mt["u"].fld), C.byref(utm)), "mhh_mean")
            um, utm = f.comm.allreduce_sum([um.value, utm.value], g.device).tolist()
            fbody = (self.uflux - um - self.utrans) / dt - utm
            L.check(g.ctx, lib.mhh_force_pressure(g.ctx, L.ptr(f.mt["u"].fld), C.c_double(fbody)), "mhh_force_pressure")
        elif self.swlspres == "uflux":
            L.check(g.ctx, lib.mhh_force_fixed_flux(g.ctx, L.ptr(f.mt["u"].fld), L.ptr(f.mp["u"].fld), C.c_double(self.uflux),
                                                    C.c_double(self.utrans), C.c_double(dt), None), "mhh_force_fixed_flux")
        elif self.swlspres == "dpdx":
            L.check(g.ctx, lib.mhh_force_pressure(g.ctx, L.ptr(f.mt["u"].fld), C.c_double(-1. * self.dpdx)), "mhh_force_pressure")
        elif self.swlspres == "geo" and not skip_coriolis:
            L.check(g.ctx, lib.mhh_force_coriolis(g.ctx, L.ptr(f.mt["u"].fld), L.ptr(f.mt["v"].fld), L.ptr(f.mp["u"].fld),
                                                  L.ptr(f.mp["v"].fld), L.ptr(self.ug), L.ptr(self.vg), C.c_double(self.fc),
                                                  C.c_double(self.utrans), C.c_double(self.vtrans)), "mhh_force_coriolis")
        for name in self.lslist:
            L.check(g.ctx, lib.mhh_force_large_scale_source(g.ctx, L.ptr(f.at[name].fld), L.ptr(self.lsprofs[name])),
                    "mhh_force_large_scale_source")
        if self.swwls != "0":
            mean = self.swwls == "mean"
            targets = (["u", "v"] if self.swwls_mom else []) + list(f.sp)
            for name in targets:
                s = f.ap[name].fld_mean if mean else f.ap[name].fld
                L.check(g.ctx, lib.mhh_force_subsidence(g.ctx, 0 if mean else 1, L.ptr(f.at[name].fld), L.ptr(s), L.ptr(self.wls)),
                        "mhh_force_subsidence")
            if self.swwls_mom and not mean:
                L.check(g.ctx, lib.mhh_force_subsidence(g.ctx, 2, L.ptr(f.mt["w"].fld), L.ptr(f.mp["w"].fld), L.ptr(self.wls)),
                        "mhh_force_subsidence")
        for name in self.nudgelist:
            L.check(g.ctx, lib.mhh_force_nudge(g.ctx, L.ptr(f.at[name].fld), L.ptr(f.ap[name].fld_mean),
                                               L.ptr(self.nudgeprofs[name]), L.ptr(self.nudge_factor)), "mhh_force_nudge")


def _surface_args_type(real):
    """mhh_surface_args / mhh_surface_args_f32 (the three alpha_* scalars have the working type)"""
    class Args(C.Structure):
        _fields_ = _SURFACE_FIELDS[:-3] + [(n, real) for n in ("alpha_m", "alpha_ch", "alpha_h")]
    return Args


_SURFACE_FIELDS = ([("mbcbot", C.c_int), ("thermo", C.c_int), ("thermobc", C.c_int)]
                + [(n, C.c_void_p) for n in ("u", "v", "ubot", "vbot", "ufluxbot", "vfluxbot", "ugradbot", "vgradbot", "th",
                                             "thbot", "thgradbot", "thfluxbot", "ustar", "obuk", "nobuk", "z0m", "z0h",
                                             "dutot", "dudz", "dvdz", "dbdz")]
                + [("iterative_solver", C.c_int), ("charnock", C.c_int), ("alpha_m", C.c_double), ("alpha_ch", C.c_double),
                   ("alpha_h", C.c_double)])


class Boundary_surface(Boundary):
    """Boundary_surface (include/boundary_surface.h, src/boundary_surface.cxx): swboundary=surface, the Monin-Obukhov
    surface layer; mbcbot in {noslip, ustar}.  swconstantz0=true: homogeneous z0m/z0h and the z/L lookup solver;
    swconstantz0=false: z0m / z0h may be 2-D arrays (jcells x icells) and the iterative solver runs; swcharnock=true (with
    swconstantz0=false) updates them from u* every step with alpha_m, alpha_ch, alpha_h (src/boundary_surface.cxx:462-480)."""

    def __init__(self, grid, fields, z0m=0.1, z0h=0.1, ustar=None, swconstantz0=True, swcharnock=False, alpha_m=None,
                 alpha_ch=None, alpha_h=None, **kw):
        mbcbot = kw.pop("mbcbot", "noslip")
        super().__init__(grid, fields, mbcbot=mbcbot, swboundary="surface", **kw)
        if self.mbcbot == L.BC_NEUMANN:
            raise L.MhhError("Neumann bc is not supported in surface model")           # src/boundary_surface.cxx:483-487
        for bc in self.sbc.values():
            if bc.bcbot == L.BC_NEUMANN:
                raise L.MhhError("Fixed Gradient bc is not supported in surface model")
            if bc.bcbot == L.BC_DIRICHLET and self.mbcbot == L.BC_USTAR:
                raise L.MhhError("Fixed Ustar bc in combination with Dirichlet bc for scalars is not supported")
        if self.mbcbot == L.BC_USTAR and ustar is None:
            raise L.MhhError('Item "ustar" in group "boundary" does not exist')
        self.ustarin = ustar
        self._args_t = _surface_args_type(grid.lib.real)
        self.swconstantz0, self.swcharnock = bool(swconstantz0), bool(swcharnock)
        if self.swcharnock and self.swconstantz0:
            raise L.MhhError('"swcharnock=true" requires "swconstantz0=false"')                 # src/boundary_surface.cxx:472-473
        if self.swcharnock and None in (alpha_m, alpha_ch, alpha_h):
            raise L.MhhError('Item "alpha_m" in group "boundary" does not exist')
        self.alpha = tuple(float(a or 0.) for a in (alpha_m, alpha_ch, alpha_h))
        self.thermobc = self.sbc["th"].bcbot if "th" in self.sbc else L.BC_FLUX          # :514-521
        if self.swconstantz0:
            self.z0m_hom, self.z0h_hom = float(z0m), float(z0h)

        def z2(v=0.):
            if np.ndim(v) == 2:
                return torch.from_numpy(np.ascontiguousarray(v, dtype=grid.npdtype)).to(grid.device)
            return torch.full(grid.shape2d, float(v), dtype=grid.dtype, device=grid.device)
        # init_surface, src/boundary_surface.cxx:536-572
        self.obuk, self.ustar = z2(1e-9), z2(1e-2)
        self.dudz_mo, self.dvdz_mo, self.dbdz_mo = z2(), z2(), z2()
        self.z0m, self.z0h = z2(z0m), z2(z0h)
        self.nobuk = torch.zeros(grid.shape2d, dtype=torch.int32, device=grid.device)
        self._dutot = grid.zeros2d()      # exchanged between y-slab ranks: from the grid's arena when there is one

    def set_values(self):
        """Boundary_surface::set_values, src/boundary_surface.cxx:749-779"""
        super().set_values()
        f = self.fields
        for name, val in (("u", self.ubot), ("v", self.vbot)):
            fld = f.mp[name]
            self._set_bc(fld.fld_bot, fld.grad_bot, fld.flux_bot, L.BC_DIRICHLET, val, f.visc, 0.)
        if self.mbcbot == L.BC_USTAR:                      # set_ustar, :782-810
            for name, val in (("u", self.ubot), ("v", self.vbot)):
                fld = f.mp[name]
                self._set_bc(fld.fld_bot, fld.grad_bot, fld.flux_bot, self.mbcbot, val, f.visc, 0.)
            self.ustar.fill_(max(0.0001, self.ustarin))
        g = self.grid
        if self.swconstantz0:       # the z/L lookup table of the homogeneous-z0 solver (prepare_lut)
            L.check(g.ctx, g.lib.mhh_surface_init_solver(g.ctx, self.mbcbot, self.thermobc, C.c_double(self.z0m_hom),
                                                         C.c_double(self.z0h_hom)), "mhh_surface_init_solver")

    def planes(self):
        """The bottom-row inputs of Diff_smag2 (exec_viscosity: z0m, dudz, dvdz, dbdz)."""
        return dict(z0m=self.z0m, dudz=self.dudz_mo, dvdz=self.dvdz_mo, dbdz=self.dbdz_mo)

    def exec(self, thermo=None):
        """Boundary_surface::exec, src/boundary_surface.cxx:827-981."""
        g, f = self.grid, self.fields
        dry = thermo is not None and thermo.get_switch() != "0"
        u, v = f.mp["u"], f.mp["v"]
        a = self._args_t()
        a.mbcbot, a.thermo, a.thermobc = self.mbcbot, 1 if dry else 0, self.thermobc
        p = lambda t: t.data_ptr() if t is not None else None
        a.u, a.v = p(u.fld), p(v.fld)
        a.ubot, a.vbot, a.ufluxbot, a.vfluxbot = p(u.fld_bot), p(v.fld_bot), p(u.flux_bot), p(v.flux_bot)
        a.ugradbot, a.vgradbot = p(u.grad_bot), p(v.grad_bot)
        if dry:
            th = f.sp["th"]
            a.th, a.thbot, a.thgradbot, a.thfluxbot = p(th.fld), p(th.fld_bot), p(th.grad_bot), p(th.flux_bot)
        a.ustar, a.obuk, a.nobuk, a.z0m, a.z0h = p(self.ustar), p(self.obuk), p(self.nobuk), p(self.z0m), p(self.z0h)
        a.dutot, a.dudz, a.dvdz, a.dbdz = p(self._dutot), p(self.dudz_mo), p(self.dvdz_mo), p(self.dbdz_mo)
        a.iterative_solver, a.charnock = 0 if self.swconstantz0 else 1, 1 if self.swcharnock else 0
        a.alpha_m, a.alpha_ch, a.alpha_h = self.alpha
        if g.npy == 1:
            L.check(g.ctx, g.lib.mhh_boundary_surface_exec(g.ctx, C.byref(a)), "mhh_boundary_surface_exec")
        else:
            # y-slab rank: the cyclic fills inside Boundary_surface::exec (src/boundary_surface.cxx:850,930-931) need the
            # neighbours' rows -> the call runs in stages with the ring exchange (self.boundary_cyclic) in between
            st = g.lib.mhh_boundary_surface_exec_stages
            L.check(g.ctx, st(g.ctx, C.byref(a), 1), "mhh_boundary_surface_exec_stages 1")
            if self.swcharnock:
                self.boundary_cyclic.exec_2d(self.z0m)
                self.boundary_cyclic.exec_2d(self.z0h)
            self.boundary_cyclic.exec_2d(self._dutot)
            L.check(g.ctx, st(g.ctx, C.byref(a), 2), "mhh_boundary_surface_exec_stages 2")
            many = getattr(self.boundary_cyclic, "exec_2d_many", None)
            if many is not None:
                many([u.flux_bot, v.flux_bot])           # one ring exchange for both planes on y-slabs
            else:
                self.boundary_cyclic.exec_2d(u.flux_bot)
                self.boundary_cyclic.exec_2d(v.flux_bot)
            L.check(g.ctx, st(g.ctx, C.byref(a), 4), "mhh_boundary_surface_exec_stages 4")
        for name, fld in f.sp.items():
            if dry and name == "th":
                continue
            L.check(g.ctx, g.lib.mhh_surface_scalar(g.ctx, self.sbc[name].bcbot, L.ptr(fld.fld), L.ptr(fld.fld_bot), L.ptr(fld.grad_bot),
                                                    L.ptr(fld.flux_bot), L.ptr(self.ustar), L.ptr(self.obuk), L.ptr(self.z0h)),
                    "mhh_surface_scalar")

    def update_slave_bcs(self):
        pass        # the surface model computes the planes (src/boundary_surface.cxx:984-988)
