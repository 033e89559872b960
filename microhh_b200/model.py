"""Thermo (dry) and the in-scope part of Model::exec's time loop, on top of ``microhh_b200.host``."""
from __future__ import annotations

import ctypes as C

import os

import numpy as np

from . import lib as L
from .host import Advec, Boundary, Boundary_surface, Buffer, Diff, Fields, Force, Grid, Pres, Timeloop


class Thermo:
    """Thermo_dry pieces inside the substep (src/thermo_dry.cxx): buoyancy tendency and N2.
    swthermo in {"0", "dry"}; Boussinesq base state thref = threfh = thref0 (swbasestate=boussinesq)."""

    def __init__(self, grid, fields, swthermo="dry", thref0=300., thref=None, threfh=None):
        self.grid, self.fields, self.swthermo = grid, fields, swthermo
        self.fuse_N2 = swthermo == "dry"      # Diff.exec_viscosity evaluates N2 in-kernel (mhh_smag2_exec_viscosity_dry)
        if swthermo not in ("0", "dry"):
            raise L.MhhError(f"{swthermo} is an illegal value for swthermo (this path supports 0 and dry)")
        if swthermo == "dry":
            self.thref = np.ascontiguousarray(np.full(grid.kcells, thref0) if thref is None else thref, dtype=grid.npdtype)
            self.threfh = np.ascontiguousarray(np.full(grid.kcells, thref0) if threfh is None else threfh, dtype=grid.npdtype)
            L.check(grid.ctx, grid.lib.mhh_thermo_set_ref(grid.ctx, self.thref.ctypes.data_as(C.c_void_p),
                                                          self.threfh.ctypes.data_as(C.c_void_p)), "mhh_thermo_set_ref")
            self._n2 = grid.zeros()

    def get_switch(self):
        return self.swthermo

    def exec(self, dt=None, stats=None):
        if self.swthermo == "0":
            return
        g, f = self.grid, self.fields
        L.check(g.ctx, g.lib.mhh_thermo_dry_exec(g.ctx, L.ptr(f.mt["w"].fld), L.ptr(f.sp["th"].fld)), "mhh_thermo_dry_exec")

    def get_thermo_field(self, name):
        if name != "N2":
            raise L.MhhError("only N2 is provided on this path")
        g, f = self.grid, self.fields
        L.check(g.ctx, g.lib.mhh_thermo_dry_N2(g.ctx, L.ptr(f.sp["th"].fld), L.ptr(self._n2)), "mhh_thermo_dry_N2")
        return self._n2


class Model:
    """The in-scope part of Model::exec's time loop (src/model.cxx:336-545), one call per RK substep.

    ``ini`` is a dict of dicts with the reference's .ini blocks/keys ([grid], [advec], [diff], [pres],
    [boundary], [fields], [time], [thermo]); the operators are selected through the same keys
    (swadvec=2/2i5/4, swdiff=smag2/2/4, swpres=2/4)."""

    def __init__(self, ini, z, device=None, surface=None, npy=1, rank_y=0, profs=None, mo=False, peer_arena=False,
                 precision="f64"):
        gi = ini["grid"]
        order = str(gi.get("swspatialorder", "2"))
        swadvec = str(ini.get("advec", {}).get("swadvec", order))
        gc = tuple(gi.get("gc", (0, 0, 0)))        # optional extra ghost cells (the library takes the scheme's minimum at least)
        if swadvec == "2i5":
            gc = (3, 3, 2 if ini.get("advec", {}).get("fluxlimit_list") else 1)   # src/advec_2i5.cxx:42-45
        self.grid = Grid(gi["itot"], gi["jtot"], gi["ktot"], gi["xsize"], gi["ysize"], gi["zsize"], z,
                         swspatialorder=order, gc=gc, device=device, npy=npy, rank_y=rank_y, peer_arena=peer_arena,
                         precision=precision)
        fi = ini.get("fields", {})
        slist = list(fi.get("slist", []))
        th = ini.get("thermo", {})
        swthermo = str(th.get("swthermo", "0"))
        if swthermo == "dry" and "th" not in slist:
            slist = ["th"] + slist
        self.fields = Fields(self.grid, slist=slist, visc=fi.get("visc", 0.), svisc=fi.get("svisc", 0.))
        bi = ini.get("boundary", {})
        bkw = dict(mbcbot=bi.get("mbcbot", "noslip"), mbctop=bi.get("mbctop", "noslip"), ubot=bi.get("ubot", 0.), utop=bi.get("utop", 0.),
                   vbot=bi.get("vbot", 0.), vtop=bi.get("vtop", 0.), sbc=bi.get("sbc"))
        # swboundary=surface: with mo=True the Monin-Obukhov surface layer runs on the device every substep
        # (Boundary_surface); otherwise the caller prescribes the surface planes (z0m, dudz, dvdz, dbdz, fluxes) via `surface`.
        self.mo = bool(mo) and bi.get("swboundary", "default") == "surface"
        if self.mo:
            self.boundary = Boundary_surface(self.grid, self.fields, z0m=bi.get("z0m", 0.1), z0h=bi.get("z0h", 0.1),
                                             ustar=bi.get("ustar"), swconstantz0=bi.get("swconstantz0", True),
                                             swcharnock=bi.get("swcharnock", False), alpha_m=bi.get("alpha_m"),
                                             alpha_ch=bi.get("alpha_ch"), alpha_h=bi.get("alpha_h"), **bkw)
        else:
            self.boundary = Boundary(self.grid, self.fields, swboundary=bi.get("swboundary", "default"), **bkw)
        ai = ini.get("advec", {})
        self.advec = Advec.factory(self.grid, self.fields, swadvec, cflmax=ai.get("cflmax", 1.0),
                                   fluxlimit_list=ai.get("fluxlimit_list", ()))
        di = ini.get("diff", {})
        self.diff = Diff.factory(self.grid, self.fields, self.boundary, str(di.get("swdiff", order)), dnmax=di.get("dnmax", 0.4),
                                 cs=di.get("cs", 0.23), tPr=di.get("tPr", 1. / 3.))
        self.pres = Pres.factory(self.grid, self.fields, str(ini.get("pres", {}).get("swpres", order)))
        self.thermo = Thermo(self.grid, self.fields, swthermo, thref0=th.get("thref0", 300.))
        # Thermo_dry + Advec_2 + Diff_smag2 tendencies in one pass per field (mhh_advec2_smag2_exec)
        self.fuse_tend = self.advec.swadvec == "2" and self.diff.swdiff == "smag2" and swthermo in ("0", "dry")
        self.fuse_tend4 = self.advec.swadvec == "4" and self.diff.swdiff == "4"     # mhh_advec4_diff4_exec
        # Advec_2::get_time_limit + Diff_smag2::get_time_limit inside the evisc pass (mhh_smag2_exec_viscosity_dry_limits)
        self.fuse_limits = self.advec.swadvec == "2" and self.diff.swdiff == "smag2" and swthermo == "dry"
        self.fuse_pres_rk = self.pres.swpres in ("2", "4")   # Pres::output + Timeloop::exec(u,v,w) in one kernel (mhh_pres_exec_rk)
        ti = ini.get("time", {})
        self.timeloop = Timeloop(self.grid, self.fields, rkorder=ti.get("rkorder", 3), dt=ti.get("dt", 0.1),
                                 dtmax=ti.get("dtmax", 1e9), adaptivestep=ti.get("adaptivestep", True),
                                 starttime=ti.get("starttime", 0.), endtime=ti.get("endtime", 1e9),
                                 savetime=ti.get("savetime"))
        # [buffer] / [force]; `profs` = the case's init profiles (ktot values each: u, v, th, u_geo, ...)
        bu, fo = ini.get("buffer", {}), ini.get("force", {})
        self.buffer = Buffer(self.grid, self.fields, swbuffer=bu.get("swbuffer", False), swupdate=bu.get("swupdate", False),
                             zstart=bu.get("zstart"), sigma=bu.get("sigma", 2.), beta=bu.get("beta", 2.), profs=profs)
        self.force = Force(self.grid, self.fields, swlspres=fo.get("swlspres", "0"), uflux=fo.get("uflux", 0.),
                           dpdx=fo.get("dpdx", 0.), fc=fo.get("fc", -1.), lat=gi.get("lat", 0.), swls=fo.get("swls", "0"),
                           lslist=fo.get("lslist", ()), swwls=fo.get("swwls", "0"), swwls_mom=fo.get("swwls_mom", False),
                           swnudge=fo.get("swnudge", "0"), nudgelist=fo.get("nudgelist", ()), profs=profs,
                           utrans=gi.get("utrans", 0.), vtrans=gi.get("vtrans", 0.))
        # the scalars' Buffer term and RK update ride in their tendency pass when no [force] term touches a scalar tendency
        fo_ = self.force
        scal = set(self.fields.sp)
        self.fuse_scalar_rk = (self.fuse_tend and bool(scal) and fo_.swwls == "0"
                               and not (scal & set(fo_.lslist)) and not (scal & set(fo_.nudgelist)))
        # single GPU: the scalars' fused pass on a second stream beside the pressure solver (Pres.exec); opt-in, MHH_SCALAR_OVERLAP=1: measured 40.5 / 40.8 ms against 40.6 / 40.4 without (r02ab) — both sides fill the GPU, the block scheduler runs them one after the other
        self.scalar_overlap = self.grid.npy == 1 and os.environ.get("MHH_SCALAR_OVERLAP", "0") != "0"
        # Tendencies left undefined (not zeroed) by the fused last-substep tails; the next fused tendency pass starts from 0
        self.lazy_tend = bool(self.fuse_tend and self.fuse_pres_rk and self.pres.swpres == "2"
                              and (not scal or self.fuse_scalar_rk) and ini.get("b200", {}).get("lazy_tendencies", True))
        if self.lazy_tend:
            L.check(self.grid.ctx, self.grid.lib.mhh_set_lazy_tendencies(self.grid.ctx, 1), "mhh_set_lazy_tendencies")
        # Buffer::exec and the Coriolis branch of Force::exec for u,v,w inside the fused momentum pass (same order of terms)
        bf, fo_ = self.buffer, self.force
        self.fuse_force = bool(self.fuse_tend and (bf.swbuffer or fo_.swlspres == "geo")
                               and ini.get("b200", {}).get("fuse_force", True))
        if self.fuse_force:
            prof = lambda n: (self.fields.mp[n].fld_mean if bf.swupdate else bf.bufferprofs[n]) if bf.swbuffer else None
            geo = fo_.swlspres == "geo"
            L.check(self.grid.ctx, self.grid.lib.mhh_advec2_smag2_set_force_tail(
                self.grid.ctx, L.ptr(prof("u")), L.ptr(prof("v")), L.ptr(prof("w")), 1 if geo else 0,
                C.c_double(fo_.fc if geo else 0.), L.ptr(fo_.ug if geo else None), L.ptr(fo_.vg if geo else None),
                C.c_double(fo_.utrans), C.c_double(fo_.vtrans)), "mhh_advec2_smag2_set_force_tail")
        self.surface = self.boundary.planes() if self.mo else (surface or {})
        self.boundary.set_values()
        self.pres.set_values()
        self.diff.create()

    # ---- hooks a decomposed model overrides (multigpu.SlabModel): halo fills and cross-rank reductions -------------------
    def _cyclic_prognostic(self):
        self.boundary.set_prognostic_cyclic_bcs()

    def _after_viscosity(self):
        pass

    def _reduce_max(self, values):
        return values

    def _scalar_halves(self):
        """True: the scalars' fused tendency + RK passes are issued in two k halves from inside Pres::exec (hooks), where a
        decomposed solver overlaps them with its exchanges (multigpu.SlabModel)."""
        return False

    def exec_substep(self):
        b, tl = self.boundary, self.timeloop
        self._cyclic_prognostic()                          # src/model.cxx:346
        b.set_ghost_cells()                                # :348
        self.fields.exec()                                 # :351 (mean profiles, only when buffer/force asked for them)
        set_dt = not tl.in_substep() and tl.adaptivestep
        # :354; on the first substep of the 2nd-order LES path the two dt maxima ride in the same pass over u,v,w
        lim = self.diff.exec_viscosity(self.thermo, surface=self.surface, limits=set_dt and self.fuse_limits)
        self._after_viscosity()
        if set_dt:                                         # set_time_step, :357 -> :731-751
            if lim is not None:
                cfl_dt, dnmul = self._reduce_max(lim)
                tl.set_time_step(self.advec.time_limit_from(cfl_dt * tl.dt, tl.idt), self.diff.time_limit_from(dnmul, tl.idt, tl.dt))
            else:
                cfl, = self._reduce_max([self.advec.get_cfl(tl.dt)])
                lims = [self.advec.time_limit_from(cfl, tl.idt)]
                if self.diff.swdiff != "0":
                    dnmul, = self._reduce_max([self.diff._dnmul()])
                    lims.append(self.diff.time_limit_from(dnmul, tl.idt, tl.dt))
                tl.set_time_step(*lims)
        if self.mo:             # :376-377 (Thermo::exec of :366 only adds to wt, which the surface layer does not read)
            b.exec(self.thermo)
            b.set_ghost_cells()
        hooks = None
        if self.fuse_tend:      # :366 + :388 + :392, one pass per prognostic field (set_ghost_cells_w is a 4th-order no-op)
            srk = self.fuse_scalar_rk
            th_b = self.fields.sp["th"].fld if self.thermo.swthermo == "dry" else None
            part = lambda *p: (lambda: self.diff.exec_with_advec2(th_b, buffer=self.buffer, timeloop=tl, parts=p))
            if srk and self._scalar_halves():
                nch = (self.grid.kmax + 63) // 64
                half = max(1, nch // 2)               # chunks of the first half; a second part starting at nch is a no-op
                part(1, 0, 0, 0)()                                     # the momentum pass now
                hooks = (part(0, 1, 0, half), part(0, 1, half, 0))     # the scalar passes from inside the pressure solver
            elif srk and self.scalar_overlap and self.fuse_pres_rk and self.fields.sp:
                part(1, 0, 0, 0)()
                hooks = (part(0, 1, 0, 0),)            # the scalar passes on a second stream beside the pressure solver (Pres.exec)
            else:
                self.diff.exec_with_advec2(th_b, buffer=self.buffer if srk else None, timeloop=tl if srk else None)
        elif self.fuse_tend4:   # :387-392 in one call, the momentum diffusion inside the advection passes
            self.thermo.exec()                             # :366
            b.set_ghost_cells_w(L.BCW_CONSERVATION)        # :387
            self.diff.exec_with_advec4()                   # :388-392
        else:
            self.thermo.exec()                             # :366
            b.set_ghost_cells_w(L.BCW_CONSERVATION)        # :387
            self.advec.exec()                              # :388
            b.set_ghost_cells_w(L.BCW_NORMAL)              # :389
            self.diff.exec()                               # :392
        done_s = tuple(self.fields.sp) if (self.fuse_tend and self.fuse_scalar_rk) else ()
        done_m = ("u", "v", "w") if self.fuse_force else ()     # their sponge / Coriolis terms rode in the momentum pass
        self.buffer.exec(skip=done_s + done_m)             # :395
        self.force.exec(tl.get_sub_time_step(), self.thermo, skip_coriolis=self.fuse_force)   # :404
        b.set_ghost_cells_w(L.BCW_CONSERVATION)            # :410
        # :411 + :484 for u,v,w in one pass
        fused = self.pres.exec(tl.get_sub_time_step(), timeloop=tl if self.fuse_pres_rk else None, hooks=hooks)
        if hooks:
            self.diff.swap_scalars()
        if not fused:
            b.set_ghost_cells_w(L.BCW_NORMAL)              # :412 (inside the fused call otherwise)
        tl.exec(skip=(("u", "v", "w") if fused else ()) + done_s)     # :484
        tl.step_time()                                     # :487

    # ---- the reference's restart files (src/model.cxx:257-262,470-478; formats in microhh_b200.io) ----
    def _gridshape(self):
        g = self.grid
        return dict(itot=g.itot, jtot=g.jtot, ktot=g.ktot, igc=g.igc, jgc=g.jgc, kgc=g.kgc)

    def save(self, directory=".", iotime=None):
        """Fields::save + Timeloop::save: <name>.<iotime:07d> for every prognostic field (u, v with the Galilean offset
        added back, src/fields.cxx) and time.<iotime:07d>."""
        from . import io as IO
        tl = self.timeloop
        iotime = int(tl.time) if iotime is None else iotime
        for name, fld in self.fields.ap.items():
            off = self.force.utrans if name == "u" else self.force.vtrans if name == "v" else 0.
            IO.save_field3d(fld.fld.cpu().numpy(), self._gridshape(), name, iotime, offset=off, directory=directory)
        IO.save_time(iotime, tl.itime, tl.idt, tl.iteration, directory=directory)

    def load(self, directory=".", iotime=0):
        """Fields::load + Timeloop::load (the files `microhh init` or a previous run wrote)."""
        import torch
        from . import io as IO
        tl = self.timeloop
        for name, fld in self.fields.ap.items():
            off = self.force.utrans if name == "u" else self.force.vtrans if name == "v" else 0.
            a = IO.load_field3d(self._gridshape(), name, iotime, offset=off, directory=directory)
            fld.fld.copy_(torch.from_numpy(a))
        tl.itime, tl.idt, tl.iteration = IO.load_time(iotime, directory=directory)
        tl.time, tl.dt = tl.itime / tl.ifactor, tl.idt / tl.ifactor

    @classmethod
    def from_case(cls, ini_path, directory=None, device=None, surface=None, profs=None, iotime=0):
        """`microhh run <case>` start-up: read <case>.ini, grid.0000000 and the restart files of `iotime`."""
        import os
        from . import io as IO
        directory = os.path.dirname(os.path.abspath(ini_path)) if directory is None else directory
        inp = IO.Input(ini_path)
        ini = IO.model_ini(inp)
        unused = inp.unused_items()             # Input::print_unused_items (src/input.cxx): keys nothing on this path reads
        if unused:
            import sys
            print("microhh_b200: input items not used by the in-scope dynamical core: " + ", ".join(unused), file=sys.stderr)
        gi = ini["grid"]
        z = IO.load_grid(gi["itot"], gi["jtot"], gi["ktot"], directory=directory)["z"]
        m = cls(ini, z, device=device, surface=surface, profs=profs, mo=surface is None)
        m.load(directory=directory, iotime=iotime)
        return m

    def step(self):
        """One full time step = 3 (RK3) or 5 (RK4) substeps."""
        for _ in range(3 if self.timeloop.rkorder == 3 else 5):
            self.exec_substep()

    def print_status(self):
        """The observables of the <case>.out line (src/model.cxx:801-828)."""
        tl = self.timeloop
        for fld in self.fields.mp.values():      # halos are refreshed before the status line (src/model.cxx:346,360)
            self.boundary.boundary_cyclic.exec(fld.fld)
        self.boundary.set_ghost_cells_w(L.BCW_CONSERVATION)     # src/model.cxx:815-817
        div = self.pres.check_divergence()
        self.boundary.set_ghost_cells_w(L.BCW_NORMAL)
        return dict(iter=tl.iteration, time=tl.time, dt=tl.dt, cfl=self.advec.get_cfl(tl.dt), dn=self.diff.get_dn(tl.dt),
                    div=div, mom=self.fields.check_momentum(), tke=self.fields.check_tke(),
                    mass=self.fields.check_mass())


class HostFeed:
    """Double-buffered host -> device feed of the prognostic fields for a host model that keeps its state in (pinned) host
    memory: the upload of step n+1 runs on a copy stream while step n computes.  submit() enqueues an upload into the free
    staging slot, adopt() makes the compute stream wait for the oldest submitted upload and copies it into the model's
    fields (device to device)."""

    def __init__(self, model):
        import torch
        self.model, self.torch = model, torch
        f = model.fields
        self.stage = [{n: torch.empty_like(fld.fld) for n, fld in f.ap.items()} for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=model.grid.device)
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]      # upload into slot finished
        self.free = [torch.cuda.Event(), torch.cuda.Event()]       # slot has been consumed by adopt()
        self.nsub = self.nadopt = 0

    def submit(self, pinned):
        torch = self.torch
        slot = self.nsub % 2
        with torch.cuda.stream(self.copy_stream):
            if self.nsub >= 2:
                self.copy_stream.wait_event(self.free[slot])
            for n, src in pinned.items():
                self.stage[slot][n].copy_(src, non_blocking=True)
            self.ready[slot].record(self.copy_stream)
        self.nsub += 1

    def adopt(self):
        torch = self.torch
        slot = self.nadopt % 2
        cur = torch.cuda.current_stream(self.model.grid.device)
        cur.wait_event(self.ready[slot])
        for n, fld in self.model.fields.ap.items():       # looked up now: the fused scalar pass swaps fld / fld_alt
            fld.fld.copy_(self.stage[slot][n], non_blocking=True)
        self.free[slot].record(cur)
        self.nadopt += 1
