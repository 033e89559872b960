"""TEST INFRASTRUCTURE ONLY — the reference's LES time loop for the in-scope terms, driven through
``oracle/_ref`` (the reference's own sources compiled in place).  Mirrors the ordering of
``Model::exec`` (src/model.cxx:336-545): cyclic + ghost BCs, exec_viscosity, set_time_step, thermo,
advec, diff, pres, RK.  Used by tests (multi-step parity) and by bench.py's CPU-baseline legs."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .ref import RefWorld, lib, _p

BC = {"dirichlet": 0, "neumann": 1, "flux": 2, "noslip": 0, "freeslip": 1}


class OracleLES:
    def __init__(self, itot, jtot, ktot, xsize, ysize, zsize, z, fields, *, swadvec="2", swdiff="smag2", order=2,
                 visc=1e-5, svisc=1e-5, cs=0.23, tPr=1. / 3., thref0=300., swthermo="dry", surface=None,
                 mbcbot="freeslip", mbctop="freeslip", sbc=None, rkorder=3, dt=1.0, dtmax=1e9, adaptivestep=True,
                 cflmax=1.0, dnmax=0.4, rhoref=None, rhorefh=None, fluxlimit_list=(), endtime=None, buffer=None, force=None,
                 profs=None, mo=None):
        gc = None
        if swadvec == "2i5":
            gc = (3, 3, 2 if fluxlimit_list else 1)      # src/advec_2i5.cxx:42-45
        self.fluxlimit_list = tuple(fluxlimit_list)
        self.iendtime = None if endtime is None else int(1e9 * endtime + 0.5)
        self.W = RefWorld(itot, jtot, ktot, xsize, ysize, zsize, z, order=order, gc=gc, rhoref=rhoref, rhorefh=rhorefh)
        W = self.W
        self.names = list(fields)                       # e.g. ["u","v","w","th"]
        self.scalars = [n for n in self.names if n not in ("u", "v", "w")]
        self.a = {n: np.ascontiguousarray(fields[n], dtype=np.float64).copy() for n in self.names}
        self.at = {n: np.zeros(W.shape) for n in self.names}
        self.evisc = np.zeros(W.shape)
        self.p = np.zeros(W.shape)
        self.N2 = np.zeros(W.shape)
        self.swadvec, self.swdiff, self.swthermo = swadvec, swdiff, swthermo
        self.visc, self.svisc, self.cs, self.tPr = visc, svisc, cs, tPr
        self.thref = np.full(W.kcells, thref0)
        self.threfh = np.full(W.kcells, thref0)
        self.surface = surface or {}
        self.sm = 1 if self.surface else 0
        self.mbc = (BC[mbcbot], BC[mbctop])
        self.sbc = sbc or {}
        z2 = lambda v=0.: np.full((W.jcells, W.icells), float(v))
        self.planes = {}
        for n in ("u", "v"):
            self.planes[n] = dict(bot=z2(), top=z2(), gradbot=z2(), gradtop=z2())
        for n in self.scalars:
            spec = self.sbc.get(n, {})
            bcb, bct = BC[spec.get("sbcbot", "neumann")], BC[spec.get("sbctop", "neumann")]
            pl = dict(bot=z2(), top=z2(), gradbot=z2(), gradtop=z2(), bcbot=bcb, bctop=bct)
            if bcb == 0: pl["bot"][:] = spec.get("sbot", 0.)
            else: pl["gradbot"][:] = spec.get("sbot", 0.)
            if bct == 0: pl["top"][:] = spec.get("stop", 0.)
            else: pl["gradtop"][:] = spec.get("stop", 0.)
            self.planes[n] = pl
        self.rkorder, self.adaptivestep = rkorder, adaptivestep
        self.cflmax, self.dnmax = cflmax, dnmax
        self.ifactor = 1e9
        self.dt = dt
        self.idt = int(self.ifactor * dt + 0.5)
        self.idtmax = int(self.ifactor * dtmax + 0.5)
        self.itime, self.time, self.iteration, self.substep = 0, 0., 0, 0
        self.zero2 = z2()
        # Monin-Obukhov surface layer (Boundary_surface, swconstantz0=true): mo = dict(z0m=, z0h=); the surface planes are then
        # computed every substep (src/model.cxx:376-377) instead of being prescribed through `surface`.
        self.mo = mo
        if mo:
            self.sm = 1
            self.ustar, self.obuk = z2(1e-2), z2(1e-9)
            self.nobuk = np.zeros((W.jcells, W.icells), dtype=np.int32)
            self.z0m, self.z0h, self.dutot = z2(mo["z0m"]), z2(mo["z0h"]), z2()
            self.surface = dict(z0m=self.z0m, dudz=z2(), dvdz=z2(), dbdz=z2(), ufluxbot=z2(), vfluxbot=z2(), ufluxtop=z2(),
                                vfluxtop=z2(), sfluxbot=z2(), sfluxtop=z2())
            self.ugradbot, self.vgradbot = self.planes["u"]["gradbot"], self.planes["v"]["gradbot"]
            self.thermobc = self.planes["th"]["bcbot"] if "th" in self.planes else 2
            if "th" in self.planes:
                pl = self.planes["th"]
                if pl["bcbot"] == 2:          # flux bc: sbot is the prescribed surface flux (Boundary::set_values)
                    self.surface["sfluxbot"][:] = pl["gradbot"]; pl["gradbot"][:] = 0.
                if pl["bctop"] != 0:
                    self.surface["sfluxtop"][:] = -svisc * pl["gradtop"]
            lib().ref_surface_lut(W.h, C.c_int(self.mbc[0]), C.c_int(self.thermobc), C.c_double(mo["z0m"]), C.c_double(mo["z0h"]), None, None)
        # Buffer / Force (src/buffer.cxx, src/force.cxx): dicts with the .ini keys; profs = init profiles (ktot values)
        self.buffer, self.force = buffer or {}, force or {}
        def lev(name):
            p = np.zeros(W.kcells)
            if profs and name in profs:
                p[W.kstart:W.kend] = profs[name]
            return p
        self.bufferprofs = {n: lev(n) for n in self.names}
        self.ug, self.vg = lev("u_geo"), lev("v_geo")

    def close(self):
        self.W.close()

    # -- Model::exec body for one substep ----------------------------------------------------
    def exec_substep(self):
        W, a, at = self.W, self.a, self.at
        u, v, w = a["u"], a["v"], a["w"]
        for n in self.names:
            W.call("ref_cyclic", a[n], 2)
        for n in ("u", "v"):
            pl = self.planes[n]
            W.call("ref_ghost_cells", a[n], self.mbc[0], self.mbc[1], pl["bot"], pl["top"], pl["gradbot"], pl["gradtop"])
        if W.order == 4:
            W.call("ref_ghost_cells_w", w, 0)
        for n in self.scalars:
            pl = self.planes[n]
            W.call("ref_ghost_cells", a[n], pl["bcbot"], pl["bctop"], pl["bot"], pl["top"], pl["gradbot"], pl["gradtop"])
        S = self.surface
        if self.swdiff == "smag2":
            W.call("ref_smag2_strain2", self.sm, self.evisc, u, v, w, S.get("dudz"), S.get("dvdz"))
            if self.swthermo == "dry":
                W.call("ref_thermo_N2", self.N2, a["th"], self.thref)
                W.call("ref_smag2_evisc", self.sm, self.evisc, u, v, w, self.N2, S.get("dbdz"), S.get("z0m"), self.cs, self.tPr)
            else:
                W.call("ref_smag2_evisc_neutral", self.sm, self.evisc, u, v, w, S.get("z0m"), self.cs, self.visc)
        if self.substep == 0 and self.adaptivestep:
            cfl = max(1e-5, self.get_cfl(self.dt))
            lim = [self.idtmax, int(self.idt * self.cflmax / cfl)]
            if self.swdiff == "smag2":
                dnmul = max(1e-9, W.call("ref_smag2_dnmul", self.evisc, self.tPr))
                lim.append(int(self.idt * self.dnmax / (self.dt * dnmul)))
            elif self.swdiff in ("2", "4"):
                lim.append(int(self.idt * self.dnmax / (self.dt * self.static_dnmul())))
            if self.iendtime is not None:
                lim.append(self.iendtime - self.itime)
            self.idt = max(min(lim), 1)
            self.dt = self.idt / self.ifactor
        if self.swthermo == "dry":
            W.call("ref_thermo_buoyancy_tend", at["w"], a["th"], self.threfh)
        if self.mo:        # boundary->exec + set_ghost_cells, src/model.cxx:376-377
            S = self.surface
            dry = self.swthermo == "dry"
            pth = self.planes.get("th", {})
            lib().ref_surface_exec(
                W.h, C.c_int(self.mbc[0]), C.c_int(1 if dry else 0), C.c_int(self.thermobc), _p(u), _p(v),
                _p(self.planes["u"]["bot"]), _p(self.planes["v"]["bot"]), _p(S["ufluxbot"]), _p(S["vfluxbot"]),
                _p(self.ugradbot), _p(self.vgradbot), _p(a["th"]) if dry else None, _p(pth.get("bot")), _p(pth.get("gradbot")),
                _p(S["sfluxbot"]), _p(self.thref), _p(self.threfh), _p(self.ustar), _p(self.obuk),
                self.nobuk.ctypes.data_as(C.c_void_p), _p(self.z0m), _p(self.z0h), _p(self.dutot), _p(S["dudz"]), _p(S["dvdz"]),
                _p(S["dbdz"]))
            for n in ("u", "v"):
                pl = self.planes[n]
                W.call("ref_ghost_cells", a[n], self.mbc[0], self.mbc[1], pl["bot"], pl["top"], pl["gradbot"], pl["gradtop"])
            for n in self.scalars:
                pl = self.planes[n]
                W.call("ref_ghost_cells", a[n], pl["bcbot"], pl["bctop"], pl["bot"], pl["top"], pl["gradbot"], pl["gradtop"])
        if W.order == 4:
            W.call("ref_ghost_cells_w", w, 1)
        if self.swadvec == "2":
            W.call("ref_advec2_uvw", at["u"], at["v"], at["w"], u, v, w)
            for n in self.scalars:
                W.call("ref_advec2_s", at[n], a[n], u, v, w)
        elif self.swadvec == "2i5":
            W.call("ref_advec2i5_uvw", at["u"], at["v"], at["w"], u, v, w)
            for n in self.scalars:
                W.call("ref_advec2i5_s", at[n], a[n], u, v, w, 1 if n in self.fluxlimit_list else 0)
        elif self.swadvec == "4":
            W.call("ref_advec4_uvw", at["u"], at["v"], at["w"], u, v, w)
            for n in self.scalars:
                W.call("ref_advec4_s", at[n], a[n], u, v, w)
        if W.order == 4:
            W.call("ref_ghost_cells_w", w, 0)
        z2 = self.zero2
        if self.swdiff == "smag2":
            W.call("ref_smag2_diff_uvw", self.sm, at["u"], at["v"], at["w"], u, v, w, self.evisc,
                   S.get("ufluxbot", z2), S.get("ufluxtop", z2), S.get("vfluxbot", z2), S.get("vfluxtop", z2), self.visc)
            for n in self.scalars:
                W.call("ref_smag2_diff_c", self.sm, at[n], a[n], self.evisc, S.get("sfluxbot", z2), S.get("sfluxtop", z2),
                       self.tPr, self.svisc)
        elif self.swdiff == "2":
            W.call("ref_diff2_c", at["u"], u, self.visc)
            W.call("ref_diff2_c", at["v"], v, self.visc)
            W.call("ref_diff2_w", at["w"], w, self.visc)
            for n in self.scalars:
                W.call("ref_diff2_c", at[n], a[n], self.svisc)
        elif self.swdiff == "4":
            W.call("ref_diff4_c", at["u"], u, self.visc)
            W.call("ref_diff4_c", at["v"], v, self.visc)
            W.call("ref_diff4_w", at["w"], w, self.visc)
            for n in self.scalars:
                W.call("ref_diff4_c", at[n], a[n], self.svisc)
        sub_dt = lib().ref_rk_subdt(C.c_int(self.rkorder), C.c_double(self.dt), C.c_int(self.substep))
        if self.buffer.get("swbuffer"):                 # Buffer::exec, src/model.cxx:395
            b = self.buffer
            for n in self.names:
                W.call("ref_buffer", at[n], a[n], self.bufferprofs[n], float(b["zstart"]), float(b.get("sigma", 2.)),
                       float(b.get("beta", 2.)), 1 if n == "w" else 0)
        sw = self.force.get("swlspres", "0")           # Force::exec, src/model.cxx:404
        if sw == "geo":
            W.call("ref_force_coriolis", at["u"], at["v"], u, v, self.ug, self.vg, float(self.force["fc"]), 0., 0.)
        elif sw == "uflux":
            W.call("ref_force_fixed_flux", at["u"], u, float(self.force["uflux"]), 0., float(sub_dt))
        elif sw == "dpdx":
            W.call("ref_force_pressure", at["u"], -1. * float(self.force["dpdx"]))
        if W.order == 4:
            W.call("ref_ghost_cells_w", w, 1)
            W.call("ref_pres4_exec", self.p, u, v, w, at["u"], at["v"], at["w"], sub_dt)
            W.call("ref_ghost_cells_w", w, 0)
        else:
            W.call("ref_pres2_exec", self.p, u, v, w, at["u"], at["v"], at["w"], sub_dt)
        for n in self.names:
            W.call("ref_rk", self.rkorder, a[n], at[n], self.substep, self.dt)
        self.substep = (self.substep + 1) % (3 if self.rkorder == 3 else 5)
        if self.substep == 0:
            self.itime += self.idt
            self.time = self.itime / self.ifactor
            self.iteration += 1

    def static_dnmul(self):
        """Diff_2/4::create, src/diff_2.cxx:139-151 / src/diff_4.cxx:236-246"""
        W = self.W
        viscmax = max([self.visc] + [self.svisc for _ in self.scalars])
        return max(abs(viscmax * (1. / (W.dx * W.dx) + 1. / (W.dy * W.dy) + 1. / (W.dz[k] * W.dz[k]))) for k in range(W.kstart, W.kend))

    def step(self):
        for _ in range(3 if self.rkorder == 3 else 5):
            self.exec_substep()

    def get_cfl(self, dt):
        name = {"2": "ref_advec2_cfl", "2i5": "ref_advec2i5_cfl", "4": "ref_advec4_cfl"}[self.swadvec]
        return self.W.call(name, self.a["u"], self.a["v"], self.a["w"], dt)

    def divergence(self):
        W = self.W
        f = [self.a[n].copy() for n in "uvw"]
        for x in f:
            W.call("ref_cyclic", x, 2)
        if W.order == 4:
            W.call("ref_ghost_cells_w", f[2], 1)      # Conservation, src/model.cxx:815
        return W.call("ref_pres4_divergence" if W.order == 4 else "ref_pres2_divergence", *f)
