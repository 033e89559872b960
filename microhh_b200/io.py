"""The reference's on-disk formats either side of the dynamical core, host side only.

* ``Input``: the ``<case>.ini`` reader — blocks, items, ``item[subitem]`` options, ``#`` comments, typed ``get_item`` /
  ``get_list`` with defaults and the reference's error texts (src/input.cxx:40-330, include/input_tools.h:42-84).
* raw binary files of the serial build: ``grid.0000000`` (src/grid_serial.cxx:46-111), ``<field>.<time:07d>`` = the
  ghost-free interior, i fastest (src/field3d_io.cxx:671-751), ``time.<time:07d>`` = ``unsigned long itime, idt; int
  iteration`` (src/timeloop.cxx:445-525).
* ``OutWriter``: the ``<case>.out`` status table (src/model.cxx:795-830), same header and printf formats.
* ``model_ini``: the typed dict-of-dicts ``microhh_b200.model.Model`` takes, built with the reference's keys and defaults.

Nothing here touches the GPU; the arrays are numpy.  NetCDF input (``<case>_input.nc``) is not read — there is no NetCDF
in this image — so initial profiles enter as arrays.
"""
from __future__ import annotations

import os
import struct

import numpy as np

_NODEFAULT = object()
DBIG = 1e9          # Constants::dbig, include/constants.h


class InputError(RuntimeError):
    """std::runtime_error thrown by Input (src/input.cxx)."""


def _check_item(s: str):
    """check_item<std::string>, include/input_tools.h:36-41."""
    if not s:
        raise InputError("Illegal string")


def _convert(value: str, typ):
    """get_item_from_stream<T>, include/input_tools.h:42-84."""
    v = value.strip()
    if typ is str:
        return v
    if typ is bool:
        low = v.lower()
        if low in ("1", "true"):
            return True
        if low in ("0", "false"):
            return False
        raise InputError("Item does not match type")
    parts = v.split()
    if len(parts) != 1:
        raise InputError("Item does not match type")
    try:
        if typ is int:
            return int(parts[0], 10)
        return float(parts[0])
    except ValueError:
        raise InputError("Item does not match type") from None


class Input:
    """Input (include/input.h, src/input.cxx:76-177): itemlist[block][item][subitem] = [value, used]."""

    def __init__(self, file_name=None, text=None):
        if text is None:
            try:
                with open(file_name) as fh:
                    text = fh.read()
            except OSError:
                raise InputError(f'"{file_name}" cannot be opened ') from None
        self.itemlist = {}
        block = ""
        for raw in text.splitlines():
            line = raw.split("#")[0].strip()
            parts = line.split("=")
            if len(parts) == 1:
                header = parts[0].strip()
                if not header:
                    continue
                if header[0] == "[" and header[-1] == "]":
                    block = header[1:-1]
                    _check_item(block)
                    self.itemlist.setdefault(block, {})
                else:
                    raise InputError(f'Line "{line}" is illegal')
            elif len(parts) == 2:
                if not block:
                    raise InputError("No block name found")
                left, right = parts[0].strip(), parts[1].strip()
                sub = ""
                o, c = left.find("["), left.rfind("]")
                if o != -1 and c != -1 and o < c:
                    sub = left[o + 1:c]
                    _check_item(sub)
                    left = left[:o]
                _check_item(left)
                self.itemlist[block].setdefault(left, {})[sub] = [right, False]
            else:
                raise InputError(f'Line "{line}" is illegal')

    def _get_string(self, block, item, sub):
        if block not in self.itemlist:
            raise InputError(f'Block "{block}" does not exist')
        items = self.itemlist[block]
        if item not in items:
            raise InputError(f'Item "{item}" in group "{block}" does not exist')
        subs = items[item]
        if sub not in subs:
            if "" not in subs:
                raise InputError(f'Subitem "{sub}" in item "{item}" does not exist')
            sub = ""
        subs[sub][1] = True
        return subs[sub][0]

    def get_item(self, typ, block, item, sub="", default=_NODEFAULT):
        """get_item<T>(block, item, subitem[, default]); an item that exists but does not parse is an error either way."""
        try:
            value = self._get_string(block, item, sub)
        except InputError:
            if default is _NODEFAULT:
                raise
            return default
        try:
            return _convert(value, typ)
        except InputError:
            out = f"[{block}][{item}]" + (f"[{sub}]" if sub else "")
            if default is _NODEFAULT:
                raise
            raise InputError(f"{out} = {value} is illegal.") from None

    def get_list(self, typ, block, item, sub="", default=_NODEFAULT):
        try:
            value = self._get_string(block, item, sub)
        except InputError:
            if default is _NODEFAULT:
                raise
            return list(default)
        out = [_convert(v, typ) for v in value.split(",")]
        if typ is str:
            for v in out:
                _check_item(v)
        return out

    def flag_as_used(self, block, item, sub=""):
        try:
            self.itemlist[block][item][sub][1] = True
        except KeyError:
            pass

    def unused_items(self):
        """print_unused_items, src/input.cxx:201-217: the '[block][item] is not used!' list."""
        out = []
        for b, items in self.itemlist.items():
            for i, subs in items.items():
                for s, (_, used) in subs.items():
                    if not used:
                        out.append(f"[{b}][{i}]" + (f"[{s}]" if s else ""))
        return out


def model_ini(inp: Input):
    """The [grid]/[advec]/[diff]/[pres]/[thermo]/[boundary]/[fields]/[time]/[buffer]/[force] keys of a case, typed and
    defaulted as the reference's constructors do (file:line in SURVEY 8(b)), in the layout ``Model`` takes."""
    g = {}
    for k in ("xsize", "ysize", "zsize"):
        g[k] = inp.get_item(float, "grid", k)                                   # src/grid.cxx:53-55
    for k in ("itot", "jtot", "ktot"):
        g[k] = inp.get_item(int, "grid", k)
    g["utrans"] = inp.get_item(float, "grid", "utrans", "", 0.)
    g["vtrans"] = inp.get_item(float, "grid", "vtrans", "", 0.)
    g["lat"] = inp.get_item(float, "grid", "lat", "", inp.get_item(float, "radiation", "lat", "", 0.))
    order = inp.get_item(str, "grid", "swspatialorder")
    if order not in ("2", "4"):
        raise InputError(f'"{order}" is an illegal value for swspatialorder')  # src/grid.cxx:69-76
    g["swspatialorder"] = order
    ini = {"grid": g}
    ini["advec"] = dict(swadvec=inp.get_item(str, "advec", "swadvec", "", order),
                        cflmax=inp.get_item(float, "advec", "cflmax", "", 1.),
                        fluxlimit_list=tuple(inp.get_list(str, "advec", "fluxlimit_list", "", [])))
    ini["diff"] = dict(swdiff=inp.get_item(str, "diff", "swdiff", "", order), dnmax=inp.get_item(float, "diff", "dnmax", "", 0.4),
                       cs=inp.get_item(float, "diff", "cs", "", 0.23), tPr=inp.get_item(float, "diff", "tPr", "", 1. / 3.))
    ini["pres"] = dict(swpres=inp.get_item(str, "pres", "swpres", "", order))
    swthermo = inp.get_item(str, "thermo", "swthermo", "", "0")
    ini["thermo"] = dict(swthermo=swthermo)
    if swthermo not in ("0", "dry"):
        raise InputError(f'swthermo={swthermo} is outside this path (0 and dry are built; SURVEY 2.4)')
    if swthermo == "dry":
        ini["thermo"]["thref0"] = inp.get_item(float, "thermo", "thref0")          # src/thermo_dry.cxx:424
        base = inp.get_item(str, "thermo", "swbasestate", "", "boussinesq")
        if base not in ("boussinesq", "anelastic"):
            raise InputError(f'"{base}" is an illegal value for swbasestate')       # src/thermo_dry.cxx:435-442
        if base == "anelastic":
            raise InputError("swbasestate=anelastic: the hydrostatic base state (rhoref, thref profiles) is computed by the "
                             "reference's Thermo_dry::create, which is outside this path; pass rhoref / thref to Model instead")
    slist = inp.get_list(str, "fields", "slist", "", [])
    if swthermo == "dry" and "th" not in slist:
        slist = ["th"] + slist
    visc = inp.get_item(float, "fields", "visc")
    svisc = [inp.get_item(float, "fields", "svisc", s) for s in slist]
    # svisc is per scalar (svisc[name] falls back to svisc, src/fields.cxx): a dict when the scalars differ
    sv = dict(zip(slist, svisc))
    ini["fields"] = dict(visc=visc, svisc=(svisc[0] if len(set(svisc)) <= 1 and svisc else sv) if svisc else 0., slist=[s for s in slist if s != "th"] if swthermo == "dry" else slist)
    b = dict(mbcbot=inp.get_item(str, "boundary", "mbcbot"), mbctop=inp.get_item(str, "boundary", "mbctop"),
             swboundary=inp.get_item(str, "boundary", "swboundary", "", "default"))
    for k in ("ubot", "utop", "vbot", "vtop"):
        b[k] = inp.get_item(float, "boundary", k, "", 0.)
    b["sbc"] = {s: dict(sbcbot=inp.get_item(str, "boundary", "sbcbot", s), sbctop=inp.get_item(str, "boundary", "sbctop", s),
                        sbot=inp.get_item(float, "boundary", "sbot", s), stop=inp.get_item(float, "boundary", "stop", s))
                for s in slist}
    if inp.get_item(bool, "boundary", "swtimedep", "", False):
        raise InputError("[boundary] swtimedep=1 (time-dependent surface values, Timedep) is outside this path: "
                         "the surface boundary values are constant in time here")
    if b["swboundary"] not in ("default", "surface"):
        raise InputError(f'swboundary={b["swboundary"]} is outside this path (default and surface are built)')
    if b["swboundary"] == "surface":
        b["swconstantz0"] = inp.get_item(bool, "boundary", "swconstantz0", "", True)
        b["swcharnock"] = inp.get_item(bool, "boundary", "swcharnock", "", False)
        if b["swconstantz0"]:
            b["z0m"] = inp.get_item(float, "boundary", "z0m", "", 0.1)
            b["z0h"] = inp.get_item(float, "boundary", "z0h", "", 0.1)
        if b["swcharnock"]:
            for k in ("alpha_m", "alpha_ch", "alpha_h"):
                b[k] = inp.get_item(float, "boundary", k)
        if b["mbcbot"] == "ustar":
            b["ustar"] = inp.get_item(float, "boundary", "ustar")
    ini["boundary"] = b
    dtmax = inp.get_item(float, "time", "dtmax", "", DBIG)
    ini["time"] = dict(starttime=inp.get_item(float, "time", "starttime", "", 0.), endtime=inp.get_item(float, "time", "endtime"),
                       savetime=inp.get_item(float, "time", "savetime", "", DBIG),
                       adaptivestep=inp.get_item(bool, "time", "adaptivestep", "", True), dtmax=dtmax,
                       dt=inp.get_item(float, "time", "dt", "", dtmax), rkorder=inp.get_item(int, "time", "rkorder", "", 3),
                       outputiter=inp.get_item(int, "time", "outputiter", "", 20))
    if inp.get_item(bool, "buffer", "swbuffer", "", False):
        ini["buffer"] = dict(swbuffer=True, swupdate=inp.get_item(bool, "buffer", "swupdate", "", False),
                             zstart=inp.get_item(float, "buffer", "zstart"), sigma=inp.get_item(float, "buffer", "sigma", "", 2.),
                             beta=inp.get_item(float, "buffer", "beta", "", 2.))
    for k in ("swtimedep_ls", "swtimedep_geo", "swtimedep_wls", "swtimedep_nudge"):
        if inp.get_item(bool, "force", k, "", False):
            raise InputError(f"[force] {k}=1 (time-dependent forcings, Timedep) is outside this path")
    f = dict(swlspres=inp.get_item(str, "force", "swlspres", "", "0"), swls=inp.get_item(str, "force", "swls", "", "0"),
             swwls=inp.get_item(str, "force", "swwls", "", "0"), swnudge=inp.get_item(str, "force", "swnudge", "", "0"))
    if f["swlspres"] == "uflux":
        f["uflux"] = inp.get_item(float, "force", "uflux")
    elif f["swlspres"] == "dpdx":
        f["dpdx"] = inp.get_item(float, "force", "dpdx")
    elif f["swlspres"] == "geo":
        f["fc"] = inp.get_item(float, "force", "fc", "", -1.)
    if f["swls"] == "1":
        f["lslist"] = inp.get_list(str, "force", "lslist", "", [])
    if f["swwls"] != "0":
        f["swwls_mom"] = inp.get_item(bool, "force", "swwls_mom", "", False)
    if f["swnudge"] == "1":
        f["nudgelist"] = inp.get_list(str, "force", "nudgelist", "", [])
    ini["force"] = f
    return ini


# ---- raw binary files (serial build) ------------------------------------------------------------------------------
def _fname(name, itime, directory="."):
    return os.path.join(directory, "%s.%07d" % (name, itime))


def grid_coordinates(itot, jtot, xsize, ysize):
    """x, xh, y, yh of the interior as Grid::calculate forms them (src/grid.cxx:241-262, serial: no process offset):
    x[i] = 0.5 dx + i dx (in this order of operations: the files are compared byte for byte)."""
    dx, dy = xsize / itot, ysize / jtot
    i, j = np.arange(itot, dtype=np.float64), np.arange(jtot, dtype=np.float64)
    return 0.5 * dx + i * dx, i * dx, 0.5 * dy + j * dy, j * dy


def save_grid(x, xh, y, yh, z, zh, directory="."):
    """Grid::save_grid, src/grid_serial.cxx:46-77: x, xh, y, yh, z, zh interiors back to back (fopen "wbx": no overwrite)."""
    with open(_fname("grid", 0, directory), "xb") as fh:
        for a in (x, xh, y, yh, z, zh):
            fh.write(np.ascontiguousarray(a, dtype="<f8").tobytes())


def load_grid(itot, jtot, ktot, directory="."):
    """Grid::load_grid, src/grid_serial.cxx:79-111 → dict of the six interior coordinate arrays."""
    path = _fname("grid", 0, directory)
    want = 8 * 2 * (itot + jtot + ktot)
    if not os.path.exists(path) or os.path.getsize(path) != want:
        raise RuntimeError("Error in grid")
    a = np.fromfile(path, dtype="<f8")
    out, o = {}, 0
    for name, n in (("x", itot), ("xh", itot), ("y", jtot), ("yh", jtot), ("z", ktot), ("zh", ktot)):
        out[name] = a[o:o + n].copy()
        o += n
    return out


def save_field3d(fld, gridshape, name, itime, offset=0., directory="."):
    """Field3d_io::save_field3d, src/field3d_io.cxx:671-709: (data + offset) of the interior, i fastest, no ghost cells.
    ``fld`` is the ghosted (kcells, jcells, icells) array, ``gridshape`` = dict(igc, jgc, kgc, itot, jtot, ktot)."""
    g = gridshape
    a = np.asarray(fld)[g["kgc"]:g["kgc"] + g["ktot"], g["jgc"]:g["jgc"] + g["jtot"], g["igc"]:g["igc"] + g["itot"]]
    with open(_fname(name, itime, directory), "xb") as fh:
        fh.write(np.ascontiguousarray(a + offset, dtype="<f8").tobytes())


def load_field3d(gridshape, name, itime, offset=0., directory="."):
    """Field3d_io::load_field3d, src/field3d_io.cxx:711-751 → ghosted array (ghost cells zero), offset removed."""
    g = gridshape
    path = _fname(name, itime, directory)
    n = g["itot"] * g["jtot"] * g["ktot"]
    if not os.path.exists(path) or os.path.getsize(path) != 8 * n:
        raise RuntimeError(f"Error loading field of {name}")       # Fields::load, src/fields.cxx
    a = np.fromfile(path, dtype="<f8").reshape(g["ktot"], g["jtot"], g["itot"])
    out = np.zeros((g["ktot"] + 2 * g["kgc"], g["jtot"] + 2 * g["jgc"], g["itot"] + 2 * g["igc"]))
    out[g["kgc"]:g["kgc"] + g["ktot"], g["jgc"]:g["jgc"] + g["jtot"], g["igc"]:g["igc"] + g["itot"]] = a - offset
    return out


_TIME = struct.Struct("<QQi")     # unsigned long itime, idt; int iteration (LP64), src/timeloop.cxx:462-464


def save_time(starttime, itime, idt, iteration, directory="."):
    with open(_fname("time", starttime, directory), "xb") as fh:
        fh.write(_TIME.pack(int(itime), int(idt), int(iteration)))


def load_time(starttime, directory="."):
    """Timeloop::load → (itime, idt, iteration); time = itime / 1e9, dt = idt / 1e9."""
    path = _fname("time", starttime, directory)
    if not os.path.exists(path):
        raise RuntimeError(f"{os.path.basename(path)} does not exist")
    with open(path, "rb") as fh:
        return _TIME.unpack(fh.read(_TIME.size))


class OutWriter:
    """<case>.out: header once, then one line per status check (src/model.cxx:795-830)."""
    HEADER = "%8s %13s %10s %11s %8s %8s %11s %16s %16s\n" % ("ITER", "TIME", "CPUDT", "DT", "CFL", "DNUM", "DIV", "MOM", "TKE")
    ROW = "%8d %13.6G %10.4f %11.3E %8.4f %8.4f %11.3E %16.8E %16.8E\n"

    def __init__(self, sim_name, directory="."):
        self.fh = open(os.path.join(directory, sim_name + ".out"), "a", buffering=1)
        self.fh.write(self.HEADER)

    @classmethod
    def format_row(cls, iteration, time, cputime, dt, cfl, dn, div, mom, tke):
        return cls.ROW % (iteration, time, cputime, dt, cfl, dn, div, mom, tke)

    def write(self, status, cputime=0.):
        """``status`` = Model.print_status(); raises like the reference when the CFL number is not finite / not in [0,10)."""
        s = status
        self.fh.write(self.format_row(s["iter"], s["time"], cputime, s["dt"], s["cfl"], s["dn"], s["div"], s["mom"], s["tke"]))
        self.fh.flush()
        if not (0. <= s["cfl"] < 10.) or not np.isfinite(s["cfl"]):
            raise RuntimeError("Simulation has non-finite numbers")

    def close(self):
        self.fh.close()


def read_out(path):
    """Parse a <case>.out back into a list of dicts (for comparing against a reference run's file)."""
    rows = []
    with open(path) as fh:
        for line in fh:
            p = line.split()
            if not p or p[0] == "ITER":
                continue
            rows.append(dict(iter=int(p[0]), time=float(p[1]), cputime=float(p[2]), dt=float(p[3]), cfl=float(p[4]),
                             dn=float(p[5]), div=float(p[6]), mom=float(p[7]), tke=float(p[8])))
    return rows
