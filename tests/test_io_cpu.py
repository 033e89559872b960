"""CPU tests of the file formats either side of the path: the .ini reader's semantics and error
texts (src/input.cxx), the raw binary grid/field/time layouts of the serial build (byte-exact), and the .out row format
checked against the C library's own printf (the reference uses std::fprintf, src/model.cxx:827)."""
import ctypes
import os
import struct

import numpy as np
import pytest

from microhh_b200 import io as IO

HERE = os.path.dirname(os.path.abspath(__file__))


def test_ini_reader_semantics():
    inp = IO.Input(os.path.join(HERE, "golden", "gabls1_like.ini"))
    assert inp.get_item(int, "grid", "ktot") == 32                       # trailing comment stripped
    assert inp.get_item(float, "boundary", "sbot", "th") == 264.5         # sub-item wins
    assert inp.get_item(float, "boundary", "sbot", "qt") == 265.          # falls back to the plain item
    assert inp.get_item(float, "boundary", "stop", "th") == 0.01
    assert inp.get_item(bool, "time", "adaptivestep", "", False) is True
    assert inp.get_item(bool, "buffer", "swbuffer") is True
    assert inp.get_item(str, "advec", "swadvec", "", "4") == "2"
    assert inp.get_item(float, "grid", "utrans", "", 0.) == 0.            # default
    assert inp.get_list(str, "fields", "slist", "", []) == []
    with pytest.raises(IO.InputError, match='Block "nope" does not exist'):
        inp.get_item(int, "nope", "x")
    with pytest.raises(IO.InputError, match='Item "x" in group "grid" does not exist'):
        inp.get_item(int, "grid", "x")
    with pytest.raises(IO.InputError, match=r"\[grid\]\[xsize\] = 400. is illegal."):
        inp.get_item(int, "grid", "xsize", "", 3)                         # present but of the wrong type
    with pytest.raises(IO.InputError, match="Item does not match type"):
        inp.get_item(bool, "grid", "itot")
    unused = inp.unused_items()
    assert "[stats][swstats]" in unused and "[fields][rndamp][th]" in unused and "[grid][ktot]" not in unused


def test_ini_reader_errors_and_lists():
    with pytest.raises(IO.InputError, match="No block name found"):
        IO.Input(text="a=1\n")
    with pytest.raises(IO.InputError, match='Line "a=b=c" is illegal'):
        IO.Input(text="[x]\na=b=c\n")
    with pytest.raises(IO.InputError, match='Line "hello" is illegal'):
        IO.Input(text="[x]\nhello\n")
    with pytest.raises(IO.InputError, match="cannot be opened"):
        IO.Input("/nonexistent/file.ini")
    inp = IO.Input(text="[fields]\nslist = s1, s2 ,s3\nvals=1,2.5,3e2\n[t]\nflags=1,false,TRUE\n")
    assert inp.get_list(str, "fields", "slist") == ["s1", "s2", "s3"]
    assert inp.get_list(float, "fields", "vals") == [1., 2.5, 300.]
    assert inp.get_list(bool, "t", "flags") == [True, False, True]
    assert inp.get_list(int, "t", "absent", "", [7]) == [7]
    with pytest.raises(IO.InputError):
        inp.get_list(int, "fields", "vals")


def test_model_ini_from_case_file():
    ini = IO.model_ini(IO.Input(os.path.join(HERE, "golden", "gabls1_like.ini")))
    assert ini["grid"]["swspatialorder"] == "2" and ini["grid"]["itot"] == 32 and ini["grid"]["lat"] == 73.
    assert ini["advec"] == dict(swadvec="2", cflmax=1.0, fluxlimit_list=())
    assert ini["diff"]["swdiff"] == "smag2" and ini["diff"]["dnmax"] == 0.3 and ini["diff"]["cs"] == 0.23
    assert ini["pres"]["swpres"] == "2"                                   # defaults to swspatialorder
    assert ini["thermo"] == dict(swthermo="dry", thref0=263.5)
    assert ini["boundary"]["sbc"]["th"] == dict(sbcbot="dirichlet", sbctop="neumann", sbot=264.5, stop=0.01)
    assert ini["boundary"]["swboundary"] == "surface" and ini["boundary"]["z0m"] == 0.1
    assert ini["buffer"] == dict(swbuffer=True, swupdate=False, zstart=300., sigma=0.00223, beta=2.)
    assert ini["force"]["swlspres"] == "geo" and ini["force"]["fc"] == 1.39e-4
    assert ini["time"]["rkorder"] == 3 and ini["time"]["dtmax"] == 60. and ini["time"]["adaptivestep"] is True
    bad = IO.Input(text="[grid]\nitot=4\njtot=4\nktot=4\nxsize=1\nysize=1\nzsize=1\nswspatialorder=3\n")
    with pytest.raises(IO.InputError, match="swspatialorder"):
        IO.model_ini(bad)


def test_binary_layouts(tmp_path):
    d = str(tmp_path)
    itot, jtot, ktot, gc = 6, 4, 3, (2, 1, 3)
    gs = dict(itot=itot, jtot=jtot, ktot=ktot, igc=gc[0], jgc=gc[1], kgc=gc[2])
    rng = np.random.default_rng(0)
    fld = rng.uniform(-1, 1, (ktot + 2 * gc[2], jtot + 2 * gc[1], itot + 2 * gc[0]))
    IO.save_field3d(fld, gs, "u", 3600, offset=0.5, directory=d)
    raw = np.fromfile(os.path.join(d, "u.0003600"), dtype="<f8")
    assert raw.size == itot * jtot * ktot
    # i fastest, then j, then k; ghost cells dropped; offset added (src/field3d_io.cxx:686-704)
    assert raw[1] == fld[gc[2], gc[1], gc[0] + 1] + 0.5
    assert raw[itot] == fld[gc[2], gc[1] + 1, gc[0]] + 0.5
    assert raw[itot * jtot] == fld[gc[2] + 1, gc[1], gc[0]] + 0.5
    back = IO.load_field3d(gs, "u", 3600, offset=0.5, directory=d)
    inner = (slice(gc[2], gc[2] + ktot), slice(gc[1], gc[1] + jtot), slice(gc[0], gc[0] + itot))
    assert np.allclose(back[inner], fld[inner], rtol=0, atol=1e-15) and back[0].max() == 0.
    with pytest.raises(FileExistsError):                                  # fopen(..., "wbx")
        IO.save_field3d(fld, gs, "u", 3600, directory=d)
    with pytest.raises(RuntimeError):
        IO.load_field3d(gs, "v", 0, directory=d)

    x = (np.arange(itot) + 0.5); xh = np.arange(itot) * 1.; y = np.arange(jtot) + 0.5; yh = np.arange(jtot) * 1.
    z = np.array([0.1, 0.4, 0.9]); zh = np.array([0., 0.25, 0.65])
    IO.save_grid(x, xh, y, yh, z, zh, directory=d)
    assert os.path.getsize(os.path.join(d, "grid.0000000")) == 8 * 2 * (itot + jtot + ktot)
    g = IO.load_grid(itot, jtot, ktot, directory=d)
    assert np.array_equal(g["z"], z) and np.array_equal(g["xh"], xh) and np.array_equal(g["yh"], yh)
    with pytest.raises(RuntimeError, match="Error in grid"):
        IO.load_grid(itot + 1, jtot, ktot, directory=d)

    IO.save_time(3600, 3600 * 10**9, 6 * 10**9, 1234, directory=d)
    b = open(os.path.join(d, "time.0003600"), "rb").read()
    assert len(b) == 20 and struct.unpack("<Q", b[:8])[0] == 3600 * 10**9 and struct.unpack("<i", b[16:])[0] == 1234
    assert IO.load_time(3600, directory=d) == (3600 * 10**9, 6 * 10**9, 1234)


def test_out_row_matches_c_printf(tmp_path):
    libc = ctypes.CDLL(None)
    vals = (120, 1234.56789, 0.0123, 6.0, 0.98765, 0.01234, 3.2e-15, 7.99998, 32.2384)
    buf = ctypes.create_string_buffer(512)
    libc.snprintf(buf, 512, IO.OutWriter.ROW.encode(), ctypes.c_int(vals[0]), *[ctypes.c_double(v) for v in vals[1:]])
    assert IO.OutWriter.format_row(*vals) == buf.value.decode()
    w = IO.OutWriter("case", directory=str(tmp_path))
    st = dict(iter=vals[0], time=vals[1], dt=vals[3], cfl=vals[4], dn=vals[5], div=vals[6], mom=vals[7], tke=vals[8], mass=0.)
    w.write(st, cputime=vals[2])
    with pytest.raises(RuntimeError, match="non-finite"):
        w.write(dict(st, cfl=float("nan")))
    w.close()
    lines = open(os.path.join(str(tmp_path), "case.out")).read().splitlines()
    assert lines[0].split() == ["ITER", "TIME", "CPUDT", "DT", "CFL", "DNUM", "DIV", "MOM", "TKE"]
    rows = IO.read_out(os.path.join(str(tmp_path), "case.out"))
    assert rows[0]["iter"] == 120 and abs(rows[0]["tke"] - 32.2384) < 1e-7


def test_binary_files_against_the_reference_writers(tmp_path, monkeypatch):
    """The restart files either side of the path, byte for byte against files the reference's own code writes
    (oracle/ref_harness/k_io.cxx, k_timeloop.cxx: Field3d_io::save_field3d / load_field3d src/field3d_io.cxx:671-751,
    Grid::save_grid src/grid_serial.cxx:46-77, Timeloop::save / load src/timeloop.cxx:445-525 compiled in place), and each
    side reading what the other wrote."""
    import ctypes as C
    from oracle.ref import RefWorld, lib
    from tests.helpers import make_case
    c = make_case(itot=12, jtot=10, ktot=9, alpha=1.3)
    W = RefWorld(c["itot"], c["jtot"], c["ktot"], c["xsize"], c["ysize"], c["zsize"], c["z"])
    shape = dict(itot=12, jtot=10, ktot=9, igc=1, jgc=1, kgc=1)
    refdir, mydir = tmp_path / "ref", tmp_path / "mine"
    refdir.mkdir(); mydir.mkdir()
    L = lib()
    u = c["u"].copy()
    # --- fields (with a Galilean offset, as Fields::save passes utrans)
    monkeypatch.chdir(refdir)
    assert L.ref_save_field3d(W.h, u.ctypes.data_as(C.c_void_p), b"u.0000600", C.c_double(1.25)) == 0
    IO.save_field3d(u, shape, "u", 600, offset=1.25, directory=str(mydir))
    assert (refdir / "u.0000600").read_bytes() == (mydir / "u.0000600").read_bytes()
    assert L.ref_save_field3d(W.h, u.ctypes.data_as(C.c_void_p), b"u.0000600", C.c_double(0.)) == 1      # "wbx": never overwrites
    with pytest.raises(FileExistsError):
        IO.save_field3d(u, shape, "u", 600, directory=str(mydir))
    mine = IO.load_field3d(shape, "u", 600, offset=1.25, directory=str(refdir))                           # I read the reference's file
    back = np.zeros_like(u)
    monkeypatch.chdir(mydir)
    assert L.ref_load_field3d(W.h, back.ctypes.data_as(C.c_void_p), b"u.0000600", C.c_double(1.25)) == 0   # it reads mine
    assert np.array_equal(mine[1:-1, 1:-1, 1:-1], back[1:-1, 1:-1, 1:-1])
    assert np.abs(back[1:-1, 1:-1, 1:-1] - u[1:-1, 1:-1, 1:-1]).max() <= 4e-16 * 2.25                      # (u + o) - o
    assert L.ref_load_field3d(W.h, back.ctypes.data_as(C.c_void_p), b"nope.0000000", C.c_double(0.)) == 1
    # --- grid
    monkeypatch.chdir(refdir)
    L.ref_save_grid(W.h)
    x, xh, y, yh = IO.grid_coordinates(12, 10, c["xsize"], c["ysize"])
    IO.save_grid(x, xh, y, yh, W.z[1:-1], W.zh[1:-1], directory=str(mydir))
    assert (refdir / "grid.0000000").read_bytes() == (mydir / "grid.0000000").read_bytes()
    got = IO.load_grid(12, 10, 9, directory=str(refdir))
    assert np.array_equal(got["z"], c["z"]) and np.array_equal(got["xh"], xh)
    # --- time
    L.ref_time_save.argtypes = [C.c_void_p, C.c_int, C.c_ulong, C.c_ulong, C.c_int]
    L.ref_time_save(W.h, 600, 600123456789, 1500000000, 4242)
    IO.save_time(600, 600123456789, 1500000000, 4242, directory=str(mydir))
    assert (refdir / "time.0000600").read_bytes() == (mydir / "time.0000600").read_bytes()
    assert IO.load_time(600, directory=str(refdir)) == (600123456789, 1500000000, 4242)
    out = (C.c_ulong * 3)()
    monkeypatch.chdir(mydir)
    L.ref_time_load(W.h, 600, out)
    assert tuple(out) == (600123456789, 1500000000, 4242)
    W.close()


class _RefInput:
    """The reference's own Input class (src/input.cxx compiled in place, oracle/_ref/liboracle_input.so)."""

    def __init__(self, path):
        import ctypes as C
        so = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "liboracle_input.so")
        if not os.path.exists(so):
            pytest.skip("oracle/_ref/liboracle_input.so not built (needs /root/reference)")
        self.C, self.lib = C, C.CDLL(so)
        self.lib.refin_open.restype = C.c_void_p
        err = C.create_string_buffer(512)
        self.h = self.lib.refin_open(str(path).encode(), err, 512)
        self.error = None if self.h else err.value.decode()

    def get(self, typ, block, item, sub="", default=None):
        C = self.C
        out = C.create_string_buffer(1024)
        code = {int: b"i", float: b"d", bool: b"b", str: b"s"}[typ]
        d = "" if default is None else ("true" if default is True else "false" if default is False else str(default))
        rc = self.lib.refin_get(C.c_void_p(self.h), C.c_char(code), block.encode(), item.encode(), sub.encode(),
                                0 if default is None else 1, d.encode(), out, 1024)
        return rc, out.value.decode()

    def get_list(self, typ, block, item, sub=""):
        C = self.C
        out = C.create_string_buffer(1024)
        code = {int: b"i", float: b"d", bool: b"b", str: b"s"}[typ]
        rc = self.lib.refin_get_list(C.c_void_p(self.h), C.c_char(code), block.encode(), item.encode(), sub.encode(), out, 1024)
        return rc, out.value.decode()


def _mine(inp, typ, block, item, sub="", default=None):
    try:
        v = inp.get_item(typ, block, item, sub) if default is None else inp.get_item(typ, block, item, sub, default)
    except IO.InputError as e:
        return 1, str(e)
    return 0, ("true" if v is True else "false" if v is False else repr(float(v)) if typ is float else str(v))


def test_ini_reader_against_the_reference_reader(tmp_path):
    """IO.Input against the reference's Input (src/input.cxx:76-330, include/input_tools.h) on the same files: values,
    fall-back from sub-item to item, defaults, type errors and every error text."""
    path = os.path.join(HERE, "golden", "gabls1_like.ini")
    ref, mine = _RefInput(path), IO.Input(path)
    assert ref.error is None
    queries = [(int, "grid", "ktot"), (float, "grid", "xsize"), (str, "advec", "swadvec"), (bool, "buffer", "swbuffer"),
               (float, "boundary", "sbot", "th"), (float, "boundary", "sbot", "qt"), (float, "boundary", "stop", "th"),
               (bool, "time", "adaptivestep", "", False), (float, "grid", "utrans", "", 0.5), (int, "time", "rkorder", "", 3),
               (str, "thermo", "swbasestate", "", "x"), (int, "nope", "x"), (int, "grid", "x"), (int, "grid", "xsize"),
               (bool, "grid", "itot"), (int, "grid", "xsize", "", 3), (float, "fields", "visc"), (str, "boundary", "sbcbot", "th")]
    for q in queries:
        r, m = ref.get(*q), _mine(mine, *q)
        assert r[0] == m[0], (q, r, m)
        if r[0] == 0 and q[0] is float:
            assert float(r[1]) == float(m[1]), (q, r, m)
        else:
            assert r[1].strip() == m[1].strip(), (q, r, m)
    # lists and malformed files
    good = tmp_path / "lists.ini"
    good.write_text("[fields]\nslist = s1, s2 ,s3\nvals=1,2.5,3e2   # a comment\n[t]\nflags=1,false,TRUE\nn = 4,5\n")
    ref2, mine2 = _RefInput(good), IO.Input(str(good))
    for typ, b, i in ((str, "fields", "slist"), (float, "fields", "vals"), (bool, "t", "flags"), (int, "t", "n"), (int, "fields", "vals"),
                      (int, "t", "absent")):
        r = ref2.get_list(typ, b, i)
        try:
            v = mine2.get_list(typ, b, i)
            m = (0, v)
        except IO.InputError as e:
            m = (1, str(e))
        assert r[0] == m[0], (typ, b, i, r, m)
        if r[0] == 0:
            conv = {str: str, float: float, int: int, bool: lambda t: t == "true"}[typ]
            assert [conv(t) for t in r[1].split(",")] == m[1], (r, m)
        else:
            assert r[1].strip() == m[1].strip(), (r, m)
    for n, text in enumerate(("a=1\n", "[x]\na=b=c\n", "[x]\nhello\n", "[x]\na[b=1\n", "[bad name]\n")):
        bad = tmp_path / f"bad{n}.ini"
        bad.write_text(text)
        r = _RefInput(bad)
        try:
            IO.Input(str(bad))
            m = None
        except IO.InputError as e:
            m = str(e)
        assert (r.error is None) == (m is None), (text, r.error, m)
        if m is not None:
            assert r.error.strip() == m.strip(), (text, r.error, m)
    assert _RefInput(tmp_path / "missing.ini").error.strip().startswith('"' + str(tmp_path / "missing.ini") + '" cannot be opened')


def test_model_ini_rejects_switches_outside_the_path():
    """Case files whose keys this path cannot honour raise instead of running different physics silently (cases/gabls1 as
    shipped sets swtimedep=1 for th_sbot)."""
    base = ("[grid]\nitot=8\njtot=8\nktot=8\nxsize=1\nysize=1\nzsize=1\nswspatialorder=2\n[fields]\nvisc=1e-5\nsvisc=1e-5\n"
            "[time]\nendtime=10\n[boundary]\nmbcbot=noslip\nmbctop=freeslip\nsbcbot=dirichlet\nsbctop=neumann\nsbot=1\nstop=0\n")
    assert IO.model_ini(IO.Input(text=base))["boundary"]["swboundary"] == "default"
    for extra, msg in (("swtimedep=1\n", "swtimedep=1"), ("swboundary=surface_bulk\n", "outside this path")):
        with pytest.raises(IO.InputError, match=msg):
            IO.model_ini(IO.Input(text=base + extra))
    with pytest.raises(IO.InputError, match="anelastic"):
        IO.model_ini(IO.Input(text=base + "[thermo]\nswthermo=dry\nthref0=300\nswbasestate=anelastic\n"))
    with pytest.raises(IO.InputError, match="illegal value for swbasestate"):
        IO.model_ini(IO.Input(text=base + "[thermo]\nswthermo=dry\nthref0=300\nswbasestate=foo\n"))
    with pytest.raises(IO.InputError, match="swthermo=moist"):
        IO.model_ini(IO.Input(text=base + "[thermo]\nswthermo=moist\n"))
    with pytest.raises(IO.InputError, match="swtimedep_geo"):
        IO.model_ini(IO.Input(text=base + "[force]\nswtimedep_geo=1\n"))
    two = IO.model_ini(IO.Input(text=base.replace("svisc=1e-5", "svisc=1e-5\nsvisc[q]=2e-5\nslist=q,s")))
    assert two["fields"]["svisc"] == {"q": 2e-5, "s": 1e-5}
