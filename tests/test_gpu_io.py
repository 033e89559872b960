"""The case-file entry point on the GPU: Model.from_case reads <case>.ini + grid.0000000 + the raw
binary fields/time a `microhh init` would leave, runs, writes <case>.out and restart files; a restart from those files
continues bit-identically (the files hold the full prognostic state and the integer clock)."""
import os
import shutil

import numpy as np
import pytest

from microhh_b200 import io as IO

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_case_files_roundtrip_and_restart(cuda_lib, tmp_path):
    import torch
    from microhh_b200.model import Model
    d = str(tmp_path)
    shutil.copy(os.path.join(HERE, "golden", "gabls1_like.ini"), os.path.join(d, "gabls1.ini"))
    n, size = 32, 400.
    z = (np.arange(n) + 0.5) * size / n
    zh = np.arange(n) * size / n
    x = (np.arange(n) + 0.5) * size / n
    xh = np.arange(n) * size / n
    IO.save_grid(x, xh, x, xh, z, zh, directory=d)
    gs = dict(itot=n, jtot=n, ktot=n, igc=1, jgc=1, kgc=1)
    rng = np.random.default_rng(3)
    shape = (n + 2, n + 2, n + 2)
    zf = np.zeros(n + 2); zf[1:-1] = z
    th = 265. + 0.01 * np.clip(zf - 100., 0., None)
    init = dict(u=np.full(shape, 8.), v=np.zeros(shape), w=np.zeros(shape),
                th=th[:, None, None] + 0.1 * rng.uniform(-1, 1, shape))
    for name, a in init.items():
        IO.save_field3d(a, gs, name, 0, directory=d)
    IO.save_time(0, 0, 6 * 10**9, 0, directory=d)
    s2 = (n + 2, n + 2)
    surface = dict(z0m=np.full(s2, 0.1), dudz=np.full(s2, 0.03), dvdz=np.zeros(s2), dbdz=np.full(s2, 1e-5),
                   ufluxbot=np.full(s2, -0.01), vfluxbot=np.zeros(s2), sfluxbot=np.full(s2, -0.007),
                   ufluxtop=np.zeros(s2), vfluxtop=np.zeros(s2), sfluxtop=np.zeros(s2))
    dev = torch.device("cuda:0")
    sdev = {k: torch.from_numpy(v).to(dev) for k, v in surface.items()}
    profs = dict(u=np.full(n, 8.), v=np.zeros(n), th=th[1:-1], u_geo=np.full(n, 8.), v_geo=np.zeros(n))

    def start(iotime):
        m = Model.from_case(os.path.join(d, "gabls1.ini"), device=dev, surface=sdev, profs=profs, iotime=iotime)
        f = m.fields
        f.mp["u"].flux_bot.copy_(sdev["ufluxbot"]); f.mp["v"].flux_bot.copy_(sdev["vfluxbot"]); f.sp["th"].flux_bot.copy_(sdev["sfluxbot"])
        return m

    m = start(0)
    assert m.timeloop.idt == 6 * 10**9 and m.buffer.swbuffer and m.force.swlspres == "geo" and m.diff.swdiff == "smag2"
    got = m.fields.sp["th"].fld.cpu().numpy()
    assert np.array_equal(got[1:-1, 1:-1, 1:-1], init["th"][1:-1, 1:-1, 1:-1])
    out = IO.OutWriter("gabls1", directory=d)
    for _ in range(3):
        m.step()
    out.write(m.print_status(), cputime=0.)
    m.save(directory=d, iotime=1)
    for _ in range(2):
        m.step()
    out.write(m.print_status(), cputime=0.)
    out.close()
    want = {nme: f.fld.cpu().numpy().copy() for nme, f in m.fields.ap.items()}
    tend = (m.timeloop.itime, m.timeloop.idt, m.timeloop.iteration)
    m.grid.close()

    rows = IO.read_out(os.path.join(d, "gabls1.out"))
    assert [r["iter"] for r in rows] == [3, 5] and all(0. < r["cfl"] < 1.2 and r["div"] < 1e-10 for r in rows)

    m2 = start(1)
    assert m2.timeloop.iteration == 3
    # the RK tendencies are zero after a full step (cA[0] = 0), so fields + clock are the whole state
    for _ in range(2):
        m2.step()
    assert (m2.timeloop.itime, m2.timeloop.idt, m2.timeloop.iteration) == tend
    inner = (slice(1, -1),) * 3
    for nme, f in m2.fields.ap.items():
        assert np.array_equal(f.fld.cpu().numpy()[inner], want[nme][inner]), nme
    m2.grid.close()
