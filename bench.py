#!/usr/bin/env python
"""bench.py — grid-point-steps/s of a full RK3 LES step (BASELINE.json metric) on N B200s.

Workload (config.workload): cases/gabls1 as shipped (advec_2, diff_smag2, thermo dry, swboundary=surface with the
Monin-Obukhov surface layer solved on the device every substep, swbuffer=1, swlspres=geo fc=1.39e-4, pres_2, RK3;
only the time-dependent surface temperature, Timedep, is left out) scaled to a 512^3 synthetic grid per GPU
(BASELINE.json configs[2]); for N > 1 the domain is split into y-slabs (weak scaling).
One "step" = 3 RK substeps, each = cyclic + ghost BCs, evisc, surface layer, buoyancy, advection, diffusion,
sponge layer, Coriolis force, pressure solve, RK update; the CFL/diffusion-number maxima once per step.

  python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
  python bench.py --impl reference ...                      # the reference's own CPU code (oracle/_ref)

Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

# algorithmic bytes per grid-point (FP64), LES config with one scalar
BYTES_PER_POINT = {
    "evisc_kernel": 40,              # R u,v,w,N2 -> W evisc
    "advec2_uvw_kernel": 72,         # R u,v,w,ut,vt,wt -> W ut,vt,wt
    "advec2_s_kernel": 48,
    "smag2_diff_uvw_kernel": 80,
    "smag2_diff_c_kernel": 32,
    # one-pass tendencies (no SURVEY row: they replace thermo + advec + diff of the rows above)
    "advec2_smag2_uvw_kernel": 88,   # R u,v,w,evisc,th ; R+W ut,vt,wt
    "advec2_smag2_c_kernel": 64,     # R s,evisc,u,v,w ; R+W st ; W s_new (the scalar's RK update rides in the pass)
    "pres2_in_kernel": 56,
    "fft_fwd_x": 16, "fft_fwd_y": 16, "fft_bwd_y": 16, "fft_bwd_x": 16,
    "pres2_tdma_kernel": 32,
    "pres2_out_kernel": 56,
    "pres2_out_rk_kernel": 104,      # R p,ut,vt,wt,u,v,w -> W ut,vt,wt,u,v,w (Pres_2::output + RK of u,v,w fused)
    "rk_kernel": 32,                 # the scalar: R a,at ; W a,at
    "n2_kernel": 16, "buoyancy_tend_kernel": 24,
    "advec2_cfl_kernel": 24, "smag2_dnmul_kernel": 8,
}
# the same launches on SURVEY 8(d)'s unfused rows (roofline.frac_survey)
SURVEY_BYTES_PER_POINT = {"advec2_smag2_uvw_kernel": 72 + 80, "advec2_smag2_c_kernel": 48 + 32 + 32, "evisc_kernel": 40,
                          "pres2_out_rk_kernel": 56 + 96, "pres2_in_kernel": 56, "pres2_tdma_kernel": 32,
                          "fft_fwd_x": 16, "fft_fwd_y": 16, "fft_bwd_y": 16, "fft_bwd_x": 16}
BYTES_PER_STEP = 3 * 608 + 32       # 1856 B / grid-point-step (the reference's operator-by-operator passes)
# what this implementation's fused passes move: evisc 40, momentum 88, scalar + its RK 64, pres_in 56, 4 FFT passes 64,
# tdma 28 (two sweeps over p + pivot checkpoints), output+RK 104 per substep; the cfl / dnmul maxima ride in the evisc pass;
# lazy tendencies: the last substep writes no ut,vt,wt (24) and no zeroed tht (8), the first reads none of them (32)
BYTES_PER_STEP_FUSED = 3 * 444 - 64


def ncu_traffic(kernel, n):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (GB) from the committed ncu --set full capture of this
    workload (profiles/r01_traffic.json), or None."""
    try:
        path = os.path.join(ROOT, "profiles", "r02_traffic.json")
        if not os.path.exists(path):
            path = os.path.join(ROOT, "profiles", "r01_traffic.json")
        with open(path) as f:
            t = json.load(f)
        return t["gb_per_launch"].get(kernel) if t.get("n") == n else None
    except (OSError, ValueError, KeyError):
        return None


def weak_scaling_shape(n, world):
    """Global grid with n^3 points per GPU, y-slabs.  The point count per GPU is fixed (weak scaling); the extra points go
    to y first, then x, then z, which keeps the transform lengths at <= 2n: 512^3, 512x1024x512, 1024x1024x512
    (BASELINE configs[4]'s named shape, on 4 GPUs), 1024x1024x1024."""
    shape = [n, n, n]               # itot, jtot, ktot
    order = [1, 0, 2]
    w, a = world, 0
    while w > 1 and w % 2 == 0:
        shape[order[a % 3]] *= 2
        w //= 2; a += 1
    shape[1] *= w                   # a non-power-of-two remainder goes to y
    return tuple(shape)


def les_ini(itot, jtot, ktot):
    size = 12.5
    return {
        "grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=size * itot, ysize=size * jtot, zsize=size * ktot, swspatialorder="2"),
        "advec": dict(swadvec="2", cflmax=1.0),
        "diff": dict(swdiff="smag2", cs=0.1, dnmax=0.3, tPr=1. / 3.),
        "pres": dict(swpres="2"),
        "thermo": dict(swthermo="dry", thref0=263.5),
        # cases/gabls1/gabls1.ini: [boundary] (without swtimedep), [buffer] (zstart = 300 of 400 m), [force]
        "boundary": dict(mbcbot="noslip", mbctop="freeslip", swboundary="surface", z0m=0.1, z0h=0.1,
                         sbc={"th": dict(sbcbot="dirichlet", sbctop="flux", sbot=265., stop=0.)}),
        "buffer": dict(swbuffer=True, zstart=0.75 * size * ktot, sigma=0.00223, beta=2.),
        "force": dict(swlspres="geo", fc=1.39e-4),
        "fields": dict(visc=1e-5, svisc=1e-5),
        "time": dict(rkorder=3, dt=1.0, dtmax=60., adaptivestep=True),
    }


def les_profs(z):
    """cases/gabls1/gabls1_input.py: u = u_geo = 8, th = 265 up to 100 m, then 0.01 K/m."""
    n = len(z)
    return dict(u=np.full(n, 8.), v=np.zeros(n), th=265. + 0.01 * np.clip(z - 100., 0., None), u_geo=np.full(n, 8.), v_geo=np.zeros(n))


def synthetic_fields(shape, ks, ke, z, seed, jslice=None):
    """u,v ~ U(-1,1)+8, w ~ 0.2 U(-1,1) (0 at walls), th = gabls1's profile + U(-0.05,0.05)."""
    rng = np.random.default_rng(seed)
    kc = shape[0]
    zf = np.zeros(kc)
    zf[ks:ke] = z
    u = 8. + rng.uniform(-1, 1, shape)
    v = rng.uniform(-1, 1, shape)
    w = 0.2 * rng.uniform(-1, 1, shape)
    w[:ks + 1] = 0.
    w[ke:] = 0.
    th = (265. + 0.01 * np.clip(zf - 100., 0., None))[:, None, None] + rng.uniform(-0.05, 0.05, shape)
    return dict(u=u, v=v, w=w, th=th)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for nm, val in zip(names, out[2:]):
                    if "Active" in val and "Not" not in val:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------------
class quiet_stdout:
    """The reference's Master/FFT print status lines to stdout (src/master_serial.cxx, src/fft.cxx);
    bench.py must print exactly ONE JSON line, so fd 1 points at /dev/null while reference code runs."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *exc):
        os.dup2(self.saved, 1)
        os.close(self.saved)
        os.close(self.null)


def cpu_reference_run(n, steps, warmup):
    with quiet_stdout():
        return _cpu_reference_run(n, steps, warmup)


def _cpu_reference_run(n, steps, warmup):
    """The reference's own CPU code (oracle/_ref timing build, 1 thread: its non-MPI build is single-threaded,
    src/model.cxx:325) on a bounded n^3 sample of the same gabls1 configuration (surface layer, sponge, Coriolis included)."""
    import oracle.ref as R
    R.use_perf_build()
    from oracle.les import OracleLES
    ini = les_ini(n, n, n)
    g = ini["grid"]
    z = (np.arange(n) + 0.5) * (g["zsize"] / n)
    shape = (n + 2, n + 2, n + 2)
    f = synthetic_fields(shape, 1, n + 1, z, seed=666)
    bi = ini["boundary"]
    o = OracleLES(n, n, n, g["xsize"], g["ysize"], g["zsize"], z, f, swadvec="2", swdiff="smag2", visc=1e-5, svisc=1e-5,
                  cs=0.1, tPr=1. / 3., thref0=263.5, mbcbot="noslip", mbctop="freeslip", sbc=bi["sbc"], dt=1.0, dtmax=60.,
                  dnmax=0.3, buffer=ini["buffer"], force=ini["force"], profs=les_profs(z), mo=dict(z0m=bi["z0m"], z0h=bi["z0h"]))
    for _ in range(warmup):
        o.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        o.step()
    dt = time.perf_counter() - t0
    o.close()
    return n ** 3 * steps / dt, dt / steps


def _cpu_worker(n, steps, warmup, barrier, q):
    with quiet_stdout():
        import oracle.ref as R
        R.use_perf_build()
        R.lib()                     # loads oracle/_ref before the barrier
        barrier.wait()
        q.put(_cpu_reference_run(n, steps, warmup))


def cpu_reference_run_all_cores(n, steps, warmup):
    """The reference's CPU path on every host core: its MPI build is absent here (and its serial build is single-threaded),
    so P independent serial replicas of the n^3 sample run concurrently, one per core — an upper bound on what the MPI
    decomposition of one domain could reach (no halo exchange, no transposes), sharing the host's memory bandwidth.
    Returns (aggregate grid-point-steps/s from the slowest replica, seconds per step, P)."""
    import multiprocessing as mp
    import oracle.ref as R
    R.use_perf_build()
    with quiet_stdout():
        R.lib()                     # also in the parent: the driver's loaded-library record then lists liboracle_ref*.so
    P = max(1, min(len(os.sched_getaffinity(0)), 64))
    if P == 1:
        v, sps = cpu_reference_run(n, steps, warmup)
        return v, sps, 1
    ctx = mp.get_context("fork")
    barrier, q = ctx.Barrier(P), ctx.Queue()
    procs = [ctx.Process(target=_cpu_worker, args=(n, steps, warmup, barrier, q)) for _ in range(P)]
    for p_ in procs:
        p_.start()
    res = [q.get(timeout=3600) for _ in range(P)]
    for p_ in procs:
        p_.join()
    sps = max(r[1] for r in res)
    return P * n ** 3 / sps, sps, P


WORKLOAD = ("cases/gabls1 LES {shape} (advec_2, smag2, thermo dry, MO surface layer on device, swbuffer=1, swlspres=geo, pres_2, RK3; "
            "Timedep surface temperature left out), {n}^3 grid points per GPU, y-slabs")


class HostDrain:
    """Device -> host return of the prognostic fields for a host model that keeps its state in (pinned) host memory: the
    download of step n runs on its own copy stream while step n+1 computes (two device staging slots)."""

    def __init__(self, model, pinned_out):
        import torch
        self.model, self.torch, self.out = model, torch, pinned_out
        f = model.fields
        self.stage = [{n: torch.empty_like(fld.fld) for n, fld in f.ap.items()} for _ in range(2)]
        self.stream = torch.cuda.Stream(device=model.grid.device)
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self.n = 0

    def submit(self):
        torch = self.torch
        slot = self.n % 2
        cur = torch.cuda.current_stream(self.model.grid.device)
        if self.n >= 2:
            cur.wait_event(self.done[slot])
        for n, fld in self.model.fields.ap.items():
            self.stage[slot][n].copy_(fld.fld, non_blocking=True)
        self.ready[slot].record(cur)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.ready[slot])
            for n, dst in self.out.items():
                dst.copy_(self.stage[slot][n], non_blocking=True)
            self.done[slot].record(self.stream)
        self.n += 1

    def finish(self):
        self.stream.synchronize()


def secondary_benches(dev, steps=3):
    """Driver-visible numbers for the other BASELINE configs on one GPU (ms per full step, CUDA events, after 2 warm-up steps):
    drycblles 64^3 (configs[0], launch-bound), moser180 256x192x128 and a 512x512x256 channel (configs[3]: advec_4, diff_4,
    pres_4, RK3, tanh-stretched z, uflux forcing)."""
    import torch
    from microhh_b200.model import Model
    out = {}

    def run(tag, ini, z, init, mo=False, profs=None):
        m = Model(ini, z, device=dev, mo=mo, profs=profs)
        g, f = m.grid, m.fields
        for n, a in init(g).items():
            f.ap[n].fld.copy_(torch.from_numpy(a).to(dev))
        for _ in range(2):
            m.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = g.launch_count()
        e0.record()
        for _ in range(steps):
            m.step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        pts = g.itot * g.jtot * g.ktot
        st = m.print_status()
        out[tag] = {"ms_per_step": round(ms, 4), "grid_point_steps_per_s": round(pts / (ms * 1e-3), 1),
                    "launches_per_step": (g.launch_count() - l0) // steps, "div": st["div"], "cfl": st["cfl"]}
        g.close()

    def dns_ini(itot, jtot, ktot):
        return {"grid": dict(itot=itot, jtot=jtot, ktot=ktot, xsize=6.28, ysize=3.14, zsize=2., swspatialorder="4"),
                "advec": dict(swadvec="4", cflmax=1.0), "diff": dict(swdiff="4", dnmax=0.4), "pres": dict(swpres="4"),
                "boundary": dict(mbcbot="noslip", mbctop="noslip"), "fields": dict(visc=1e-5),
                "force": dict(swlspres="uflux", uflux=0.0282), "time": dict(rkorder=3, dt=1e-3, dtmax=0.1, adaptivestep=True)}

    def dns_z(ktot, alpha=2.0):
        zh = 0.5 * 2. * (1. + np.tanh(alpha * (2. * np.arange(ktot + 1) / ktot - 1.)) / np.tanh(alpha))
        return 0.5 * (zh[1:] + zh[:-1])

    def dns_init(z):
        def init(g):
            rng = np.random.default_rng(1)
            zf = np.zeros(g.kcells); zf[g.kstart:g.kend] = z
            r = {}
            for n, amp, mean in (("u", 0.01, 0.11), ("v", 0.01, 0.), ("w", 0.01, 0.)):
                a = (zf * (2. - zf))[:, None, None] * mean + amp * rng.uniform(-1, 1, g.shape)
                if n == "w":
                    a[:g.kstart + 1] = 0.; a[g.kend:] = 0.
                r[n] = a
            return r
        return init

    for tag, dims in (("dns4_256x192x128", (256, 192, 128)), ("dns4_512x512x256", (512, 512, 256))):
        z = dns_z(dims[2])
        run(tag, dns_ini(*dims), z, dns_init(z))
    n = 64
    ini = les_ini(n, n, n)
    z = (np.arange(n) + 0.5) * (ini["grid"]["zsize"] / n)
    run("les_gabls1_64x64x64", ini, z, lambda g: synthetic_fields(g.shape, g.kstart, g.kend, z, seed=5), mo=True, profs=les_profs(z))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--size", type=int, default=512, help="itot = jtot(per GPU) = ktot")
    ap.add_argument("--cpu-size", type=int, default=None, help="edge of the CPU sample grid (default 96 for cpu_baseline, 192 for --impl reference)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="experiments only: skip the host-resident leg (e2e value null)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.size

    if args.impl == "reference":
        if rank != 0:
            return
        # DRAM-resident sample (192^3: 57 MB per field, far beyond a core's cache share, like 512^3); at most 1 warm-up and
        # 5 timed steps per replica so that the arm ends within a few minutes whatever --steps says
        ns = args.cpu_size or 192
        steps = max(1, min(args.steps, 5))
        value, s_per_step, cores = cpu_reference_run_all_cores(ns, steps, min(args.warmup, 1))
        import oracle.ref as R
        line = {"impl": "reference", "metric": "grid-point-steps/s (full RK3 step)", "value": value, "unit": "grid-point-steps/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * s_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": WORKLOAD.format(shape=f"{n}^3", n=n), "sample": f"{ns}^3 per core",
                           "timed_steps_per_replica": steps,
                           "differences_from_the_gpu_arm": f"{ns}^3 grid per core instead of one {n}^3 grid (the reference's multi-core mode is "
                                                           "MPI, absent here), FFTW replaced by the oracle's DFT shim"},
                "cpu_baseline": {"value": value, "unit": "grid-point-steps/s", "cores": cores, "kind": "reference",
                                 "build": os.path.basename(R.LIB_PATH) + " (g++ -O3 -march=x86-64-v3; the reference's default is -O3 -march=native)",
                                 "sample": f"{steps} full RK3 steps of the same gabls1 configuration on a {ns}^3 grid per core, "
                                           f"{cores} concurrent serial replicas (the reference's MPI build is absent; its serial build is "
                                           "single-threaded), reference sources compiled in place (oracle/_ref), FFTW replaced by the "
                                           "oracle's DFT shim"},
                "e2e": {"value": value, "unit": "grid-point-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    from microhh_b200 import lib as L
    from microhh_b200.model import Model

    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    if world > 1:
        from microhh_b200.multigpu import nccl_env_defaults
        nccl_env_defaults()
        dist.init_process_group("nccl", device_id=dev)

    # ---- multi-GPU values, driver-visible: the slab decomposition against one GPU on a small case, before anything is timed
    slab_parity = None
    if world > 1:
        from tests.slab_parity import parity_check
        slab_parity = parity_check(rank, world, dev, dims=(64, 64 * world, 32))
        dist.barrier()

    itot, jtot, ktot = weak_scaling_shape(n, world)
    ini = les_ini(itot, jtot, ktot)
    if os.environ.get("MHH_FUSE_FORCE") == "0":     # experiments: sponge / Coriolis as separate passes
        ini["b200"] = dict(fuse_force=False)
    z = (np.arange(ktot) + 0.5) * (ini["grid"]["zsize"] / ktot)
    profs = les_profs(z)
    if world > 1:
        from microhh_b200.multigpu import SlabModel
        model = SlabModel(ini, z, device=dev, rank=rank, world=world, profs=profs, mo=True)
    else:
        model = Model(ini, z, device=dev, profs=profs, mo=True)
    g = model.grid
    host = synthetic_fields(g.shape, g.kstart, g.kend, z, seed=666 + rank)
    f = model.fields
    # pinned host staging of the prognostic fields (e2e leg copies from / to here every step)
    pinned = {k: torch.from_numpy(v).pin_memory() for k, v in host.items()}
    for k in pinned:
        f.ap[k].fld.copy_(pinned[k], non_blocking=True)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        model.step()
    barrier()

    # ---- timed region: exactly K steps, CUDA events on the launch stream, per-kernel events live ----------
    sampler = ClockSampler(local_rank)
    sampler.start()
    g.profile_enable(True)
    launches0 = g.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        model.step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    slab_timing = model.pres.timing_report() if getattr(getattr(model, "pres", None), "timing", None) is not None else None
    launches = g.launch_count() - launches0
    prof = g.profile_report()
    g.profile_enable(False)
    sampler.stop_flag = True
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    points = n * n * n * world
    value = points * args.steps / (ms * 1e-3)
    status_timed = model.print_status()

    # ---- e2e: the same step for a host model that keeps its state in HOST memory (the reference's CPU build): every step
    # uploads the four prognostic fields from pinned memory, runs the step through the public API (Model.step), and
    # downloads the four updated fields plus the status line.  Upload of step n+1 and download of step n-1 overlap the
    # compute of step n on two copy streams; every copy, including the first upload and the last download, is timed.
    h2d = sum(v.numel() * 8 for v in pinned.values())
    from microhh_b200.model import HostFeed
    e2e_steps = 0 if args.no_e2e else max(2, min(args.steps, 10))
    pinned_out = {k: torch.empty_like(v).pin_memory() for k, v in pinned.items()}
    feed, drain = HostFeed(model), HostDrain(model, pinned_out)
    barrier()
    ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ee0.record()
    status = None
    if e2e_steps:
        feed.submit(pinned)
    for s_ in range(e2e_steps):
        feed.adopt()
        if s_ + 1 < e2e_steps:
            feed.submit(pinned)
        model.timeloop.dt, model.timeloop.idt = 1.0, int(1e9)
        model.step()
        drain.submit()
        status = model.print_status()        # device reductions + D2H of 6 scalars
    drain.finish()
    ee1.record()
    barrier()
    ems = ee0.elapsed_time(ee1)
    t = torch.tensor([ems], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = points * e2e_steps / (float(t.item()) * 1e-3) if e2e_steps else None
    del feed, drain, pinned_out

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (largest share of the timed region) --------------------------------
    peak, peak_kind = measured_hbm_peak()
    shares = sorted(((v[0], k, v[1]) for k, v in prof.items()), reverse=True)
    total_kernel_ms = sum(v[0] for v in prof.values())
    dom = next((s for s in shares if s[1] in BYTES_PER_POINT), shares[0])
    dom_ms, dom_name, dom_count = dom
    pts_local = n * n * n
    bpp = dict(BYTES_PER_POINT)
    if getattr(model, "lazy_tend", False):
        # lazy tendencies: per RK3 step one of the three launches of these kernels skips the tendency arrays —
        # momentum 64/88/88, output+RK 104/104/80, scalar 56/64/56 B per point: the per-launch AVERAGE is what is timed
        bpp.update({"advec2_smag2_uvw_kernel": 80, "pres2_out_rk_kernel": 96, "advec2_smag2_c_kernel": 58.67})
    algo_bytes = bpp.get(dom_name, 0) * pts_local
    achieved = algo_bytes / (dom_ms / dom_count * 1e-3) / 1e9 if dom_count else 0.
    survey_bytes = SURVEY_BYTES_PER_POINT.get(dom_name, 0) * pts_local
    kernels = {k: {"ms_per_launch": round(v[0] / v[1], 4), "launches": v[1], "share": round(v[0] / total_kernel_ms, 4),
                   "gbs": round(bpp[k] * pts_local / (v[0] / v[1] * 1e-3) / 1e9, 1) if k in bpp else None}
               for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    step_s = ms / args.steps * 1e-3

    line = {
        "metric": "grid-point-steps/s (full RK3 step)", "value": value, "unit": "grid-point-steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD.format(shape=f"{itot}x{jtot}x{ktot}", n=n),
                   "l2": "inputs larger than L2 (each 3-D field 1.09 GB)",
                   "bytes_per_gridpoint_step_moved": BYTES_PER_STEP_FUSED, "bytes_per_gridpoint_step_survey_8d": BYTES_PER_STEP},
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_value, "unit": "grid-point-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": h2d + 6 * 8,
                "steps": e2e_steps,
                "host_model": "state in pinned host memory: 4 prognostic fields up and back down every step (double-buffered on two "
                              "copy streams) + the status line; PCIe-bound.  A device-resident host (the reference's USECUDA build, "
                              "src/model.cxx:561-579) moves only the status line and runs at `value`."},
        "gpu_launches": int(launches),
        # the fraction of the measured HBM peak on the bytes the fused passes actually move (headline), and — labelled — the
        # same time expressed in the reference's operator-by-operator traffic of SURVEY 8(d) (an equivalent-work speed-up)
        "step_hbm_frac": round(BYTES_PER_STEP_FUSED * pts_local / step_s / 1e9 / peak, 4),
        "step_equivalent_reference_traffic_frac": round(BYTES_PER_STEP * pts_local / step_s / 1e9 / peak, 4),
        "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": round(achieved, 1), "peak": peak, "peak_kind": peak_kind,
                     "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "frac_survey": round(survey_bytes / (dom_ms / dom_count * 1e-3) / 1e9 / peak, 4) if dom_count else None,
                     "frac_survey_note": "the same launch time on SURVEY 8(d)'s unfused rows (advec uvw 72 + diff_smag2 uvw 80 B/pt)",
                     "traffic": ncu_traffic(dom_name, n) if world == 1 else None,
                     "traffic_unit": "GB per launch (ncu dram read+write; the captured launch is one that reads the tendencies, "
                                     "88 B/pt = 11.81 GB algorithmic)",
                     "algorithmic_gb_per_launch": round(algo_bytes / 1e9, 3)},
        "kernels": kernels,
        "status": status_timed,
    }
    if slab_parity is not None:
        line["slab_parity"] = slab_parity
        line["config"]["slab_path"] = getattr(model, "path", None)
        line["config"]["slab_a2a"] = getattr(model.pres, "a2a", None)
        if slab_timing:
            line["slab_exchange_timing"] = slab_timing
    if world == 1 and not args.no_secondary:
        del model
        torch.cuda.empty_cache()
        line["secondary"] = secondary_benches(dev)
    if not args.no_cpu_baseline and world == 1:
        ns = args.cpu_size or 96
        v, s_per_step = cpu_reference_run(ns, 3, 1)
        line["cpu_baseline"] = {"value": v, "unit": "grid-point-steps/s", "cores": 1, "kind": "reference",
                                "sample": f"3 full RK3 steps of the same gabls1 configuration on a {ns}^3 grid "
                                          "(oracle/_ref timing build, -O3 -march=x86-64-v3)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
