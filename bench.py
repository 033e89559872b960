#!/usr/bin/env python
"""bench.py — grid-point-steps/s of a full RK3 LES step (BASELINE.json metric) on N B200s.

Workload (config.workload): gabls1-like LES (advec_2, diff_smag2 with surface-model rows, thermo dry,
pres_2, RK3) scaled to a 512^3 synthetic grid per GPU (BASELINE.json configs[2]); for N > 1 the domain
is itot x (N*jtot) x ktot split into y-slabs (weak scaling).
One "step" = 3 RK substeps, each = cyclic + ghost BCs, evisc, buoyancy, advection, diffusion,
pressure solve, RK update; the CFL/diffusion-number maxima once per step.

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
BYTES_PER_STEP = 3 * 608 + 32       # 1856 B / grid-point-step (the reference's operator-by-operator passes)
# what this implementation's fused passes move: evisc 40, momentum 88, scalar + its RK 64, pres_in 56, 4 FFT passes 64,
# tdma 28 (two sweeps over p + pivot checkpoints), output+RK 104 per substep; the cfl / dnmul maxima ride in the evisc pass;
# lazy tendencies: the last substep writes no ut,vt,wt (24) and no zeroed tht (8), the first reads none of them (32)
BYTES_PER_STEP_FUSED = 3 * 444 - 64


def ncu_traffic(kernel, n):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (GB) from the committed ncu --set full capture of this
    workload (profiles/r01_traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
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
        "boundary": dict(mbcbot="noslip", mbctop="freeslip", swboundary="surface",
                         sbc={"th": dict(sbcbot="dirichlet", sbctop="neumann", sbot=265., stop=0.01)}),
        "fields": dict(visc=1e-5, svisc=1e-5),
        "time": dict(rkorder=3, dt=1.0, dtmax=60., adaptivestep=True),
    }


def synthetic_fields(shape, ks, ke, z, seed, jslice=None):
    """u,v ~ U(-1,1)+8, w ~ 0.2 U(-1,1) (0 at walls), th = 263.5 + 0.003 z + U(-0.05,0.05)."""
    rng = np.random.default_rng(seed)
    kc = shape[0]
    zf = np.zeros(kc)
    zf[ks:ke] = z
    u = 8. + rng.uniform(-1, 1, shape)
    v = rng.uniform(-1, 1, shape)
    w = 0.2 * rng.uniform(-1, 1, shape)
    w[:ks + 1] = 0.
    w[ke:] = 0.
    th = 263.5 + 0.003 * zf[:, None, None] + rng.uniform(-0.05, 0.05, shape)
    return dict(u=u, v=v, w=w, th=th)


def surface_planes(shape2d, rng):
    return dict(z0m=np.full(shape2d, 0.1), dudz=rng.uniform(0.05, 0.1, shape2d), dvdz=rng.uniform(-0.01, 0.01, shape2d),
                dbdz=rng.uniform(-1e-5, 1e-5, shape2d), ufluxbot=rng.uniform(-0.02, 0., shape2d),
                vfluxbot=rng.uniform(-0.002, 0.002, shape2d), sfluxbot=rng.uniform(-0.01, 0., shape2d))


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
    """The reference's own CPU code (oracle/_ref, 1 thread: its non-MPI build is single-threaded,
    src/model.cxx:325) on a bounded n^3 sample of the same LES configuration."""
    from oracle.les import OracleLES
    ini = les_ini(n, n, n)
    g = ini["grid"]
    z = (np.arange(n) + 0.5) * (g["zsize"] / n)
    shape = (n + 2, n + 2, n + 2)
    f = synthetic_fields(shape, 1, n + 1, z, seed=666)
    S = surface_planes(shape[1:], np.random.default_rng(7))
    S.update(ufluxtop=np.zeros(shape[1:]), vfluxtop=np.zeros(shape[1:]), sfluxtop=np.zeros(shape[1:]))
    o = OracleLES(n, n, n, g["xsize"], g["ysize"], g["zsize"], z, f, swadvec="2", swdiff="smag2", visc=1e-5, svisc=1e-5,
                  cs=0.1, tPr=1. / 3., thref0=263.5, surface=S, mbcbot="noslip", mbctop="freeslip",
                  sbc=ini["boundary"]["sbc"], dt=1.0, dtmax=60., dnmax=0.3)
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
        from oracle.les import OracleLES  # noqa: F401  (loads oracle/_ref before the barrier)
        barrier.wait()
        q.put(_cpu_reference_run(n, steps, warmup))


def cpu_reference_run_all_cores(n, steps, warmup):
    """The reference's CPU path on every host core: its MPI build is absent here (and its serial build is single-threaded),
    so P independent serial replicas of the n^3 sample run concurrently, one per core — an upper bound on what the MPI
    decomposition of one domain could reach (no halo exchange, no transposes), sharing the host's memory bandwidth.
    Returns (aggregate grid-point-steps/s from the slowest replica, seconds per step, P)."""
    import multiprocessing as mp
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--size", type=int, default=512, help="itot = jtot(per GPU) = ktot")
    ap.add_argument("--cpu-size", type=int, default=96)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.size

    if args.impl == "reference":
        if rank != 0:
            return
        ns = args.cpu_size
        value, s_per_step, cores = cpu_reference_run_all_cores(ns, max(args.steps, 1), min(args.warmup, 1))
        line = {"impl": "reference", "metric": "grid-point-steps/s (full RK3 step)", "value": value, "unit": "grid-point-steps/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * s_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"gabls1-like LES {n}^3 per GPU (advec_2, smag2+surface rows, thermo dry, pres_2, RK3)",
                           "sample": f"{ns}^3"},
                "cpu_baseline": {"value": value, "unit": "grid-point-steps/s", "cores": cores, "kind": "reference",
                                 "sample": f"{args.steps} full RK3 steps of the same LES configuration on a {ns}^3 grid per core, "
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

    itot, jtot, ktot = weak_scaling_shape(n, world)
    ini = les_ini(itot, jtot, ktot)
    z = (np.arange(ktot) + 0.5) * (ini["grid"]["zsize"] / ktot)
    if world > 1:
        from microhh_b200.multigpu import SlabModel
        model = SlabModel(ini, z, device=dev, rank=rank, world=world)
    else:
        model = Model(ini, z, device=dev)
    g = model.grid
    host = synthetic_fields(g.shape, g.kstart, g.kend, z, seed=666 + rank)
    S = surface_planes(g.shape2d, np.random.default_rng(7 + rank))
    model.surface = {k: torch.from_numpy(v).to(dev) for k, v in S.items()}
    f = model.fields
    f.mp["u"].flux_bot.copy_(model.surface["ufluxbot"]); f.mp["v"].flux_bot.copy_(model.surface["vfluxbot"])
    f.sp["th"].flux_bot.copy_(model.surface["sfluxbot"])
    # pinned host staging of the prognostic fields (e2e leg copies from here every step)
    pinned = {k: torch.from_numpy(v).pin_memory() for k, v in host.items()}
    prog = {"u": f.mp["u"].fld, "v": f.mp["v"].fld, "w": f.mp["w"].fld, "th": f.sp["th"].fld}
    for k in prog:
        prog[k].copy_(pinned[k], non_blocking=True)
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

    # ---- e2e: same step through the public API with HOST buffers: H2D of the prognostic fields from pinned
    # memory every step, D2H of the step's status line (cfl, dn, div, mom, tke, mass) -----------------------
    h2d = sum(v.numel() * 8 for v in pinned.values())
    # The upload of step n+1 (copy stream, second staging slot) overlaps the compute of step n (model.HostFeed); every
    # step's upload, including the first, lies inside the timed region.
    from microhh_b200.model import HostFeed
    e2e_steps = max(2, min(args.steps, 10))
    feed = HostFeed(model)
    barrier()
    ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ee0.record()
    status = None
    feed.submit(pinned)
    for s_ in range(e2e_steps):
        feed.adopt()
        if s_ + 1 < e2e_steps:
            feed.submit(pinned)
        model.timeloop.dt, model.timeloop.idt = 1.0, int(1e9)
        model.step()
        status = model.print_status()        # device reductions + D2H of 6 scalars
    ee1.record()
    barrier()
    ems = ee0.elapsed_time(ee1)
    t = torch.tensor([ems], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = points * e2e_steps / (float(t.item()) * 1e-3)

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
    kernels = {k: {"ms_per_launch": round(v[0] / v[1], 4), "launches": v[1], "share": round(v[0] / total_kernel_ms, 4),
                   "gbs": round(bpp[k] * pts_local / (v[0] / v[1] * 1e-3) / 1e9, 1) if k in bpp else None}
               for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}

    line = {
        "metric": "grid-point-steps/s (full RK3 step)", "value": value, "unit": "grid-point-steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"gabls1-like LES {itot}x{jtot}x{ktot} (advec_2, smag2+surface rows, thermo dry, pres_2, RK3), "
                               f"{n}^3 grid points per GPU, y-slabs", "l2": "inputs larger than L2 (each 3-D field 1.09 GB)",
                   "bytes_per_gridpoint_step": BYTES_PER_STEP},
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_value, "unit": "grid-point-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 6 * 8,
                "steps": e2e_steps},
        "gpu_launches": int(launches),
        "step_hbm_frac": round(BYTES_PER_STEP * pts_local / (ms / args.steps * 1e-3) / 1e9 / peak, 4),
        "step_hbm_frac_fused_passes": round(BYTES_PER_STEP_FUSED * pts_local / (ms / args.steps * 1e-3) / 1e9 / peak, 4),
        "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": round(achieved, 1), "peak": peak, "peak_kind": peak_kind,
                     "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": ncu_traffic(dom_name, n) if world == 1 else None,
                     "traffic_unit": "GB per launch (ncu dram read+write; the captured launch is one that reads the tendencies, "
                                     "88 B/pt = 11.81 GB algorithmic)",
                     "algorithmic_gb_per_launch": round(algo_bytes / 1e9, 3)},
        "kernels": kernels,
        "status": status,
    }
    if not args.no_cpu_baseline and world == 1:
        ns = args.cpu_size
        v, s_per_step = cpu_reference_run(ns, 3, 1)
        line["cpu_baseline"] = {"value": v, "unit": "grid-point-steps/s", "cores": 1, "kind": "reference",
                                "sample": f"3 full RK3 steps of the same LES configuration on a {ns}^3 grid (oracle/_ref)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
