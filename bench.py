#!/usr/bin/env python
"""bench.py -- spin-flip attempts/s of the Metropolis sweep hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--layout int8|bits]

One "step" = one full checkerboard sweep of the workload (two kernel launches, one per colour).
N = 1: BASELINE.json configs[1], a single 16384 x 16384 Ising lattice (2^28 sites) at kT = T_c.
N > 1: the replica-exchange workload -- 256 replicas of 1024^2 per GPU (again 2^28 sites per GPU, weak
scaling), one temperature ladder over all GPUs, energies all-gathered with NCCL per exchange.
``--workload`` selects others: ``slab`` (one 16384^2 lattice split by rows, strong scaling), ``ising16k``,
``latgas``, ``discs``.
Prints ONE JSON line on rank 0 (contract in the task statement / DESIGN.md section "Measurement").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "spin-flip attempts/s"
UNIT = "attempts/s"
TC = 2.269185314213022
SEED = 20260611


def measured_peak_hbm():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def profiled_traffic(kernel_name, nsites_expected):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    ncu --set full summary of this workload (profiles/); None if there is no capture for this kernel/size."""
    import csv
    table = {"sweep_fused_i8_kernel": ("r1_fused_final_summary.csv", 268435456),
             "sweep_bits_rows_kernel": ("r1_bits_final_summary.csv", 268435456),
             # 256 replicas of 1024^2 int8, 10 sweeps per launch (the default replica-exchange workload)
             "resident_kernel": ("r1_resident_rx_summary.csv", 268435456)}
    if kernel_name not in table or table[kernel_name][1] != nsites_expected:
        return None, None
    path = os.path.join(ROOT, "profiles", table[kernel_name][0])
    try:
        vals = {}
        for row in csv.reader(open(path)):
            if row and row[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[row[1]]
                nums = [float(v.replace(",", "")) * scale for v in row[2:] if v]
                vals[row[0]] = sum(nums) / len(nums)
        return vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"], "profiles/" + table[kernel_name][0]
    except Exception:
        return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's MCstep loop (oracle/cpu_port.py), one replica per host core
# ------------------------------------------------------------------------------------------------
def cpu_baseline(seconds_target=12.0):
    from oracle import cpu_port
    cores = os.cpu_count() or 1
    # size the sample from a short calibration so that it is about seconds_target of CPU work per core
    calib, _ = cpu_port.time_all_cores(1024, -1.0, TC, 100000, 1)
    ntrials = int(max(calib, 5e4) * seconds_target)
    value, wall = cpu_port.time_all_cores(1024, -1.0, TC, ntrials, cores)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "wall_s": wall,
            "sample": "%d single-site MCstep trials per core on a 1024x1024 float64 lattice at kT=Tc, one replica "
                      "per core (%d cores, %.1f s); per-trial cost of the reference loop does not depend on "
                      "lattice size" % (ntrials, cores, wall)}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    # each "step" of this arm is a bounded sample of the same workload: cap the count so the run ends in minutes
    args.steps = min(args.steps, 4)
    args.warmup = min(args.warmup, 1)
    t_per_step = max(2.0, 20.0 / max(1, args.steps + args.warmup))
    cb = None
    vals, walls = [], []
    for i in range(args.warmup + args.steps):
        c = cpu_baseline(seconds_target=t_per_step)
        if i >= args.warmup:
            vals.append(c["value"])
            walls.append(c["wall_s"])
            cb = c
    value = sum(vals) / len(vals)
    cb["value"] = value
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(walls) / len(walls),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args)}, "cpu_baseline": cb,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


KERNELS = {"auto": 0, "generic": 1, "two_pass": 2, "resident": 3}  # PYMC_KERNEL_* of include/pymc_b200.h


def pick_workload(args, world):
    if args.workload != "auto":
        return args.workload
    return "ising16k" if world == 1 else "rx"


def workload_name(args, world=None):
    world = args.gpus if world is None else world
    if pick_workload(args, world) == "rx":
        name = ("ising2d" if args.rx_model == "ising" else "hardcore_lattice_gas_mu=%g" % args.rx_mu)
        return ("%s_replica_exchange_%d_replicas_of_%dx%d_per_gpu_kT=%.2f..%.2f_exchange_every_%d_sweeps"
                % (name, args.rx_replicas, args.rx_L, args.rx_L, args.rx_kT_lo, args.rx_kT_hi, args.rx_every))
    return "ising2d_%dx%d_single_replica_checkerboard_kT=%s_J=-1" % (args.L, args.L, "Tc" if args.kT == TC else "%g" % args.kT)


# ------------------------------------------------------------------------------------------------
def run_rx(args):
    """Replica exchange, weak scaling: every GPU holds rx_replicas lattices of rx_L^2; the ladder spans all
    replicas of all GPUs; energies are all-gathered with NCCL (in place, on the kernels' stream) and the
    exchange is decided on every GPU by the same kernel.  No host synchronisation in the timed region."""
    import numpy as np
    import torch
    rank, local_rank, world = dist_env()
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    from py_mc_b200.mc_mpi import _DeviceArray
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L, nloc = args.rx_L, args.rx_replicas
    R = nloc * world
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    kTs = np.linspace(args.rx_kT_lo, args.rx_kT_hi, R)
    if args.rx_model == "hardcore":
        # BASELINE config 5 read as a lattice model: nearest-neighbour exclusion, ladder over kT at fixed mu
        # (e.g. --rx-model hardcore --rx-replicas 8 on 8 GPUs = 64 temperatures)
        from py_mc_b200.latgas import hardcore_latticegas
        model = hardcore_latticegas(mu=args.rx_mu)
    else:
        model = ising2D(-1.0)
    lat = _cabi.Lattices(model._pmc_model, L, L, nreplicas=nloc, ntemps=R, seed=SEED, layout=layout,
                         device=local_rank, replica_offset=rank * nloc, nreplicas_global=R, kernel=KERNELS[args.kernel])
    stream = torch.cuda.current_stream()
    lat.set_stream(stream.cuda_stream)
    lat.init_random()
    lat.set_accept_tables(np.array([model._pmc_thresholds(kT) for kT in kTs]))
    lat.rx_setup(1.0 / kTs, *model._pmc_energy_coeffs())
    eptr, _ = lat.rx_buffers()
    e_all = torch.as_tensor(_DeviceArray(eptr, R, "<f8"), device="cuda")
    e_mine = e_all[rank * nloc:(rank + 1) * nloc]
    state = {"scheme": 0, "done": 0, "exchanges": 0, "calls": 0}

    def steps(n):
        while n > 0:  # the sweeps up to the next exchange are one call (one launch on the resident kernel)
            chunk = min(n, args.rx_every - state["done"] % args.rx_every)
            lat.sweep(chunk)
            state["calls"] += 1
            state["done"] += chunk
            n -= chunk
            if state["done"] % args.rx_every == 0:
                lat.rx_energies()
                if dist is not None:
                    dist.all_gather_into_tensor(e_all, e_mine)
                lat.rx_exchange(state["scheme"])
                state["scheme"] ^= 1
                state["exchanges"] += 1

    steps(args.warmup)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ex0, calls0 = state["exchanges"], state["calls"]
    ev0.record(stream)
    steps(args.steps)
    ex1, ncalls = state["exchanges"], state["calls"] - calls0
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    if dist is not None:
        dist.barrier()
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if ms < 400.0:
        # keep the same loop running (untimed) until the sampler has seen ~0.4 s under load; the count comes
        # from the all-reduced time so that every rank issues the same number of collectives
        steps(int(args.steps * (400.0 - ms) / max(ms, 1e-3)) + 1)
        torch.cuda.synchronize()
    clocks = sampler.stop()
    clocks["window"] = "timed region" if ms >= 400.0 else "timed region + untimed repeat of the same loop (0.4 s in all)"
    nsites = nloc * L * L
    value = world * nsites * args.steps / (ms * 1e-3)
    nex = ex1 - ex0

    # end to end: host configurations in, sweeps + exchanges, observables and the map out, every step
    host = [lat.download(r) for r in range(min(nloc, 4))]
    img = np.stack([lat.download(r) for r in range(nloc)]) if args.e2e_full else None
    reps = 3
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for r in range(nloc):
            lat.upload(r, img[r] if img is not None else host[r % len(host)])
        steps(args.e2e_sweeps)
        b, m = lat.tracked()
        mp = lat.get_replica_to_T()
    t_e2e = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([t_e2e], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    e2e_value = world * nsites * args.e2e_sweeps * reps / t_e2e
    tried, acc = lat.rx_stats()

    bytes_per_site = 1.0 if layout == _cabi.LAYOUT_INT8 else 0.125
    lps, kname = lat.sweep_kernel_info()
    nlaunch = ncalls if kname == "resident_kernel" else lps * args.steps  # resident: one launch per call
    # algorithmic bytes: 2 x bytes_per_site per attempt, whatever the kernel really moves
    alg = 2.0 * bytes_per_site * nsites * args.steps / nlaunch
    launch_ms = ms / nlaunch  # exchange kernels are ~microseconds; stated in config
    peak, peak_src = measured_peak_hbm()
    achieved = alg / (launch_ms * 1e-3) / 1e9
    traffic = traffic_src = None
    if (kname == "resident_kernel" and layout == _cabi.LAYOUT_INT8 and args.rx_model == "ising" and (L, nloc) == (1024, 256)
            and args.rx_every == 10):  # the capture is of exactly this launch shape
        traffic, traffic_src = profiled_traffic(kname, nsites)
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8" if layout == _cabi.LAYOUT_INT8 else "u64 (1 bit/site)",
            "data": "synthetic",
            "config": {"workload": workload_name(args, world), "layout": args.layout, "sites_per_gpu": nsites,
                       "replicas_per_gpu": nloc, "replicas_total": R, "lattice": [L, L],
                       "step": "one sweep of every replica (2 colour passes) + exchange when due",
                       "exchanges_in_timed_region": nex, "seed": SEED,
                       "rx_sweeps_per_s": world * nloc * args.steps / (ms * 1e-3),
                       "exchange_acceptance": float(acc.sum()) / max(1.0, float(tried.sum())),
                       "l2": "working set %d MiB per GPU %s L2" % (int(nsites * bytes_per_site) >> 20,
                                                                  "exceeds" if nsites * bytes_per_site > 126e6 else "fits in"),
                       "parallelism": "replicas block-partitioned over %d GPU(s); NCCL all-gather of %d energies per "
                                      "exchange" % (world, R)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nsites,
                    "d2h_bytes_per_step": 16 * nloc + 4 * R, "sweeps_per_step": args.e2e_sweeps,
                    "note": "C-ABI: upload every replica from host int8, sweeps+exchanges, read back tracked "
                            "observables and the replica->T map"},
            "gpu_launches": nlaunch + 2 * nex,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "kernel": kname, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg},
        }
        if kname == "resident_kernel":
            line["roofline"]["note"] = ("not HBM-bound: the lattice stays in shared memory for the %d sweeps of a launch; "
                                        "HBM moves 2 x bytes_per_site per launch, not per sweep" % args.rx_every)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_latgas(args):
    """BASELINE config 4: pair-interaction lattice gas, 512 batched replicas of 256 x 256 (size not given by
    BASELINE; stated here), J = -1, Eads - mu = 0, kT = 1.  One step = one flip sweep + one full
    particle-hole exchange sweep of every replica.  Single GPU."""
    import numpy as np
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas_move import latticegas
    L, R = args.lg_L, args.lg_replicas
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    model = latticegas(Eads=0.0, J=-1.0, mu=0.0)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, L, L, nreplicas=R, ntemps=1, seed=SEED, layout=layout,
                         kernel=KERNELS[args.kernel])
    lat.init_random()
    lat.set_accept_tables(model._pmc_thresholds(1.0)[None, :])
    lat.set_pair_tables(model._pmc_pair_thresholds(1.0)[None, :])
    nsites = R * L * L
    out = {}
    for name, fn, attempts in (("flip", lambda n: lat.sweep(n), nsites), ("pair_exchange", lambda n: lat.pair_sweep(n), 2 * nsites)):
        fn(args.warmup)
        lat.synchronize()
        lat.timer_start()
        fn(args.steps)
        ms = lat.timer_stop()
        out[name] = {"value": attempts * args.steps / (ms * 1e-3), "unit": "attempts/s", "ms_per_step": ms / args.steps}
    b, n = lat.tracked()
    b2, n2 = lat.reduce_observables()
    assert np.array_equal(b, b2) and np.array_equal(n, n2)
    lps, kname = lat.sweep_kernel_info()
    line = {"metric": "lattice-gas attempts/s", "value": out["flip"]["value"], "unit": "attempts/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": out["flip"]["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "latgas_%d_replicas_of_%dx%d_J=-1_kT=1" % (R, L, L), "layout": args.layout,
                       "flip_kernel": kname, "seed": SEED},
            "pair_exchange": out["pair_exchange"],
            "gpu_launches": 2 * ((args.steps + 4095) // 4096) if kname == "resident_kernel" else (lps + 16) * args.steps}
    print(json.dumps(line), flush=True)


def run_discs(args):
    """BASELINE config 5 (re-interpreted, SURVEY.md section 8 notes): batched independent hard-disc replicas,
    64 x 128 = 8192 replicas of the shipped system (20 discs, L = 1, r = 1/14, maxstep = r/2)."""
    from py_mc_b200.hardcore import DiscBatch
    R = args.disc_replicas
    b = DiscBatch(R, 1.0, 20, 1.0 / 14.0, 1.0 / 28.0, isotropic=False, seed=SEED)
    per = args.disc_trials
    b._d.timed_run(per)
    ms = [b._d.timed_run(per) for _ in range(max(3, args.steps // 10))]
    ms_best = sorted(ms)[len(ms) // 2]
    assert not b.any_overlap().any()
    line = {"metric": "hard-disc move attempts/s", "value": R * per / (ms_best * 1e-3), "unit": "attempts/s", "n_gpus": 1,
            "steps": len(ms), "warmup": 1, "ms_per_step": ms_best, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "hard_discs_%d_replicas_20_discs_%d_trials_per_launch" % (R, per), "seed": SEED,
                       "acceptance": float(b.acceptance().mean())},
            "gpu_launches": len(ms)}
    print(json.dumps(line), flush=True)


def run_slab(args):
    """Strong scaling of ONE 16384 x 16384 lattice split by rows over the GPUs: halo rows are
    stored into the neighbours' memory by the sweep kernel itself (CUDA IPC peer mappings), one flag barrier
    kernel per sweep, no NCCL in the timed region."""
    import torch
    rank, local_rank, world = dist_env()
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    from py_mc_b200.slab import SlabLattice
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = args.L
    sl = SlabLattice(_cabi.MODEL_ISING, L, L, SEED, dist, device=local_rank)
    sl.init_random()
    sl.set_accept_tables(ising2D(-1.0)._pmc_thresholds(TC)[None, :])
    sl.sweep(args.warmup)
    sl.lat.synchronize()
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    sl.lat.timer_start()
    sl.sweep(args.steps)
    ms = sl.lat.timer_stop()
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if ms < 400.0:  # every slab must issue the same number of sweeps: decide from the all-reduced time
        sl.sweep(int(args.steps * (400.0 - ms) / max(ms, 1e-3)) + 1)
        sl.lat.synchronize()
    clocks = sampler.stop()
    sl.check()
    b, m = sl.observables()
    if rank == 0:
        nsites = L * L
        peak, peak_src = measured_peak_hbm()
        achieved = 2.0 * nsites / world / (ms / args.steps * 1e-3) / 1e9
        print(json.dumps({
            "metric": METRIC, "value": nsites * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "ising2d_%dx%d_one_lattice_in_%d_row_slabs_kT=Tc_J=-1" % (L, L, world),
                       "layout": "int8", "rows_per_gpu": L // world, "seed": SEED,
                       "parallelism": "row slabs; halo rows stored into peer memory by the sweep kernel, flag barrier "
                                      "per sweep, no NCCL in the timed region",
                       "bond_sum": b, "magnetisation": m},
            "clocks": clocks, "gpu_launches": (1 if world == 1 else 2) * args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "kernel": "sweep_fused_i8_kernel", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": 2.0 * nsites / world},
        }), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_ours(args):
    import numpy as np
    rank, local_rank, world = dist_env()
    if args.workload == "slab":
        return run_slab(args)
    if args.workload == "latgas":
        return run_latgas(args)
    if args.workload == "discs":
        return run_discs(args)
    if pick_workload(args, world) == "rx":
        return run_rx(args)
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    L = args.L
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    thr = ising2D(-1.0)._pmc_thresholds(args.kT)[None, :]
    nsites = L * L

    def make(lay):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=1, ntemps=1, seed=SEED, layout=lay,
                             device=local_rank, replica_offset=rank, nreplicas_global=world)
        lat.init_random()
        lat.set_accept_tables(thr)
        return lat

    def barrier():
        if dist is not None:
            dist.barrier()

    def timed(lat, steps, warmup):
        lat.sweep(warmup)
        lat.synchronize()
        barrier()
        lat.timer_start()
        lat.sweep(steps)
        ms = lat.timer_stop()
        barrier()
        return ms

    lat = make(layout)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed(lat, args.steps, args.warmup)
    if ms < 400.0:  # keep the same kernel running (untimed) until the sampler has seen ~0.4 s under load
        lat.sweep(int(args.steps * (400.0 - ms) / max(ms, 1e-3)) + 1)
        lat.synchronize()
    clocks = sampler.stop()
    clocks["window"] = "timed region" if ms >= 400.0 else "timed region + untimed repeat of the same loop (0.4 s in all)"
    if dist is not None:
        import torch
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * nsites * args.steps / (ms * 1e-3)

    # end to end through the C ABI with HOST buffers: upload, sweeps, download, every step
    host = _cabi.pinned_empty((L, L), np.int8)
    lat.download(0, out=host)
    e2e_sweeps = args.e2e_sweeps
    reps = 5
    lat.upload(0, host)
    lat.sweep(1)
    lat.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        lat.upload(0, host)
        lat.sweep(e2e_sweeps)
        lat.download(0, out=host)
        b, m = lat.tracked()
    t_e2e = time.perf_counter() - t0
    if dist is not None:
        import torch
        t = torch.tensor([t_e2e], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    e2e_value = world * nsites * e2e_sweeps * reps / t_e2e

    bytes_per_site = 1.0 if layout == _cabi.LAYOUT_INT8 else 0.125
    lps, kname = lat.sweep_kernel_info()
    persistent = kname == "persistent_kernel"
    # every site a launch updates is read once and written once (the fused kernel updates both colours).  The
    # persistent kernel runs all sweeps of a call (<= 4096) in ONE cooperative launch: its algorithmic bytes are
    # 2 x bytes_per_site per attempt all the same, although DRAM moves the lattice once per launch.
    nlaunch = (args.steps + 4095) // 4096 if persistent else lps * args.steps
    alg_bytes_per_launch = 2.0 * bytes_per_site * nsites * args.steps / nlaunch
    launch_ms = ms / nlaunch
    peak, peak_src = measured_peak_hbm()
    achieved = alg_bytes_per_launch / (launch_ms * 1e-3) / 1e9
    traffic, traffic_src = profiled_traffic(kname, nsites)

    extra = {}
    if args.also_bits and layout == _cabi.LAYOUT_INT8:
        lat.close()
        try:  # an extra, never allowed to take the headline line down with it
            lat2 = make(_cabi.LAYOUT_BITS)
            ms2 = timed(lat2, args.steps, args.warmup)
            extra["bitpacked"] = {"value": world * nsites * args.steps / (ms2 * 1e-3), "unit": UNIT,
                                  "ms_per_step": ms2 / args.steps, "kernel": lat2.sweep_kernel_info()[1],
                                  "note": "same workload on the 1-bit/site layout (32 MiB: not HBM-bound)"}
            lat2.close()
        except Exception as exc:
            extra["bitpacked"] = {"error": repr(exc)}
    if persistent:  # the per-sweep kernel on the same workload, for comparison (one launch per sweep, 2 B/attempt of DRAM)
        try:
            lat3 = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=1, ntemps=1, seed=SEED, layout=layout,
                                  device=local_rank, replica_offset=rank, nreplicas_global=world)
            lat3.init_random()
            lat3.set_accept_tables(thr)
            lat3.sweep(1)
            lat3.synchronize()
            n3 = max(50, args.steps // 10)
            lat3.timer_start()
            for _ in range(n3):
                lat3.sweep(1)  # single-sweep calls take the fused kernel
            ms3 = lat3.timer_stop()
            extra["one_sweep_per_launch"] = {"value": world * nsites * n3 / (ms3 * 1e-3), "unit": UNIT,
                                             "ms_per_step": ms3 / n3, "kernel": "sweep_fused_i8_kernel",
                                             "roofline_frac": 2.0 * bytes_per_site * nsites / (ms3 / n3 * 1e-3) / 1e9 /
                                             measured_peak_hbm()[0]}
            lat3.close()
        except Exception as exc:
            extra["one_sweep_per_launch"] = {"error": repr(exc)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8" if layout == _cabi.LAYOUT_INT8 else "u64 (1 bit/site)",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "layout": args.layout, "sites_per_gpu": nsites,
                       "step": "one full sweep (both colours)", "seed": SEED,
                       "l2": ("lattice (%d MiB in HBM) is loaded once per launch and stays in the shared memory of the SMs "
                              "as bits for all %d sweeps of the timed launch, then is stored once; no cache is relied upon"
                              % (int(nsites * bytes_per_site) >> 20, args.steps)) if persistent else
                             "working set %d MiB %s L2" % (int(nsites * bytes_per_site) >> 20,
                                                           "exceeds" if nsites * bytes_per_site > 126e6 else "fits in"),
                       "parallelism": "replicas only" if world > 1 else "single GPU"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nsites, "d2h_bytes_per_step": nsites + 16,
                    "sweeps_per_step": e2e_sweeps,
                    "note": "C-ABI upload(pinned host int8) + sweeps + download(host int8) + tracked observables per step"},
            "gpu_launches": nlaunch,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "kernel": kname, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes_per_launch,
                         "note": ("not HBM-bound: the lattice stays in the shared memory of the %d SMs for all sweeps of the "
                                  "launch; DRAM moves 2 x bytes_per_site per launch, not per sweep; the integer ALU pipe "
                                  "limits the kernel" % 148) if persistent else
                                 "not HBM-bound: the integer ALU pipe limits this kernel (profiles/README.md)"},
        }
        line.update(extra)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--layout", default="int8", choices=["int8", "bits"])
    ap.add_argument("--L", type=int, default=16384)
    ap.add_argument("--kT", type=float, default=TC, help="temperature of the single-lattice workload (default T_c; "
                    "SURVEY 8d also asks for 2.0 and 5.0)")
    ap.add_argument("--e2e-sweeps", type=int, default=64)
    ap.add_argument("--also-bits", action="store_true", default=True)
    ap.add_argument("--no-also-bits", dest="also_bits", action="store_false")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workload", default="auto", choices=["auto", "ising16k", "rx", "latgas", "discs", "slab"])
    ap.add_argument("--lg-replicas", type=int, default=512)
    ap.add_argument("--lg-L", type=int, default=256)
    ap.add_argument("--disc-replicas", type=int, default=8192)
    ap.add_argument("--disc-trials", type=int, default=2000)
    ap.add_argument("--rx-replicas", type=int, default=256, help="replicas per GPU (weak scaling)")
    ap.add_argument("--rx-L", type=int, default=1024)
    ap.add_argument("--rx-every", type=int, default=10, help="sweeps between exchanges")
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "two_pass", "resident"],
                    help="sweep kernel of the handle (PYMC_KERNEL_*); resident = one block keeps one replica in shared memory")
    ap.add_argument("--rx-model", default="ising", choices=["ising", "hardcore"])
    ap.add_argument("--rx-mu", type=float, default=1.0, help="chemical potential of the hard-core lattice gas")
    ap.add_argument("--rx-kT-lo", type=float, default=2.0)
    ap.add_argument("--rx-kT-hi", type=float, default=2.6)
    ap.add_argument("--e2e-full", action="store_true", help="rx e2e uploads a distinct host image per replica")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
