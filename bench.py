#!/usr/bin/env python
"""bench.py -- spin-flip attempts/s of the Metropolis sweep / replica-exchange hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--layout int8|bits]

One "step" = one full checkerboard sweep of every lattice of the workload (plus the exchange when one is due).

N = 1   BASELINE.json configs[1]: one 16384 x 16384 Ising lattice (2^28 sites) at kT = T_c, int8.  The line also
        carries the same workload bit-packed, one-sweep-per-launch, and an "rx" block: BASELINE configs[2] (256
        replicas of 1024^2 with replica exchange) on this one GPU.
N > 1   BASELINE.json configs[2] as written: 256 replicas of 1024^2 IN TOTAL, block-partitioned over the N GPUs
        (strong scaling), one temperature ladder, energies all-gathered per exchange, decisions on every GPU.  Extra
        keys: "weak" (256 replicas PER GPU), "scaling_check" (rank 0 times the whole 256-replica workload alone, in
        the same job, so that speed-up and efficiency can be read against the same workload), "exchange_us",
        "parity".
``--workload`` selects others: ``slab`` (one 16384^2 lattice split by rows), ``rx``, ``latgas`` (config 4),
``discs`` / ``--rx-model hardcore`` (config 5).  ``--impl reference`` times the UNMODIFIED reference (oracle/_ref) on
the host cores.  Prints ONE JSON line on rank 0 (contract: task statement / DESIGN.md section "Measurement").
"""
import argparse
import json
import math
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
MIN_TIMED_MS = 200.0  # a timed region shorter than this is repeated inside the events ("repeats")

KERNELS = {"auto": 0, "generic": 1, "two_pass": 2, "resident": 3, "persistent": 4, "pbatch": 5, "rowstream": 6}  # PYMC_KERNEL_*


def measured_peak_hbm():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# committed ncu --set full summaries (profiles/): (kernel, workload tag) -> (file, launch shape the capture is of).
# A line only quotes a capture of its own kernel on its own workload.
PROFILED = {
    ("sweep_fused_i8_kernel", "single16k"): ("r1_fused_final_summary.csv", "16384x16384 int8, one sweep per launch"),
    ("sweep_bits_rows_kernel", "single16k"): ("r1_bits_final_summary.csv", "16384x16384 bits, one colour pass per launch"),
    ("resident_kernel", "rx256"): ("r1_resident_rx_summary.csv", "256 x 1024^2 int8, 10 sweeps per launch"),
    ("rowstream_kernel", "single16k"): ("r2_rowstream_summary.csv", "16384x16384 int8, one sweep per launch"),
    ("persistent_kernel", "single16k"): ("r2_persist_summary.csv", "16384x16384 int8, 20 sweeps per launch"),
    ("pbatch_kernel", "single16k"): ("r2_pbatch_summary.csv", "16384x16384 int8, 20 sweeps per launch"),
    ("pbatch_kernel", "rx256"): ("r2_pbatch_rx_summary.csv", "256 x 1024^2 int8, 20 sweeps + 2 exchanges per launch"),
    ("resident_kernel", "latgas512_pair"): ("r2_latgas_resident_summary.csv", "512 x 256^2 int8 lattice gas, 20 pair-exchange sweeps per launch"),
    ("pair_class_kernel", "latgas512_pair"): ("r2_latgas_pair_summary.csv", "512 x 256^2 int8 lattice gas, one bond class per launch"),
    ("discs_run_kernel", "discs8192"): ("r2_discs_summary.csv", "8192 replicas x 20 discs, 2000 trials per launch"),
}


def profiled(kernel_name, tag):
    """Per-launch DRAM bytes and pipe utilisation of the dominant kernel from the committed ncu summary."""
    import csv
    if (kernel_name, tag) not in PROFILED:
        return None
    fname, shape = PROFILED[(kernel_name, tag)]
    path = os.path.join(ROOT, "profiles", fname)
    want = {"dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
            "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
            "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed": "fmaheavy_pipe_pct",
            "gpu__time_duration.sum": "duration"}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "%": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "": 1.0}
    out = {}
    try:
        for row in csv.reader(open(path)):
            if row and row[0] in want:
                nums = [float(v.replace(",", "")) * scale.get(row[1], 1.0) for v in row[2:] if v]
                out[want[row[0]]] = nums[-1]  # the last captured launch (the first may be a short warm-up call)
    except Exception:
        return None
    if "dram_read" not in out or "dram_write" not in out:
        return None
    out["traffic"] = out.pop("dram_read") + out.pop("dram_write")
    out["source"] = "profiles/" + fname
    out["capture_shape"] = shape
    return out


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
        return self

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


class Group:
    """The process group of the run (NCCL under torchrun, nothing on one GPU) with the few collectives bench needs."""

    def __init__(self, local_rank, world):
        self.world, self.dist, self.torch = world, None, None
        if world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.dist, self.torch = dist, torch

    def barrier(self):
        if self.dist:
            self.dist.barrier()

    def max(self, x):
        if not self.dist:
            return float(x)
        t = self.torch.tensor([float(x)], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def min(self, x):
        return -self.max(-float(x))

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# CPU baseline: the UNMODIFIED reference (oracle/_ref, see oracle/make_ref.py) on the host cores; the restated loop
# (oracle/cpu_port.py) only if the reference files are not on this machine
# ------------------------------------------------------------------------------------------------
def cpu_baseline(seconds_target=12.0, legs=False):
    from oracle import refarm
    cores = os.cpu_count() or 1
    if refarm.available():
        calib, _ = refarm.time_mcstep("ising", 1024, TC, 30000, 1)
        ntrials = int(max(calib, 5e4) * seconds_target)
        value, wall = refarm.time_mcstep("ising", 1024, TC, ntrials, cores)
        out = {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "wall_s": wall,
               "sample": "%d calls of the reference's CanonicalMonteCarlo.MCstep (py_mc/mc.py:170-185 on "
                         "examples/ising2D.py) per core on a 1024x1024 float64 lattice at kT=Tc, one replica per core "
                         "(%d cores, %.1f s; imported unmodified from oracle/_ref); the per-trial cost does not depend "
                         "on lattice size" % (ntrials, cores, wall)}
        if legs:
            out["legs"] = refarm.all_legs(seconds=2.0, cores=cores)
        return out
    from oracle import cpu_port
    calib, _ = cpu_port.time_all_cores(1024, -1.0, TC, 100000, 1)
    ntrials = int(max(calib, 5e4) * seconds_target)
    value, wall = cpu_port.time_all_cores(1024, -1.0, TC, ntrials, cores)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "wall_s": wall,
            "sample": "%d single-site MCstep trials per core (restated loop, oracle/cpu_port.py: the reference files "
                      "are not on this machine) on a 1024x1024 float64 lattice at kT=Tc, %d cores, %.1f s"
                      % (ntrials, cores, wall)}


def pick_workload(args, world):
    if args.workload != "auto":
        return args.workload
    return "ising16k" if world == 1 else "rx"


def headline_config(args, world):
    """The ``config`` object of the line: the same for our arm and for the reference arm."""
    wl = pick_workload(args, world)
    if wl == "rx":
        R = args.rx_replicas_total if args.rx_replicas_total else args.rx_replicas * world
        name = "ising2d" if args.rx_model == "ising" else "hardcore_lattice_gas_mu=%g" % args.rx_mu
        return {"workload": "%s_replica_exchange_%d_replicas_of_%dx%d_in_total_kT=%.2f..%.2f_exchange_every_%d_sweeps"
                            % (name, R, args.rx_L, args.rx_L, args.rx_kT_lo, args.rx_kT_hi, args.rx_every),
                "layout": args.layout, "replicas_total": R, "lattice": [args.rx_L, args.rx_L], "seed": SEED}
    if wl == "slab":
        return {"workload": "ising2d_%dx%d_one_lattice_in_row_slabs_kT=Tc_J=-1" % (args.L, args.L), "layout": "int8", "seed": SEED}
    if wl == "latgas":
        return {"workload": "latgas_%d_replicas_of_%dx%d_J=-1_kT=1" % (args.lg_replicas, args.lg_L, args.lg_L),
                "layout": args.layout, "seed": SEED}
    if wl == "discs":
        return {"workload": "hard_discs_%d_replicas_20_discs" % args.disc_replicas, "seed": SEED}
    return {"workload": "ising2d_%dx%d_single_replica_checkerboard_kT=%s_J=-1"
                        % (args.L, args.L, "Tc" if args.kT == TC else "%g" % args.kT),
            "layout": args.layout, "sites_per_gpu": args.L * args.L, "seed": SEED}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    nsamples = args.warmup + args.steps
    t_per_step = min(6.0, max(1.0, 60.0 / max(1, nsamples)))  # bounded samples: the whole arm ends within minutes
    vals, walls, cb = [], [], None
    for i in range(nsamples):
        c = cpu_baseline(seconds_target=t_per_step)
        if i >= args.warmup:
            vals.append(c["value"])
            walls.append(c["wall_s"])
            cb = c
    value = sum(vals) / len(vals)
    cb["value"] = value
    try:  # BASELINE.md section 4, legs 3-4, once
        from oracle import refarm
        if refarm.available():
            cb["legs"] = refarm.all_legs(seconds=2.0, cores=cb["cores"])
    except Exception as exc:
        cb["legs"] = {"error": repr(exc)}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(walls) / len(walls),
            "higher_is_better": True, "scaling": "weak" if pick_workload(args, args.gpus) != "rx" else "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": headline_config(args, args.gpus), "cpu_baseline": cb,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "each step is a bounded sample of the workload's Metropolis trials run by the reference's own "
                    "CPU code on all host cores (one replica per core, as the reference uses cores)"}
    print(json.dumps(line), flush=True)


def roofline_block(kname, alg_bytes_per_launch, launch_ms, on_chip, tag=None):
    """The contract's roofline object: algorithmic bytes per launch over the measured launch time against the measured
    HBM copy peak (SURVEY 8d accounting).  For the kernels that keep the lattice in shared memory that figure is an
    accounting figure, not DRAM traffic: ``bound`` says what really limits them and ``dram_gbs`` what DRAM moves."""
    peak, peak_src = measured_peak_hbm()
    achieved = alg_bytes_per_launch / (launch_ms * 1e-3) / 1e9
    prof = profiled(kname, tag) or {}
    r = {"bound": "alu" if on_chip else "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
         "traffic": prof.get("traffic"), "traffic_source": prof.get("source"), "traffic_capture": prof.get("capture_shape"),
         "kernel": kname, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes_per_launch,
         "accounting": "achieved = algorithmic bytes (2 x bytes/site per attempt, SURVEY 8d) / CUDA-event time of the launch"}
    for k in ("issue_active_pct", "alu_pipe_pct", "fmaheavy_pipe_pct"):
        if k in prof:
            r[k] = prof[k]
    if prof.get("traffic") and prof.get("duration"):
        r["dram_gbs_in_capture"] = prof["traffic"] / (prof["duration"] * 1e-3) / 1e9
    if on_chip:
        r["note"] = ("not HBM-bound: the lattices stay in the shared memory of the SMs for all sweeps of a launch, DRAM moves "
                     "them once per launch (`traffic`); the limiter is instruction issue on the integer ALU / FMA-heavy "
                     "pipes (Philox4x32-10 + bit-serial Metropolis test), see profiles/README.md")
    else:
        r["note"] = "crosses HBM every sweep; measured DRAM traffic equals the algorithmic bytes; issue-limited (profiles/README.md)"
    return r


def timed_repeats(grp, one_block, sync, start, stop, min_ms=MIN_TIMED_MS):
    """Time ``one_block()`` (the K timed steps) with CUDA events; if it is shorter than min_ms, repeat it inside the
    events and report the mean.  All ranks use the same repeat count (taken from the all-reduced calibration)."""
    sync()
    grp.barrier()
    start()
    one_block()
    ms1 = grp.max(stop())
    repeats = max(1, int(math.ceil(min_ms / max(ms1, 1e-3))))
    sync()
    grp.barrier()
    start()
    for _ in range(repeats):
        one_block()
    ms = grp.max(stop())
    return ms / repeats, repeats


# ------------------------------------------------------------------------------------------------
# replica exchange on `world` GPUs
# ------------------------------------------------------------------------------------------------
def make_rx(args, grp, rank, local_rank, nloc, R, group_size, offset, kernel="auto"):
    import numpy as np
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    L = args.rx_L
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    kTs = np.linspace(args.rx_kT_lo, args.rx_kT_hi, R)
    if args.rx_model == "hardcore":
        from py_mc_b200.latgas import hardcore_latticegas
        model = hardcore_latticegas(mu=args.rx_mu)
    else:
        model = ising2D(-1.0)
    lat = _cabi.Lattices(model._pmc_model, L, L, nreplicas=nloc, ntemps=R, seed=SEED, layout=layout,
                         device=local_rank, replica_offset=offset, nreplicas_global=R, kernel=KERNELS[kernel])
    if group_size > 1:
        lat.set_stream(grp.torch.cuda.current_stream().cuda_stream)
    lat.init_random()
    lat.set_accept_tables(np.array([model._pmc_thresholds(kT) for kT in kTs]))
    lat.rx_setup(1.0 / kTs, *model._pmc_energy_coeffs())
    return lat


def rx_measure(args, grp, rank, local_rank, nloc, R, distributed, kernel="auto"):
    """Time K steps of replica exchange on a batch of nloc local replicas out of R.  distributed: the R replicas are
    spread over all ranks (energies all-gathered per exchange); otherwise this rank holds all R by itself."""
    import numpy as np
    from py_mc_b200.mc_mpi import _DeviceArray
    world = grp.world if distributed else 1
    lat = make_rx(args, grp, rank, local_rank, nloc, R, world, rank * nloc if distributed else 0, kernel)
    every = args.rx_every
    st = {"scheme": 0, "done": 0, "exchanges": 0, "launches": 0}
    lps, kname = lat.sweep_kernel_info()
    peer = distributed and world > 1 and args.rx_comm == "peer" and kname == "pbatch_kernel"
    if peer:  # the exchange goes through peer memory inside the sweep kernel: one launch per call, like on one GPU
        lat.rx_connect_group(grp.dist)
    if distributed and world > 1 and not peer:
        torch, dist = grp.torch, grp.dist
        eptr, _ = lat.rx_buffers()
        e_all = torch.as_tensor(_DeviceArray(eptr, R, "<f8"), device="cuda")
        e_mine = e_all[rank * nloc:(rank + 1) * nloc]

        def steps(n, exchange=True):
            while n > 0:  # the sweeps up to the next exchange are one call
                chunk = min(n, every - st["done"] % every)
                lat.sweep(chunk)
                st["launches"] += 1 if kname in ("pbatch_kernel", "resident_kernel", "persistent_kernel") and chunk >= 2 else lps * chunk
                st["done"] += chunk
                n -= chunk
                if st["done"] % every == 0 and exchange:
                    lat.rx_energies()
                    dist.all_gather_into_tensor(e_all, e_mine)
                    lat.rx_exchange(st["scheme"])
                    st["scheme"] ^= 1
                    st["exchanges"] += 1
                    st["launches"] += 2
    else:
        def steps(n, exchange=True):
            # one GPU holds all replicas: sweeps, energies and decisions are one launch per call (pymc_rx_cycle)
            head = min(n, (every - st["done"] % every) % every)
            if head:
                if (st["done"] + head) % every == 0 and exchange:  # up to the next exchange, and that exchange
                    lat.rx_cycle(1, head, st["scheme"])
                    st["scheme"] ^= 1
                    st["exchanges"] += 1
                else:
                    lat.sweep(head)
                st["launches"] += 1
                st["done"] += head
                n -= head
            ncyc = n // every
            if ncyc and exchange:
                lat.rx_cycle(ncyc, every, st["scheme"])
                st["scheme"] ^= ncyc & 1
                st["exchanges"] += ncyc
                st["launches"] += (1 if kname == "pbatch_kernel" else ncyc * (2 + (1 if every >= 2 and lps == 1 else lps * every)))
            elif ncyc:
                lat.sweep(ncyc * every)  # the same sweeps in the same single launch, without the exchange steps
                st["launches"] += 1
            st["done"] += ncyc * every
            tail = n - ncyc * every
            if tail:
                lat.sweep(tail)
                st["launches"] += 1
                st["done"] += tail

    steps(args.warmup)
    # one untimed exchange after the barrier: first-call and skew costs stay outside, and the timed blocks start on an
    # exchange boundary (one launch each where the exchange is inside the kernel)
    lat.synchronize()
    grp.barrier()
    steps(every - st["done"] % every)
    K = args.steps
    ex0, l0 = st["exchanges"], st["launches"]
    ms, repeats = timed_repeats(grp, lambda: steps(K), lat.synchronize, lat.timer_start, lat.timer_stop)
    nex = (st["exchanges"] - ex0) / (repeats + 1.0)
    nlaunch = (st["launches"] - l0) / (repeats + 1.0)
    # the same launches without the exchange step: what an exchange costs on top of pure sweeping
    ms_plain, _ = timed_repeats(grp, lambda: steps(K, exchange=False), lat.synchronize, lat.timer_start, lat.timer_stop)
    if st["done"] % every:  # realign so that later calls start on an exchange boundary
        steps(every - st["done"] % every, exchange=False)
    nsites = nloc * args.rx_L * args.rx_L
    tried, acc = lat.rx_stats()
    out = {"lat": lat, "steps": steps, "ms": ms, "repeats": repeats, "kernel": kname, "nsites_local": nsites,
           "comm": "peer" if peer else ("nccl" if distributed and world > 1 else "none"),
           "value": world * nsites * K / (ms * 1e-3), "ms_per_step": ms / K, "exchanges_per_block": nex,
           "launches_per_block": nlaunch, "exchange_us": (1e3 * (ms - ms_plain) / nex) if nex else None,
           "acceptance": float(acc.sum()) / max(1.0, float(tried.sum())),
           "rx_sweeps_per_s": world * nloc * K / (ms * 1e-3)}
    # parity: tracked == recomputed sums on this rank; identical replica -> temperature map on every rank
    b, m = lat.tracked()
    b = b.copy()
    m = m.copy()
    rb, rm = lat.reduce_observables()
    ok = bool(np.array_equal(b, rb) and np.array_equal(m, rm))
    mp = lat.get_replica_to_T()
    ok = ok and sorted(mp.tolist()) == list(range(R))
    if distributed and world > 1:
        t = grp.torch.tensor(mp, device="cuda", dtype=grp.torch.int32)
        lo, hi = t.clone(), t.clone()
        grp.dist.all_reduce(lo, op=grp.dist.ReduceOp.MIN)
        grp.dist.all_reduce(hi, op=grp.dist.ReduceOp.MAX)
        ok = ok and bool((lo == hi).all().item())
        ok = grp.min(1.0 if ok else 0.0) > 0.5
    out["parity"] = "ok" if ok else "FAILED"
    return out


def run_rx(args):
    rank, local_rank, world = dist_env()
    grp = Group(local_rank, world)
    import numpy as np  # noqa: F401
    R = args.rx_replicas_total if args.rx_replicas_total else args.rx_replicas * world
    if R % world:
        raise SystemExit("%d replicas cannot be split evenly over %d GPUs" % (R, world))
    nloc = R // world
    sampler = ClockSampler(local_rank).start()  # before the barriers: its start-up cost is not inside a timed region
    head = rx_measure(args, grp, rank, local_rank, nloc, R, True, args.kernel)
    clocks = sampler.stop()
    lat = head["lat"]
    layout_bytes = 1.0 if args.layout == "int8" else 0.125
    nsites = head["nsites_local"]
    K = args.steps

    # end to end: host configurations in, sweeps + exchanges, observables and the map out, every step
    import numpy as np
    from py_mc_b200 import _cabi
    host = []
    for r in range(min(nloc, 4)):  # pinned host images: the uploads below are enqueued, not waited for one by one
        a = _cabi.pinned_empty((args.rx_L, args.rx_L), np.int8)
        lat.download(r, out=a)
        host.append(a)
    e2e = {}
    for spp in sorted(set([1, args.e2e_sweeps])):
        reps = 2
        lat.synchronize()
        grp.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            for r in range(nloc):
                lat.upload_async(r, host[r % len(host)])
            head["steps"](spp)
            lat.tracked()
            lat.get_replica_to_T()
        lat.synchronize()
        t = grp.max(time.perf_counter() - t0)
        e2e[spp] = world * nsites * spp * reps / t
    lat.close()

    extra = {}
    if world > 1 and not args.no_extras:
        # weak scaling: the same per-GPU batch as on one GPU (256 replicas each), one ladder over all of them
        try:
            w = rx_measure(args, grp, rank, local_rank, args.rx_replicas, args.rx_replicas * world, True, args.kernel)
            w["lat"].close()
            extra["weak"] = {"value": w["value"], "unit": UNIT, "ms_per_step": w["ms_per_step"], "replicas_per_gpu": args.rx_replicas,
                             "replicas_total": args.rx_replicas * world, "exchange_us": w["exchange_us"], "kernel": w["kernel"],
                             "exchange_comm": w["comm"],
                             "parity": w["parity"], "repeats": w["repeats"], "scaling": "weak"}
        except Exception as exc:
            extra["weak"] = {"error": repr(exc)}
        # the SAME 256-replica workload on one GPU, timed by rank 0 in this job
        n1 = None
        if rank == 0:
            try:
                one = rx_measure(args, Group(local_rank, 1), 0, local_rank, R, R, False, args.kernel)
                one["lat"].close()
                n1 = one
            except Exception as exc:
                extra["scaling_check"] = {"error": repr(exc)}
        grp.barrier()
        if rank == 0 and n1:
            extra["scaling_check"] = {
                "n1_value": n1["value"], "n1_ms_per_step": n1["ms_per_step"], "n1_kernel": n1["kernel"],
                "n1_exchange_us": n1["exchange_us"],
                "speedup": head["value"] / n1["value"], "efficiency": head["value"] / (world * n1["value"]),
                "weak_efficiency": (extra["weak"]["value"] / (world * n1["value"])) if "value" in extra.get("weak", {}) else None,
                "note": "n1 = the same %d-replica workload on one GPU, measured by rank 0 in this job" % R}

    if rank == 0:
        on_chip = head["kernel"] in ("pbatch_kernel", "resident_kernel", "persistent_kernel")
        alg = 2.0 * layout_bytes * nsites * K / max(1.0, head["launches_per_block"] - 2 * head["exchanges_per_block"])
        launch_ms = head["ms"] / max(1.0, head["launches_per_block"] - 2 * head["exchanges_per_block"])
        cfg = headline_config(args, world)
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "repeats": head["repeats"], "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if args.rx_replicas_total else "weak", "vs_baseline": None,
            "dtype": "u8" if args.layout == "int8" else "u64 (1 bit/site)", "data": "synthetic", "config": cfg,
            "workload_detail": {"replicas_per_gpu": nloc, "sites_per_gpu": nsites,
                                "step": "one sweep of every replica (2 colour passes) + exchange when due",
                                "exchanges_per_timed_block": head["exchanges_per_block"], "rx_sweeps_per_s": head["rx_sweeps_per_s"],
                                "exchange_acceptance": head["acceptance"],
                                "l2": "no cache is relied upon: the lattices sit in shared memory between exchanges" if on_chip else
                                      "working set %d MiB per GPU" % (int(nsites * layout_bytes) >> 20),
                                "parallelism": (("replicas block-partitioned over %d GPUs; per exchange every GPU stores the exact sums of "
                                                 "its replicas into the others' memory over NVLink from inside the sweep kernel and "
                                                 "all decide alike: one launch per GPU, no NCCL, no host in the loop" % world)
                                                if head["comm"] == "peer" else
                                                ("replicas block-partitioned over %d GPUs; all-gather of %d energies (NCCL, in place, on "
                                                 "the kernels' stream) per exchange; identical decision kernel on every GPU" % (world, R)))
                                if world > 1 else "one GPU holds all replicas: sweeps, energies and pair decisions in one launch"},
            "exchange_us": head["exchange_us"], "exchange_comm": head["comm"], "parity": head["parity"], "clocks": clocks,
            "e2e": {"value": e2e[args.e2e_sweeps], "unit": UNIT, "h2d_bytes_per_step": nsites, "d2h_bytes_per_step": 16 * nloc + 4 * R,
                    "sweeps_per_step": args.e2e_sweeps, "at_1_sweep_per_step": e2e[1],
                    "note": "C-ABI: upload every local replica from host int8, sweeps + exchanges, read back the exact sums and "
                            "the replica->T map; bytes are per GPU"},
            "gpu_launches": int(round(head["launches_per_block"])),
            "roofline": roofline_block(head["kernel"], alg, launch_ms, on_chip,
                                       "rx256" if (nloc, args.rx_L, args.layout, args.rx_model) == (256, 1024, "int8", "ising") else None),
        }
        line.update(extra)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    grp.close()


# ------------------------------------------------------------------------------------------------
def run_latgas(args):
    """BASELINE config 4: pair-interaction lattice gas, 512 batched replicas of 256 x 256 (size not given by
    BASELINE; stated here), J = -1, Eads - mu = 0, kT = 1.  Flip sweeps and particle-hole exchange sweeps.  Single GPU."""
    import numpy as np
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas_move import latticegas
    L, R = args.lg_L, args.lg_replicas
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    bps = 1.0 if args.layout == "int8" else 0.125
    model = latticegas(Eads=0.0, J=-1.0, mu=0.0)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, L, L, nreplicas=R, ntemps=1, seed=SEED, layout=layout, kernel=KERNELS[args.kernel])
    lat.init_random()
    lat.set_accept_tables(model._pmc_thresholds(1.0)[None, :])
    lat.set_pair_tables(model._pmc_pair_thresholds(1.0)[None, :])
    nsites = R * L * L
    grp = Group(0, 1)
    sampler = ClockSampler(0).start()
    out = {}
    lps, kname = lat.sweep_kernel_info()
    for name, fn, attempts in (("flip", lambda: lat.sweep(args.steps), nsites), ("pair_exchange", lambda: lat.pair_sweep(args.steps), 2 * nsites)):
        fn()
        ms, rep = timed_repeats(grp, fn, lat.synchronize, lat.timer_start, lat.timer_stop)
        out[name] = {"value": attempts * args.steps / (ms * 1e-3), "unit": "attempts/s", "ms_per_step": ms / args.steps, "repeats": rep}
    clocks = sampler.stop()
    b, n = lat.tracked()
    b2, n2 = lat.reduce_observables()
    parity = bool(np.array_equal(b, b2) and np.array_equal(n, n2))
    # end to end: host images in, one flip sweep + one pair sweep per step, exact sums out
    host = [lat.download(r) for r in range(4)]
    e2e = {}
    for spp in (1, args.e2e_sweeps):
        t0 = time.perf_counter()
        for r in range(R):
            lat.upload(r, host[r % 4])
        lat.sweep(spp)
        lat.pair_sweep(spp)
        lat.tracked()
        e2e[spp] = nsites * spp / (time.perf_counter() - t0)
    one_launch = kname in ("resident_kernel", "pbatch_kernel")
    nl = max(1, (args.steps + 4095) // 4096) if one_launch else lps * args.steps
    flip_ms = out["flip"]["ms_per_step"] * args.steps
    roof = roofline_block(kname, 2.0 * bps * nsites * args.steps / nl, flip_ms / nl, one_launch)
    pair_ms = out["pair_exchange"]["ms_per_step"] * args.steps
    peak, _ = measured_peak_hbm()
    out["pair_exchange"]["roofline"] = {"bound": "alu" if one_launch else "latency", "unit": "GB/s", "peak": peak,
                                        "achieved": 1.0 * bps * 2 * nsites * args.steps / (pair_ms * 1e-3) / 1e9,
                                        "accounting": "1 B (int8) per pair attempt: one read + one write of the lattice per 2N bonds"}
    out["pair_exchange"]["roofline"]["frac"] = out["pair_exchange"]["roofline"]["achieved"] / peak
    pk = "resident_kernel" if one_launch and kname == "resident_kernel" else "pair_class_kernel"
    pp = profiled(pk, "latgas512_pair") if (R, L, args.layout) == (512, 256, "int8") else None
    out["pair_exchange"]["roofline"].update({"kernel": pk, "traffic": pp["traffic"] if pp else None,
                                             "traffic_source": pp["source"] if pp else None,
                                             "issue_active_pct": pp.get("issue_active_pct") if pp else None,
                                             "alu_pipe_pct": pp.get("alu_pipe_pct") if pp else None})
    line = {"metric": "lattice-gas flip attempts/s", "value": out["flip"]["value"], "unit": "attempts/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "repeats": out["flip"]["repeats"], "ms_per_step": out["flip"]["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8" if args.layout == "int8" else "u64 (1 bit/site)",
            "data": "synthetic", "config": headline_config(args, 1), "flip_kernel": kname, "pair_exchange": out["pair_exchange"],
            "parity": "ok" if parity else "FAILED", "clocks": clocks,
            "e2e": {"value": e2e[args.e2e_sweeps], "unit": "flip attempts/s", "h2d_bytes_per_step": nsites, "d2h_bytes_per_step": 16 * R,
                    "sweeps_per_step": args.e2e_sweeps, "at_1_sweep_per_step": e2e[1],
                    "note": "C-ABI: upload every replica, flip sweeps + pair sweeps, exact sums back"},
            "gpu_launches": nl, "roofline": roof}
    if not args.no_cpu:
        from oracle import refarm
        if refarm.available():
            f = refarm.time_mcstep("latgas", 256, 1.0, 200000, 1)[0]
            mv = refarm.time_mcstep("latgas_move", 64, 1.0, 20000, 1)[0]
            line["cpu_baseline"] = {"value": f, "unit": "attempts/s", "cores": 1, "kind": "reference",
                                    "sample": "2e5 MCstep calls on examples/latgas.py latticegas (flip), 256x256, one core",
                                    "flip_or_move": {"value": mv, "sample": "2e4 MCstep calls on examples/latgas_move.py (flip or "
                                                                            "non-local move, O(N) per trial), 64x64, one core"}}
    print(json.dumps(line), flush=True)


def run_discs(args):
    """BASELINE config 5 (re-interpreted, SURVEY.md section 8 notes): batched independent hard-disc replicas of the
    shipped system (20 discs, L = 1, r = 1/14, maxstep = r/2), split evenly over the GPUs of the job."""
    rank, local_rank, world = dist_env()
    grp = Group(local_rank, world)
    import numpy as np
    from py_mc_b200.hardcore import DiscBatch
    R = args.disc_replicas // world
    b = DiscBatch(R, 1.0, 20, 1.0 / 14.0, 1.0 / 28.0, isotropic=False, seed=SEED, device=local_rank, replica_offset=rank * R)
    per = args.disc_trials
    sampler = ClockSampler(local_rank).start()
    b._d.timed_run(per)
    grp.barrier()
    ms = [b._d.timed_run(per) for _ in range(max(3, args.steps // 10))]
    ms_med = grp.max(sorted(ms)[len(ms) // 2])
    clocks = sampler.stop()
    ok = not b.any_overlap().any()
    # end to end: coordinates in, trials, coordinates and g(r) counts out
    t0 = time.perf_counter()
    co = [b._d.download(r) for r in range(4)]
    for r in range(min(R, 256)):
        b._d.upload(r, co[r % 4])
    b._d.run(per)
    b._d.observe(0.01, 49)
    e2e = grp.max(time.perf_counter() - t0)
    if rank == 0:
        line = {"metric": "hard-disc move attempts/s", "value": world * R * per / (ms_med * 1e-3), "unit": "attempts/s", "n_gpus": world,
                "steps": len(ms), "warmup": 1, "ms_per_step": ms_med, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": headline_config(args, world),
                "workload_detail": {"replicas_per_gpu": R, "trials_per_launch": per, "acceptance": float(b.acceptance().mean())},
                "parity": "ok" if ok else "FAILED", "clocks": clocks, "gpu_launches": len(ms),
                "e2e": {"value": world * R * per / e2e, "unit": "attempts/s", "h2d_bytes_per_step": 320 * min(R, 256),
                        "d2h_bytes_per_step": 4 * 320 + R * (4 + 49 * 8), "note": "upload coordinates, one launch of trials, overlap flags and g(r) counts back"},
                "roofline": {"bound": "issue", "achieved": None, "peak": None, "unit": "GB/s", "frac": None,
                             "traffic": (profiled("discs_run_kernel", "discs8192") or {}).get("traffic") if args.disc_replicas == 8192 else None,
                             "issue_active_pct": (profiled("discs_run_kernel", "discs8192") or {}).get("issue_active_pct"),
                             "alu_pipe_pct": (profiled("discs_run_kernel", "discs8192") or {}).get("alu_pipe_pct"),
                             "traffic_source": "profiles/r2_discs_summary.csv",
                             "note": "320 B of state per replica live in registers: no HBM roof (SURVEY 8d); one warp per replica, "
                                     "float64 with explicit rounding; the figure of merit is issue-slot utilisation"}}
        if not args.no_cpu:
            from oracle import refarm
            if refarm.available():
                v = refarm.time_mcstep("hc2d", 0, 1.0, 50000, 1)[0]
                line["cpu_baseline"] = {"value": v, "unit": "attempts/s", "cores": 1, "kind": "reference",
                                        "sample": "5e4 MCstep calls on examples/2D_hardcore HC_2D (20 discs), one core"}
        print(json.dumps(line), flush=True)
    grp.close()


def run_slab(args):
    """Strong scaling of ONE 16384 x 16384 lattice split by rows over the GPUs: halo rows are
    stored into the neighbours' memory by the sweep kernel itself (CUDA IPC peer mappings), neighbour barrier inside
    the kernel, no NCCL in the timed region."""
    rank, local_rank, world = dist_env()
    grp = Group(local_rank, world)
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    from py_mc_b200.slab import SlabLattice
    L = args.L
    sl = SlabLattice(_cabi.MODEL_ISING, L, L, SEED, grp.dist, device=local_rank)
    sl.init_random()
    sl.set_accept_tables(ising2D(-1.0)._pmc_thresholds(TC)[None, :])
    sl.sweep(args.warmup)
    sampler = ClockSampler(local_rank).start()
    ms, rep = timed_repeats(grp, lambda: sl.sweep(args.steps), sl.lat.synchronize, sl.lat.timer_start, sl.lat.timer_stop)
    clocks = sampler.stop()
    sl.check()
    b, m = sl.observables()
    kname = sl.lat.sweep_kernel_for(args.steps)[1]
    on_chip = kname == "pbatch_kernel"
    if rank == 0:
        nsites = L * L
        print(json.dumps({
            "metric": METRIC, "value": nsites * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "repeats": rep, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": headline_config(args, world),
            "workload_detail": {"rows_per_gpu": L // world, "bond_sum": b, "magnetisation": m,
                                "parallelism": ("row slabs kept in the shared memory of each GPU for all sweeps of a call; the first / last row "
                                                "of every colour pass goes to the neighbouring GPUs as {data, pass number} packets stored "
                                                "into their exchange buffers over NVLink: one launch per GPU, no NCCL, no host in the loop")
                                               if on_chip else
                                               ("row slabs; halo rows stored into peer memory by the sweep kernel, neighbour "
                                                "barrier inside the kernel, no NCCL in the timed region")},
            "clocks": clocks, "gpu_launches": 1 if on_chip else args.steps,
            "roofline": roofline_block(kname, 2.0 * nsites / world * (args.steps if on_chip else 1),
                                       ms if on_chip else ms / args.steps, on_chip),
        }), flush=True)
    grp.close()


def run_single(args):
    import numpy as np
    rank, local_rank, world = dist_env()
    grp = Group(local_rank, world)
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    L = args.L
    layout = _cabi.LAYOUT_BITS if args.layout == "bits" else _cabi.LAYOUT_INT8
    thr = ising2D(-1.0)._pmc_thresholds(args.kT)[None, :]
    nsites = L * L
    K = args.steps

    def make(lay, kernel=None):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=1, ntemps=1, seed=SEED, layout=lay, device=local_rank,
                             replica_offset=rank, nreplicas_global=world, kernel=KERNELS[args.kernel] if kernel is None else kernel)
        lat.init_random()
        lat.set_accept_tables(thr)
        return lat

    def timed(lat, one_block):
        lat.sweep(args.warmup)
        return timed_repeats(grp, one_block, lat.synchronize, lat.timer_start, lat.timer_stop)

    lat = make(layout)
    sampler = ClockSampler(local_rank).start()
    ms, repeats = timed(lat, lambda: lat.sweep(K))
    clocks = sampler.stop()
    value = world * nsites * K / (ms * 1e-3)

    # end to end through the C ABI with HOST buffers: upload, sweeps, download, exact sums -- every step.
    # "serial": one handle, every call waits for the one before.  Headline: the same jobs through
    # py_mc_b200.pipeline.SweepPipeline (the public way to stream host lattices through the GPU): `depth` handles, each
    # on its own stream, so that the PCIe copies of one lattice overlap the sweeps of another.  Every job still uploads
    # its own 256 MiB from pinned host memory and downloads its configuration and exact sums inside the timed region.
    from py_mc_b200.pipeline import SweepPipeline
    host = _cabi.pinned_empty((L, L), np.int8)
    lat.download(0, out=host)
    e2e, e2e_serial = {}, {}
    depth = args.e2e_depth
    bufs = [_cabi.pinned_empty((L, L), np.int8) for _ in range(depth)]
    for k, b in enumerate(bufs):  # distinct inputs: the lattice a few sweeps apart
        if k:
            lat.sweep(2)
        lat.download(0, out=b)
    for spp in sorted(set([1, args.e2e_sweeps])):
        reps = 3
        lat.upload(0, host)
        lat.sweep(spp)
        lat.synchronize()
        grp.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            lat.upload(0, host)
            lat.sweep(spp)
            lat.download(0, out=host)
            lat.tracked()
        e2e_serial[spp] = world * nsites * spp * reps / grp.max(time.perf_counter() - t0)
        with SweepPipeline(_cabi.MODEL_ISING, L, L, thr[0], spp, depth=depth, seed=SEED, layout=layout, device=local_rank,
                           kernel=KERNELS[args.kernel]) as pipe:
            for k in range(depth):  # warm-up: first-call allocations of every handle
                pipe.submit(bufs[k], bufs[k], first_sweep=0)
            pipe.drain()
            njobs, sums = 8 * depth, []  # fill and drain of the pipeline are inside the timed region
            grp.barrier()
            t0 = time.perf_counter()
            for j in range(njobs):  # job j continues the lattice of job j - depth: host -> GPU -> host every time
                r = pipe.submit(bufs[j % depth], bufs[j % depth], first_sweep=(1 + j // depth) * spp)
                if r is not None:
                    sums.append(r[2:])
            sums += [r[2:] for r in pipe.drain()]
            e2e[spp] = world * nsites * spp * njobs / grp.max(time.perf_counter() - t0)
            assert len(sums) == njobs

    bytes_per_site = 1.0 if layout == _cabi.LAYOUT_INT8 else 0.125
    lps, kname = lat.sweep_kernel_info()
    on_chip = kname in ("persistent_kernel", "pbatch_kernel", "resident_kernel")
    nlaunch = (K + 4095) // 4096 if on_chip else lps * K
    roof = roofline_block(kname, 2.0 * bytes_per_site * nsites * K / nlaunch, ms / nlaunch, on_chip,
                          "single16k" if (L, args.layout) == (16384, "int8") or (L == 16384 and kname == "sweep_bits_rows_kernel") else None)

    extra = {}
    if not args.no_extras:
        lat.close()
        if args.also_bits and layout == _cabi.LAYOUT_INT8:
            try:  # an extra, never allowed to take the headline line down with it
                lat2 = make(_cabi.LAYOUT_BITS)
                ms2, _ = timed(lat2, lambda: lat2.sweep(K))
                extra["bitpacked"] = {"value": world * nsites * K / (ms2 * 1e-3), "unit": UNIT, "ms_per_step": ms2 / K,
                                      "kernel": lat2.sweep_kernel_info()[1],
                                      "note": "same workload on the 1-bit/site layout (32 MiB)"}
                lat2.close()
            except Exception as exc:
                extra["bitpacked"] = {"error": repr(exc)}
        try:  # the per-sweep kernel on the same workload (one launch per sweep, 2 B/attempt of DRAM)
            lat3 = make(layout)
            n3 = max(50, K)
            ms3, _ = timed(lat3, lambda: [lat3.sweep(1) for _ in range(n3)])  # single-sweep calls: the fused kernel
            k3 = lat3.sweep_kernel_for(1)[1]
            extra["one_sweep_per_launch"] = {"value": world * nsites * n3 / (ms3 * 1e-3), "unit": UNIT, "ms_per_step": ms3 / n3, "kernel": k3,
                                             "roofline": roofline_block(k3, 2.0 * bytes_per_site * nsites / (1 if layout == _cabi.LAYOUT_INT8 else 2),
                                                                        ms3 / n3 / (1 if layout == _cabi.LAYOUT_INT8 else 2), False,
                                                                        "single16k" if L == 16384 else None),
                                             "roofline_frac": 2.0 * bytes_per_site * nsites / (ms3 / n3 * 1e-3) / 1e9 / measured_peak_hbm()[0]}
            lat3.close()
        except Exception as exc:
            extra["one_sweep_per_launch"] = {"error": repr(exc)}
        if world == 1:
            try:  # BASELINE configs[2] on this one GPU: 256 x 1024^2, exchange every rx_every sweeps
                rx = rx_measure(args, grp, 0, local_rank, args.rx_replicas, args.rx_replicas, False, args.kernel)
                rx["lat"].close()
                extra["rx"] = {"value": rx["value"], "unit": UNIT, "ms_per_step": rx["ms_per_step"], "repeats": rx["repeats"],
                               "kernel": rx["kernel"], "exchange_us": rx["exchange_us"], "parity": rx["parity"],
                               "rx_sweeps_per_s": rx["rx_sweeps_per_s"], "exchange_acceptance": rx["acceptance"],
                               "workload": "ising2d_replica_exchange_%d_replicas_of_%dx%d_kT=%.2f..%.2f_exchange_every_%d_sweeps"
                                           % (args.rx_replicas, args.rx_L, args.rx_L, args.rx_kT_lo, args.rx_kT_hi, args.rx_every)}
            except Exception as exc:
                extra["rx"] = {"error": repr(exc)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup, "repeats": repeats,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8" if layout == _cabi.LAYOUT_INT8 else "u64 (1 bit/site)", "data": "synthetic",
            "config": headline_config(args, world),
            "workload_detail": {"step": "one full sweep (both colours)",
                                "l2": ("no cache is relied upon: the lattice (%d MiB in HBM) is loaded once per launch, stays in the shared "
                                       "memory of the SMs as bits for the %d sweeps of the launch and is stored once"
                                       % (int(nsites * bytes_per_site) >> 20, K)) if on_chip else
                                      "working set %d MiB %s L2" % (int(nsites * bytes_per_site) >> 20,
                                                                    "exceeds" if nsites * bytes_per_site > 126e6 else "fits in"),
                                "parallelism": "replicas only" if world > 1 else "single GPU"},
            "clocks": clocks,
            "e2e": {"value": e2e[args.e2e_sweeps], "unit": UNIT, "h2d_bytes_per_step": nsites, "d2h_bytes_per_step": nsites + 16,
                    "sweeps_per_step": args.e2e_sweeps, "at_1_sweep_per_step": e2e[1], "pipeline_depth": depth,
                    "serial": {"value": e2e_serial[args.e2e_sweeps], "at_1_sweep_per_step": e2e_serial[1],
                               "note": "one handle, upload -> sweeps -> download -> sums strictly one after the other"},
                    "note": "py_mc_b200.pipeline.SweepPipeline over the C ABI: every step uploads its lattice from pinned host "
                            "int8, sweeps it, downloads the configuration and the exact sums; %d handles on their own "
                            "streams, so the copies of one step overlap the sweeps of another" % depth},
            "gpu_launches": nlaunch * repeats, "roofline": roof,
        }
        line.update(extra)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    grp.close()


def run_ours(args):
    _, _, world = dist_env()
    wl = pick_workload(args, world)
    if wl == "slab":
        return run_slab(args)
    if wl == "latgas":
        return run_latgas(args)
    if wl == "discs":
        return run_discs(args)
    if wl == "rx":
        if args.workload == "auto" and not args.rx_replicas_total:
            args.rx_replicas_total = 256  # BASELINE configs[2]: 256 replicas in total
        return run_rx(args)
    return run_single(args)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--layout", default="int8", choices=["int8", "bits"])
    ap.add_argument("--L", type=int, default=16384)
    ap.add_argument("--kT", type=float, default=TC, help="temperature of the single-lattice workload (default T_c; "
                    "SURVEY 8d also asks for 2.0 and 5.0)")
    ap.add_argument("--e2e-sweeps", type=int, default=64)
    ap.add_argument("--e2e-depth", type=int, default=4, help="lattices in flight in the end-to-end leg (handles / streams)")
    ap.add_argument("--also-bits", action="store_true", default=True)
    ap.add_argument("--no-also-bits", dest="also_bits", action="store_false")
    ap.add_argument("--no-extras", action="store_true", help="headline only (profiling runs)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workload", default="auto", choices=["auto", "ising16k", "rx", "latgas", "discs", "slab"])
    ap.add_argument("--lg-replicas", type=int, default=512)
    ap.add_argument("--lg-L", type=int, default=256)
    ap.add_argument("--disc-replicas", type=int, default=8192)
    ap.add_argument("--disc-trials", type=int, default=2000)
    ap.add_argument("--rx-replicas", type=int, default=256, help="replicas per GPU (weak scaling / the rx block at N = 1)")
    ap.add_argument("--rx-replicas-total", type=int, default=0, help="replicas over all GPUs (strong scaling); default at "
                    "N > 1: 256, BASELINE configs[2]")
    ap.add_argument("--rx-L", type=int, default=1024)
    ap.add_argument("--rx-every", type=int, default=10, help="sweeps between exchanges")
    ap.add_argument("--rx-comm", default="peer", choices=["peer", "nccl"], help="several GPUs: energies exchanged through peer "
                    "memory inside the sweep kernel (one launch per GPU), or by an NCCL all-gather between launches")
    ap.add_argument("--kernel", default="auto", choices=sorted(KERNELS), help="sweep kernel of the handle (PYMC_KERNEL_*)")
    ap.add_argument("--rx-model", default="ising", choices=["ising", "hardcore"])
    ap.add_argument("--rx-mu", type=float, default=1.0, help="chemical potential of the hard-core lattice gas")
    ap.add_argument("--rx-kT-lo", type=float, default=2.0)
    ap.add_argument("--rx-kT-hi", type=float, default=2.6)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
