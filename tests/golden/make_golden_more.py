"""More golden-fixture generators (replica exchange, lattice gas, hard discs); see make_golden.py."""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

from oracle import philox, ising as o_ising  # noqa: E402

RX_CASES = [
    dict(name="rx4_8x8", R=4, lenX=8, lenY=8, J=-1.0, kT_lo=1.5, kT_hi=3.5, seed=777, nsweeps=40, rx_every=2),
    dict(name="rx6_6x10", R=6, lenX=6, lenY=10, J=-1.0, kT_lo=2.0, kT_hi=2.6, seed=4242, nsweeps=30, rx_every=1),
]


class RXInjector:
    """Per-rank injector: checkerboard site order + per-site uniforms for MCstep, and the
    (round, pair) uniform for Xtrial (read from the caller's frame: myTrank in Xtrial, i in run)."""

    def __init__(self, case, rank):
        self.case, self.rank = case, rank
        self.N = case["lenX"] * case["lenY"]
        self.n = -1
        self.phase = 0
        self.sweep_loaded = -1
        self.base_sweep = 0
        self.base_round = 0
        self.exchange_draws = 0

    def _load(self, sweep):
        c = self.case
        self.sites, self.us = [], []
        for colour in (0, 1):
            u = philox.x32_to_uniform(o_ising.site_x32(c["seed"], self.rank, sweep, colour, c["lenX"], c["lenY"]))
            for i in range(c["lenX"]):
                q = (i + colour) & 1
                for k in range(c["lenY"] // 2):
                    self.sites.append((i, 2 * k + q))
                    self.us.append(u[i, k])
        self.sweep_loaded = sweep

    def randrange(self, n):
        if self.phase == 0:
            self.n += 1
            sweep = self.base_sweep + self.n // self.N
            if sweep != self.sweep_loaded:
                self._load(sweep)
            self.phase = 1
            return int(self.sites[self.n % self.N][0])
        self.phase = 0
        return int(self.sites[self.n % self.N][1])

    def random_metropolis(self):
        return float(self.us[self.n % self.N])

    def random_exchange(self):
        f = sys._getframe(1)
        assert f.f_code.co_name == "Xtrial"
        t = int(f.f_locals["myTrank"])
        run = f.f_back
        assert run.f_code.co_name == "run"
        rnd = self.base_round + run.f_locals["i"] // run.f_locals["RXtrial_frequency"] - 1
        self.exchange_draws += 1
        return float(philox.x32_to_uniform(philox.rx_x32(self.case["seed"], rnd, t)))


def _rx_rank_main(comm, case, workdir, ref_root):
    import fake_mpi
    np.int = int      # removed from numpy >= 1.24; py_mc/mc_mpi.py:79-81 still uses them
    np.float = float
    fake_mpi.install(comm)
    sys.path.insert(0, ref_root)
    sys.path.insert(0, os.path.join(ref_root, "examples"))
    import py_mc.mc as refmc
    import py_mc.mc_mpi as refmpi
    import ising2D as ref_ising

    rank = comm.Get_rank()
    inj = RXInjector(case, rank)
    ref_ising.rand = inj
    refmc.random = inj.random_metropolis
    refmpi.random = inj.random_exchange
    os.chdir(workdir)
    R, N = case["R"], case["lenX"] * case["lenY"]
    kTs = np.linspace(case["kT_lo"], case["kT_hi"], R)
    configs = []
    for r in range(R):
        cfg = ref_ising.ising2D_config(case["lenX"], case["lenY"])
        cfg.config[:, :] = o_ising.init_config(case["seed"], r, case["lenX"], case["lenY"])
        configs.append(cfg)
    model = ref_ising.ising2D(case["J"])
    rx = refmpi.TemperatureRX_MPI(comm, refmc.CanonicalMonteCarlo, model, configs, kTs)
    obs = ref_ising.observer()
    out = rx.run(case["nsweeps"] * N, case["rx_every"] * N, N, 10 ** 12, obs, True, True)
    # post-processing of the reference (examples/analyze_ising2D_mpi.py), run unmodified on the files above
    import runpy
    sys.argv = ["analyze_ising2D_mpi.py", str(R)]
    comm.coll = 0
    runpy.run_path(os.path.join(ref_root, "examples", "analyze_ising2D_mpi.py"), run_name="__main__")
    res = dict(
        rank_to_T=np.array(rx.rank_to_T, dtype=np.int64),
        config=rx.mycalc.config.config.astype(np.int8),
        obs_save=np.load(os.path.join(workdir, str(rank), "obs_save.npy")),
        Trank_hist=np.load(os.path.join(workdir, str(rank), "Trank_hist.npy")),
        kT_hist=np.load(os.path.join(workdir, str(rank), "kT_hist.npy")),
        energy=float(rx.mycalc.energy),
        exchange_draws=inj.exchange_draws,
        obs_list=np.array(out, dtype=np.float64) if rank == 0 else None,
        files=sorted(os.listdir(os.path.join(workdir, str(rank)))),
        top_files=sorted(f for f in os.listdir(workdir) if not f.isdigit() and f not in ("obs_mean.dat", "binning.dat"))
        if rank == 0 else None,
        obs_mean=open(os.path.join(workdir, "obs_mean.dat")).read() if rank == 0 else None,
        binning=open(os.path.join(workdir, "binning.dat")).read() if rank == 0 else None,
    )
    return res


def gen_rx(out):
    import fake_mpi
    ref_root = os.environ.get("PYMC_REFERENCE", "/root/reference")
    meta = []
    for case in RX_CASES:
        workdir = tempfile.mkdtemp()
        res = fake_mpi.run_ranks(case["R"], _rx_rank_main, (case, workdir, ref_root))
        n = case["name"]
        out[n + "_rank_to_T"] = res[0]["rank_to_T"]
        for r in range(case["R"]):
            assert np.array_equal(res[r]["rank_to_T"], res[0]["rank_to_T"])
        out[n + "_configs"] = np.array([r["config"] for r in res])
        out[n + "_obs_save"] = np.array([r["obs_save"] for r in res])      # [rank, sample, (E, |M|)]
        out[n + "_Trank_hist"] = np.array([r["Trank_hist"] for r in res])  # [rank, sample]
        out[n + "_kT_hist"] = np.array([r["kT_hist"] for r in res])
        out[n + "_energy"] = np.array([r["energy"] for r in res])
        out[n + "_obs_list"] = res[0]["obs_list"]                          # [T, (E, |M|)] means
        c = dict(case)
        c["files"] = res[0]["files"]
        c["top_files"] = res[0]["top_files"]
        c["exchange_draws"] = int(sum(r["exchange_draws"] for r in res))
        c["obs_mean_dat"] = res[0]["obs_mean"]
        c["binning_dat"] = res[0]["binning"]
        meta.append(c)
        print("rx", n, "final rank_to_T", res[0]["rank_to_T"], "draws", c["exchange_draws"], "files", c["files"],
              c["top_files"])
        print("   mean obs per T:", np.round(res[0]["obs_list"], 3).tolist())
    out["meta"] = np.array(json.dumps(meta))


LATGAS_CASES = [
    # name, lenX, lenY, J, Eads, mu, kT, seed, replica, nsweeps
    ("lg_5x... shipped-like J=0", 6, 6, 0.0, -0.4, -0.1, 0.25, 11, 0, 8),
    ("lg_attract", 8, 10, -1.0, -0.5, 0.25, 1.0, 12, 1, 10),
    ("lg_repulse", 8, 8, 1.0, -2.0, 0.0, 0.5, 13, 2, 10),
    ("lg_word_6x128", 6, 128, -0.5, 0.25, -0.75, 0.75, 14, 0, 5),
    ("lg_partial_4x132", 4, 132, 0.25, -0.5, 0.0, 0.3, 15, 3, 5),
    # nearest-neighbour exclusion ("hard-core lattice gas", BASELINE config 5 read as a lattice model): J so large
    # that exp(-dE/kT) underflows to 0.0 for an insertion next to a particle; random start relaxes into the
    # exclusion manifold, then only nn = 0 insertions / any removal are accepted
    ("lg_hardcore", 8, 8, 1.0e6, 0.0, 0.5, 1.0, 16, 1, 10),
]


def gen_latgas(out):
    """Flip moves of examples/latgas.py driven in checkerboard order with injected uniforms."""
    ref_root = os.environ.get("PYMC_REFERENCE", "/root/reference")
    sys.path.insert(0, ref_root)
    sys.path.insert(0, os.path.join(ref_root, "examples"))
    import py_mc.mc as refmc
    import latgas as ref_latgas
    from make_golden import SiteInjector, checkerboard_schedule
    from oracle import latgas as o_latgas
    inj = SiteInjector()
    ref_latgas.rand = inj
    refmc.random = inj.random
    meta = []
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())
    try:
        for n, (name, lenX, lenY, J, Eads, mu, kT, seed, replica, nsweeps) in enumerate(LATGAS_CASES):
            name = "lg%d" % n
            cfg = ref_latgas.latgas_config(lenX, lenY)
            init = o_latgas.init_config(seed, replica, lenX, lenY)
            cfg.config[:, :] = init
            model = ref_latgas.latticegas(Eads=Eads, J=J, mu=mu)
            calc = refmc.CanonicalMonteCarlo(model, kT, cfg)
            # dE for every site of the initial configuration, straight from trialstep
            sites = [(x, y) for x in range(lenX) for y in range(lenY)]
            inj.load(sites, [0.0] * len(sites))
            dEs = np.zeros((lenX, lenY))
            for (x, y) in sites:
                dconfig, dE = model.trialstep(cfg, 0.0)
                assert dconfig == [x, y]
                dEs[x, y] = dE
            e0 = model.energy(cfg)
            energies, dens, configs = [], [], []
            for sw in range(nsweeps):
                s2, us = checkerboard_schedule(seed, replica, sw, lenX, lenY)
                inj.load(s2, us)
                calc.run(lenX * lenY)
                energies.append(calc.energy)
                dens.append(model.density(cfg))
                configs.append(cfg.config.astype(np.int8).copy())
            out[name + "_init"] = init
            out[name + "_dE"] = dEs
            out[name + "_e0"] = np.float64(e0)
            out[name + "_energy"] = np.array(energies, dtype=np.float64)
            out[name + "_density"] = np.array(dens, dtype=np.float64)
            out[name + "_configs"] = np.array(configs, dtype=np.int8)
            meta.append(dict(name=name, lenX=lenX, lenY=lenY, J=J, Eads=Eads, mu=mu, kT=kT, seed=seed,
                             replica=replica, nsweeps=nsweeps))
            print("latgas", name, "E0", e0, "E_end", energies[-1], "N_end", dens[-1])
    finally:
        os.chdir(cwd)
    out["meta"] = np.array(json.dumps(meta))


class MoveInjector:
    """random-module stand-in for latgas_move.trialstep: scripted random(), randrange(), choice()."""

    def __init__(self):
        self.script = []

    def random(self):
        kind, v = self.script.pop(0)
        assert kind == "random"
        return v

    def randrange(self, n):
        kind, v = self.script.pop(0)
        assert kind == "randrange" and 0 <= v < n
        return v

    def choice(self, seq):
        kind, v = self.script.pop(0)
        assert kind == "choice"
        return seq[v % len(seq)]


def gen_latgas_move(out):
    """dconfig / dE of examples/latgas_move.py trialstep (flip and move branches) on fixed
    configurations with scripted draws -- including the defective move dE -- and newconfig."""
    ref_root = os.environ.get("PYMC_REFERENCE", "/root/reference")
    sys.path.insert(0, ref_root)
    sys.path.insert(0, os.path.join(ref_root, "examples"))
    import latgas_move as ref_move
    from oracle import latgas as o_latgas
    inj = MoveInjector()
    ref_move.rand = inj
    rng = np.random.default_rng(2718)
    meta = []
    for n, (lenX, lenY, J, Eads, mu, seed) in enumerate([(5, 5, -0.25, -0.4, -0.1, 21), (6, 8, 1.0, 0.5, 0.25, 22),
                                                         (4, 4, 0.0, -0.4, 0.3, 23)]):
        name = "mv%d" % n
        cfg = ref_move.latgas_config(lenX, lenY)
        init = (rng.random((lenX, lenY)) >= 0.5).astype(np.int8) if lenX % 2 else o_latgas.init_config(seed, 0, lenX, lenY)
        cfg.config[:, :] = init
        model = ref_move.latticegas(Eads=Eads, J=J, mu=mu)
        scripts, dconfigs, dEs = [], [], []
        nocc, nemp = int(init.sum()), int((1 - init).sum())
        trials = []
        for t in range(60):
            if t % 3 == 0:
                trials.append([("random", 0.25), ("randrange", int(rng.integers(lenX))), ("randrange", int(rng.integers(lenY)))])
            else:
                trials.append([("random", 0.75), ("choice", int(rng.integers(nocc))), ("choice", int(rng.integers(nemp)))])
        # make sure the quirky branches are hit: empty site in the last row / last column
        emp = np.transpose(np.where(init == 0))
        for idx, (x, y) in enumerate(emp):
            if x == lenX - 1 or y == lenY - 1:
                trials.append([("random", 0.9), ("choice", 0), ("choice", idx)])
        for tr in trials:
            inj.script = list(tr)
            dconfig, dE = model.trialstep(cfg, 0.0)
            assert not inj.script
            scripts.append([[k, float(v)] for k, v in tr])
            dconfigs.append([int(v) for v in dconfig] + [-1] * (4 - len(dconfig)))
            dEs.append(float(dE))
        assert np.array_equal(cfg.config, init)
        after = ref_move.latgas_config(lenX, lenY)
        after.config[:, :] = init
        mv = next(d for d in dconfigs if d[2] >= 0)
        model.newconfig(after, mv)
        out[name + "_init"] = init
        out[name + "_dconfig"] = np.array(dconfigs, dtype=np.int64)
        out[name + "_dE"] = np.array(dEs, dtype=np.float64)
        out[name + "_after_move"] = after.config.astype(np.int8)
        out[name + "_move"] = np.array(mv, dtype=np.int64)
        meta.append(dict(name=name, lenX=lenX, lenY=lenY, J=J, Eads=Eads, mu=mu, scripts=scripts))
        print("latgas_move", name, "trials", len(trials), "moves", sum(1 for d in dconfigs if d[2] >= 0))
    out["meta"] = np.array(json.dumps(meta))


DISC_CASES = [
    dict(name="hd_shipped", L=1.0, rdisc=1.0 / 14.0, ndisc=20, maxstep=1.0 / 28.0, seed=88, replica=0, ntrials=3000,
         dr=0.01),
    dict(name="hd_dense", L=2.0, rdisc=0.21, ndisc=16, maxstep=0.25, seed=89, replica=5, ntrials=2000, dr=0.05),
]


def gen_discs(out):
    """examples/2D_hardcore: HC_2D trialstep/newconfig under CanonicalMonteCarlo.MCstep, energy, g_r."""
    import types
    ref_root = os.environ.get("PYMC_REFERENCE", "/root/reference")
    sys.path.insert(0, os.path.join(ref_root, "py_mc"))                 # "from mc import model" (model_setup.py:10)
    sys.path.insert(0, os.path.join(ref_root, "examples", "2D_hardcore"))
    stub = types.ModuleType("matplotlib")                                # matplotlib is absent; only plot_fig uses it
    stub.pyplot = types.ModuleType("matplotlib.pyplot")
    sys.modules.setdefault("matplotlib", stub)
    sys.modules.setdefault("matplotlib.pyplot", stub.pyplot)
    import mc as refmc_flat
    import model_setup as ref_hc
    from oracle import harddisc as o_hd

    state = {}

    class NPProxy:
        """numpy with random.randint replaced (model_setup.py:45 draws the disc from numpy's global RNG)."""

        class random:
            @staticmethod
            def randint(n):
                return state["id"]

        def __getattr__(self, k):
            return getattr(np, k)

    class RandProxy:
        @staticmethod
        def uniform(a, b):
            return a + (b - a) * state["u"]

    ref_hc.np = NPProxy()
    ref_hc.rand = RandProxy()
    refmc_flat.random = lambda: state["x32"] * (1.0 / 4294967296.0)
    meta = []
    for case in DISC_CASES:
        n = case["name"]
        cfg = ref_hc.HC_2D_config(case["L"], case["ndisc"], case["rdisc"])
        cfg.prepare_ordered()
        start = cfg.coors.copy()
        model = ref_hc.HC_2D(case["maxstep"])
        grid = refmc_flat.grid_1D(case["dr"], case["dr"], case["L"] / 2.0)
        calc = refmc_flat.CanonicalMonteCarlo(model, 1, cfg, grid=grid)
        calc.energy = model.energy(cfg)
        e0 = calc.energy
        g0 = ref_hc.g_r(cfg, grid)
        snaps, nacc, ids = [], 0, []
        for t in range(case["ntrials"]):
            state["id"], state["u"], state["x32"] = o_hd.trial_draws(case["seed"], case["replica"], t, case["ndisc"])
            before = cfg.coors.copy()
            calc.MCstep()
            nacc += int(not np.array_equal(before, cfg.coors))
            ids.append(state["id"])
            if (t + 1) % 500 == 0:
                snaps.append(cfg.coors.copy())
        out[n + "_start"] = start
        out[n + "_snaps"] = np.array(snaps)
        out[n + "_g0"] = g0
        out[n + "_g_end"] = ref_hc.g_r(cfg, grid)
        out[n + "_gridx"] = grid.x
        out[n + "_ids"] = np.array(ids[:50])
        c = dict(case)
        c.update(nacc=nacc, e0=float(e0), e_end=float(model.energy(cfg)), g_pref=float(cfg.g_pref))
        meta.append(c)
        print("discs", n, "accepted", nacc, "of", case["ntrials"], "E0", e0, "nbins", len(grid.x))
    out["meta"] = np.array(json.dumps(meta))


MORE = {"rx": gen_rx, "latgas": gen_latgas, "latgas_move": gen_latgas_move, "discs": gen_discs}
