#!/usr/bin/env python
"""bench.py — throughput of the Metropolis trial-move energy path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA arm (one rank per GPU under torchrun)
    python bench.py --impl reference [--gpus N --steps K --warmup W]   the reference's own CPU code on the host cores

A "step" is one pass of the hot path over one batch: `mcb_run_steps` advancing every chain on the GPU by
--mc-steps Metropolis steps (one kernel launch).  Headline workload = BASELINE.json configs[1]: the reference's shipped
DMSO box (30 molecules + 11-atom solute), 1024 independent chains per GPU, T = 300 K, config0.ini step sizes.
`value` = evaluated site-pair interactions per second (LJ_FORMULA evaluations of trial states), whole job; `e2e` = the
same through the C ABI with page-locked host buffers, H2D and D2H copies inside the timed region.  A second timed leg,
`target_box`, is the box north_star's target is quoted on: the 2x2x2 periodic replica (240 molecules), 512 chains per GPU
= configs[2]'s per-GPU share (at --gpus 8 it is configs[2] at full size).  Each leg checks one chain per rank against the
C restatement (untimed).  The job's one collective (statistics all-reduce + per-chain all-gather) runs inside
libmcb200.so over NCCL.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "oracle"))

FLOP_PER_SITEPAIR = 42          # literal count of LJ_FORMULA incl. the difference and the two accumulates
METRIC = "site_pair_interactions_per_s"
UNIT = "site-pairs/s"


def workload_name(chains, mc_steps):
    return (f"configs[1]: shipped DMSO box (boxbin/box.txt: 30 molecules, 11-atom solute, BOXL=7.7838685668), "
            f"{chains} independent chains per GPU (seeds 1..), T=300K, LJ+Coulomb(conv=0) as the reference, "
            f"{mc_steps} Metropolis steps per chain per launch")


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); power.append(float(r[3]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU arms
def mean_sitepairs_per_move(nsteps=20000) -> float:
    """Mean LJ_FORMULA evaluations per trial move on this workload, counted by replaying the oracle."""
    import oracle_lib as ol
    box, system = ol.shipped_system()
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    o.run_steps_rng(nsteps, ol.rng_seed(1), trace=False)
    return o.stats()["sitepairs"] / nsteps


def cpu_reference_run(lj_steps: int, nproc: int):
    """The reference's own binary (oracle/_ref), nproc independent single-threaded processes (its LJ path has
    no OpenMP region, MonteCarlo.F90:792-828; multi-run parallelism in the reference is separate processes)."""
    import run_ref
    results = [None] * nproc

    def one(k):
        results[k] = run_ref.run_reference(lj_steps=lj_steps, nprint=0, ndump=0, want_matrix=False)

    t0 = time.perf_counter()
    th = [threading.Thread(target=one, args=(k,)) for k in range(nproc)]
    [t.start() for t in th]
    [t.join() for t in th]
    wall = time.perf_counter() - t0
    walls = [r.walltime for r in results if r is not None and r.walltime]
    # the binary's own `Walltime:` (MonteCarlo.F90:677) covers start-up + the LJ loop; use the slowest process
    return max(walls) if walls else wall, wall


def cpu_port_run(lj_steps: int, nproc: int):
    """The C restatement (oracle/libmc_oracle.so), one chain per thread (ctypes releases the GIL)."""
    import oracle_lib as ol
    box, system = ol.shipped_system()
    chains = []
    for k in range(nproc):
        o = ol.OracleChain(system, box.com, box.hoek)
        o.total_energy()
        chains.append((o, ol.rng_seed(k + 1)))
    t0 = time.perf_counter()
    th = [threading.Thread(target=lambda oc=oc: oc[0].run_steps_rng(lj_steps, oc[1], trace=False)) for oc in chains]
    [t.start() for t in th]
    [t.join() for t in th]
    return time.perf_counter() - t0, sum(c[0].stats()["sitepairs"] for c in chains)


def cpu_baseline(seconds_target=12.0) -> dict:
    """Bounded sample of the same workload on the host: the reference binary on one core when oracle/_ref
    travelled with the repo, else the C port on one core."""
    import run_ref
    w_per_move = mean_sitepairs_per_move()
    if run_ref.available():
        steps = int(4.0e4 * seconds_target)
        t, _ = cpu_reference_run(steps, 1)
        return {"value": steps * w_per_move / t, "unit": UNIT, "cores": 1, "kind": "reference",
                "moves_per_s": steps / t,
                "sample": f"reference binary MonteCarloOpenMP (serial LJ path), 1 chain, LJ={steps} steps of the shipped "
                          f"box, its own Walltime {t:.2f}s; site-pairs = moves x {w_per_move:.1f} (oracle replay)"}
    steps = int(1.0e5 * seconds_target)
    t, w = cpu_port_run(steps, 1)
    return {"value": w / t, "unit": UNIT, "cores": 1, "kind": "port", "moves_per_s": steps / t,
            "sample": f"C restatement (oracle/), 1 chain, {steps} steps of the shipped box, {t:.2f}s"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import run_ref
    ncores = os.cpu_count() or 1
    nproc = max(1, ncores)
    total = args.steps + args.warmup
    use_ref = run_ref.available()
    per_core_rate = 4.0e4 if use_ref else 1.0e5
    lj = int(max(2000, min(100000, per_core_rate * 150.0 / max(total, 1))))   # whole run stays within minutes
    w_per_move = mean_sitepairs_per_move()
    times = []
    for k in range(total):
        if use_ref:
            t, _ = cpu_reference_run(lj, nproc)
            w = nproc * lj * w_per_move
        else:
            t, w = cpu_port_run(lj, nproc)
        if k >= args.warmup:
            times.append((t, w))
    tsum = sum(t for t, _ in times)
    wsum = sum(w for _, w in times)
    value = wsum / tsum
    kind = "reference" if use_ref else "port"
    sample = (f"{nproc} independent single-threaded processes of the reference binary, LJ={lj} steps each per step"
              if use_ref else f"{nproc} threads of the C restatement, {lj} steps each per step")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tsum / max(len(times), 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.chains, args.mc_steps), "chains_per_gpu": args.chains},
            "trial_moves_per_s": value / w_per_move,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": nproc, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ GPU arm
def target_workload_name(chains, mc_steps):
    return (f"configs[2] per-GPU share: 2x2x2 periodic replica of the shipped box (240 molecules, BOXL=15.5677371336, one "
            f"11-atom solute), {chains} independent chains per GPU (seeds 1..), T=300K, {mc_steps} Metropolis steps per "
            f"chain per launch")


def oracle_check(eng, system, com, hoek, seed, nsteps, chain=0):
    """One chain of the full-size batch against the C restatement on the same stream (checker only, untimed)."""
    import oracle_lib as ol
    st = eng.stats(chain, 1)
    o = ol.OracleChain(system, com, hoek)
    o.total_energy()
    o.run_steps_rng(nsteps, ol.rng_seed(int(seed)), trace=False)
    so = o.stats()
    rel = abs(st["toteng"][0] - so["toteng"]) / abs(so["toteng"])
    ok = bool(st["nsuc"][0] == so["nsuc"] and st["sitepairs"][0] == so["sitepairs"] and rel <= 1e-10)
    return {"chain_seed": int(seed), "steps": int(nsteps), "accepted_equal": bool(st["nsuc"][0] == so["nsuc"]),
            "sitepairs_equal": bool(st["sitepairs"][0] == so["sitepairs"]), "toteng_rel_err": float(rel), "ok": ok}


def measure_leg(eng, system, box, nch, S, steps, warmup, seeds, stream, flush, barrier, torch):
    """One workload on this rank: oracle check of the first chain, warm-up, `steps` event-timed launches with the state
    resident in HBM (L2 flushed between them), then the same through the C ABI with host buffers."""
    com0 = np.ascontiguousarray(np.broadcast_to(box.com, (nch, box.n, 3)))
    hoek0 = np.ascontiguousarray(np.broadcast_to(box.hoek, (nch, box.n, 3)))
    eng.set_state(com0, hoek0)
    eng.seed(seeds)
    eng.total_energy()
    n_check = min(S, 2000)
    eng.run_steps(n_check, want_counts=False)
    check = oracle_check(eng, system, box.com, box.hoek, seeds[0], n_check)
    for _ in range(max(warmup, 3)):
        eng.run_steps(S, want_counts=False)
    torch.cuda.synchronize()
    st0 = eng.stats()
    l0 = eng.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    for k in range(steps):
        flush.zero_()
        ev[k][0].record(stream)
        eng.run_steps(S, want_counts=False)
        ev[k][1].record(stream)
    barrier()
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    launches = eng.launch_count - l0
    st1 = eng.stats()
    out = {"check": check, "kernel_ms": kernel_ms, "launches": launches, "st1": st1,
           "w": float((st1["sitepairs"] - st0["sitepairs"]).sum()), "moves": float((st1["step"] - st0["step"]).sum()),
           "t": sum(kernel_ms) * 1e-3}
    # ---- end to end through the C ABI with host buffers: H2D of every chain's CoM/HOEK, K-step launch,
    #      D2H of the statistics and of the final coordinates (what the driver's dump/result writers read)
    # the host buffers are page-locked, as the harness contract asks (a driver would cudaHostRegister its arrays)
    com_h = torch.empty((nch, box.n, 3), dtype=torch.float64, pin_memory=True).numpy()
    hoek_h = torch.empty((nch, box.n, 3), dtype=torch.float64, pin_memory=True).numpy()
    eng.get_state(out=(com_h, hoek_h))
    e2e_ms, w_e2e = [], 0.0
    out["h2d"] = com_h.nbytes + hoek_h.nbytes
    out["d2h"] = com_h.nbytes + hoek_h.nbytes + nch * 128
    for k in range(steps + 1):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s_before = eng.stats()["sitepairs"].sum()
        barrier()
        a.record(stream)
        eng.set_state(com_h, hoek_h)          # pinned staging + cudaMemcpy2DAsync inside the library
        eng.run_steps(S, want_counts=False)   # includes the start-up recompute the new state requires
        st = eng.stats()
        eng.get_state(out=(com_h, hoek_h))
        b.record(stream)
        barrier()
        if k > 0:
            e2e_ms.append(a.elapsed_time(b))
            w_e2e += float(st["sitepairs"].sum() - s_before)
    out.update(e2e_ms=e2e_ms, w_e2e=w_e2e, t_e2e=sum(e2e_ms) * 1e-3)
    return out


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from montecarlo_b200.engine import LJEngine, measure_fp64_peak, comm_unique_id
    from montecarlo_b200 import boxio, partition

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    inputs = ROOT / "tests" / "golden" / "inputs"
    box = boxio.read_box(inputs / "box.txt")
    system = boxio.load_system(box, inputs / "DMSO.txt", inputs / "param.txt", inputs / "par_solute.txt")
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def join_comm(eng, n_total):
        """The job's NCCL communicator lives inside libmcb200.so; torch.distributed only carries the 128-byte id."""
        if world == 1:
            return
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        eng.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()), n_total)

    sampler = ClockSampler(local)
    sampler.start()
    # ---- leg 1 = the headline: BASELINE configs[1], chains partitioned over ranks in contiguous blocks
    nch, S = args.chains, args.mc_steps
    eng = LJEngine(system, box.n, n_chains=nch, device=local, lookahead=args.lookahead)
    eng.set_stream(stream.cuda_stream)
    seeds = partition.global_seeds(rank, world, nch * world)
    join_comm(eng, nch * world)
    m1 = measure_leg(eng, system, box, nch, S, args.steps, args.warmup, seeds, stream, flush, barrier, torch)
    # the path's one collective, inside the product: all-reduce of the statistics + all-gather of per-chain finals
    eng.reduce_stats()        # (the first collective of a new communicator also sets up its connections: untimed)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    a.record(stream)
    gstats = eng.reduce_stats()
    finals = eng.gather_chains()
    b.record(stream)
    torch.cuda.synchronize()
    stats_ms = a.elapsed_time(b)
    st_end = eng.stats()
    assert finals["toteng"].shape[0] == nch * world and int(gstats["n_chains"][0]) == nch * world
    assert int(gstats["step"][0]) == int(st_end["step"].sum()) * world      # every rank made the same number of steps
    assert np.array_equal(finals["toteng"][rank * nch:(rank + 1) * nch], st_end["toteng"])
    eng.close()

    # ---- leg 2 = the box north_star's target is quoted on: 2x2x2 replica, configs[2]'s per-GPU share of chains
    m2 = None
    if args.target_chains > 0:
        b2 = boxio.replicate(box, 2)
        sys2 = boxio.System(**{**system.__dict__, "boxl": b2.boxl, "solute": b2.solute})
        nch2, S2 = args.target_chains, args.target_mc_steps
        eng2 = LJEngine(sys2, b2.n, n_chains=nch2, device=local)
        eng2.set_stream(stream.cuda_stream)
        seeds2 = partition.global_seeds(rank, world, nch2 * world)
        m2 = measure_leg(eng2, sys2, b2, nch2, S2, args.steps, args.warmup, seeds2, stream, flush, barrier, torch)
        eng2.close()
    clocks = sampler.stop()

    # ---- whole-job aggregation: max time over ranks, sum of work (torch.distributed = plumbing for the bench's own numbers)
    def aggregate(m):
        t, te, w, mv, we = m["t"], m["t_e2e"], m["w"], m["moves"], m["w_e2e"]
        ok = 1.0 if m["check"]["ok"] else 0.0
        if world > 1:
            tt = torch.tensor([t, te, -ok], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t, te, ok = tt[0].item(), tt[1].item(), -tt[2].item()
            sv = torch.tensor([w, mv, we], dtype=torch.float64, device="cuda")
            dist.all_reduce(sv, op=dist.ReduceOp.SUM)
            w, mv, we = sv.tolist()
        return t, te, w, mv, we, ok == 1.0

    agg1 = aggregate(m1)
    agg2 = aggregate(m2) if m2 else None

    if rank == 0:
        peak_tf = measure_fp64_peak(local, 150.0)
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except (OSError, ValueError):
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        traffic = {}
        tp = ROOT / "profiles" / "traffic_r02.json"
        if tp.exists():
            try:
                traffic = json.loads(tp.read_text())
            except ValueError:
                pass

        def roofline(m, kernel, alg_bytes, tkey):
            w_launch = m["w"] / max(m["launches"], 1)
            t_launch = m["t"] / max(m["launches"], 1)
            achieved_tf = w_launch * FLOP_PER_SITEPAIR / t_launch / 1e12
            tr = traffic.get(tkey, {})
            moves_launch = m["moves"] / max(m["launches"], 1)
            dram_move, l2_move, dram_launch = tr.get("dram_bytes_per_move"), tr.get("l2_bytes_per_move"), tr.get("dram_bytes_per_launch")
            if tr.get("scales_with") == "launch":      # the capture's bytes belong to a launch (image in, image out), whatever its length
                dram_move = dram_launch / moves_launch
                l2_move = tr.get("l2_bytes_per_launch", 0.0) / moves_launch
            elif tr.get("scales_with") == "move" and dram_move is not None:
                dram_launch = dram_move * moves_launch
            return {"bound": "fp64", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
                    "traffic": dram_launch, "kernel": kernel, "flop_per_site_pair": FLOP_PER_SITEPAIR,
                    "site_pairs_per_launch": w_launch, "kernel_ms": 1e3 * t_launch,
                    "dram_bytes_per_move": dram_move, "l2_bytes_per_move": l2_move,
                    "warp_instructions_per_move": tr.get("warp_instructions_per_move"),
                    "traffic_source": tr.get("source"),
                    "peak_source": "measured live: mcb_measure_fp64_peak (independent DFMA chains, all SMs, loop-invariant "
                                   "multiplier and addend: 98 % of the nominal 148 SM x 64 DFMA/clk x 1.965 GHz); "
                                   "MEASURED_PEAKS.json has no fp64 entry",
                    "hbm": {"achieved": alg_bytes / t_launch / 1e9, "peak": hbm_peak, "unit": "GB/s",
                            "frac": alg_bytes / t_launch / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                            "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}

        t_all, t_e2e_all, w_all, moves_all, w_e2e_all, ok1 = agg1
        value = w_all / t_all
        # algorithmic bytes per launch: every chain's image is read once and its mutable part written once
        g_image = (19 * 32 + 12 * 3 + 4 * 48 + 30 * 31) * 8
        alg1 = nch * (g_image + (7 * 32 + 12 * 30 + 30 * 31) * 8 + 2 * 216)
        st1 = m1["st1"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(nch, S), "chains_per_gpu": nch, "mc_steps_per_launch": S,
                       "parallelism": f"chains x{world} (no data-path collective; one NCCL all-reduce + all-gather of statistics "
                                      f"at the end, inside libmcb200.so: mcb_reduce_stats / mcb_gather_chains)",
                       "l2": "flushed between timed launches (256 MiB memset outside the event pairs)",
                       "uniforms": "built-in counter stream (mcb_seed_chains), seed = global chain index + 1"},
            "trial_moves_per_s": moves_all / t_all,
            "sitepairs_per_move": w_all / max(moves_all, 1.0),
            "acceptance": float(st1["nsuc"].sum() / max(st1["step"].sum(), 1)),
            "oracle_check": {**m1["check"], "all_ranks_ok": ok1},
            "e2e": {"value": w_e2e_all / t_e2e_all, "unit": UNIT, "h2d_bytes_per_step": m1["h2d"], "d2h_bytes_per_step": m1["d2h"],
                    "ms_per_step": 1e3 * t_e2e_all / max(len(m1["e2e_ms"]), 1),
                    "path": "mcb_set_chain_state -> mcb_run_steps -> mcb_get_stats + mcb_get_chain_state (page-locked host numpy buffers)"},
            "gpu_launches": int(m1["launches"]),
            "clocks": clocks,
            "roofline": roofline(m1, "k_run_spec", alg1, "k_run_spec"),
            "stats_collective_ms": stats_ms,
        }
        if m2:
            t2, te2, w2, mv2, we2, ok2 = agg2
            n2 = 240
            # per launch: the chain image (coordinates, site cache, tables) in and its mutable part out; the pair
            # matrix is read once for the row sums (N x MS doubles) and written only where it changes
            alg2 = args.target_chains * ((7 * n2 + 12 * (n2 + 3) + 4 * 48 + n2 * (n2 + 1)) * 8 + (7 * n2 + 12 * n2) * 8 + 2 * 216)
            line["target_box"] = {
                "workload": target_workload_name(args.target_chains, args.target_mc_steps),
                "chains_per_gpu": args.target_chains, "mc_steps_per_launch": args.target_mc_steps,
                "value": w2 / t2, "unit": UNIT, "trial_moves_per_s": mv2 / t2, "sitepairs_per_move": w2 / max(mv2, 1.0),
                "ms_per_step": 1e3 * t2 / args.steps, "gpu_launches": int(m2["launches"]),
                "acceptance": float(m2["st1"]["nsuc"].sum() / max(m2["st1"]["step"].sum(), 1)),
                "oracle_check": {**m2["check"], "all_ranks_ok": ok2},
                "e2e": {"value": we2 / te2, "unit": UNIT, "h2d_bytes_per_step": m2["h2d"], "d2h_bytes_per_step": m2["d2h"],
                        "ms_per_step": 1e3 * te2 / max(len(m2["e2e_ms"]), 1)},
                "roofline": roofline(m2, args.target_kernel_name, alg2, "target_box"),
            }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--chains", type=int, default=1024, help="chains per GPU")
    ap.add_argument("--mc-steps", type=int, default=5000, help="Metropolis steps per chain per launch")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--target-chains", type=int, default=512, help="chains per GPU of the second leg (2x2x2 replica, N = 240); 0 = skip")
    ap.add_argument("--target-mc-steps", type=int, default=4000, help="Metropolis steps per chain per launch of the second leg")
    ap.add_argument("--target-kernel-name", default="k_run_vl2", help=argparse.SUPPRESS)
    ap.add_argument("--lookahead", type=int, default=0, help="speculative look-ahead depth of mcb_run_steps (0 = library default)")
    args = ap.parse_args()
    # stdout carries the ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL prints a
    # version banner there when NCCL_DEBUG asks for it) are sent to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    try:
        if args.impl == "reference":
            run_reference_arm(args)
        else:
            run_gpu_arm(args)
    finally:
        sys.stdout.flush()


if __name__ == "__main__":
    main()
