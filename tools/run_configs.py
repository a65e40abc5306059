"""BASELINE configs[2] and configs[3] at full size, one process per GPU (torchrun), NCCL only at the end.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
        tools/run_configs.py --config 3 --steps 20000

  --config 3  2x2x2 replica of the shipped box (N = 240), 4096 chains sharded in contiguous blocks
  --config 4  dihedral sweep: 36 conformers (0..350 degrees about atoms 9-8-1-2) x 256 chains = 9216 chains; per-chain
              mean energies are all-gathered, averaged per angle, and handed to the reference's window hopping
              (postprocess.window_hopping, PostProcess.F90:260-304)

Timing: CUDA events on each rank, maximum over ranks.  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from montecarlo_b200 import boxio, dihedral, partition, postprocess  # noqa: E402
from montecarlo_b200.engine import LJEngine, comm_unique_id  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, choices=[3, 4], required=True)
    ap.add_argument("--steps", type=int, default=20000, help="Metropolis steps per chain")
    ap.add_argument("--launch", type=int, default=5000, help="steps per launch")
    ap.add_argument("--chains", type=int, default=0, help="total chains (default: 4096 for config 3, 9216 for config 4)")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    inputs = ROOT / "tests" / "golden" / "inputs"
    box = boxio.read_box(inputs / "box.txt")
    system = boxio.load_system(box, inputs / "DMSO.txt", inputs / "param.txt", inputs / "par_solute.txt")
    angles = np.arange(0.0, 360.0, 10.0)
    if args.config == 3:
        b = boxio.replicate(box, 2)
        system = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute})
        n_total = args.chains or 4096
    else:
        b = box
        n_total = args.chains or len(angles) * 256
    per_angle = n_total // len(angles)
    start, count = partition.chain_block(rank, world, n_total)
    eng = LJEngine(system, b.n, n_chains=count, device=local)
    stream = torch.cuda.current_stream()
    eng.set_stream(stream.cuda_stream)
    if args.config == 4:
        conf, _ = dihedral.conformer_sweep(box.solute, box.sol_sym, box.dihedrals[0][:2], angles)
        mine = (np.arange(start, start + count) // per_angle).clip(max=len(angles) - 1)   # chain -> angle, contiguous blocks
        eng.set_solute(conf[mine], np.broadcast_to(system.sol_tab, (count,) + system.sol_tab.shape).copy())
    eng.set_state(b.com, b.hoek)
    eng.set_mc_params(start_prod=args.steps // 2)
    eng.seed(partition.global_seeds(rank, world, n_total))
    e_start = eng.total_energy()[0]
    # warm-up: three launches (they count as equilibration: the chains simply go on afterwards)
    for _ in range(3):
        eng.run_steps(args.launch, want_counts=False)
    torch.cuda.synchronize()
    st_w = eng.stats()
    if world > 1:
        dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    left = args.steps
    while left > 0:
        n = min(left, args.launch)
        eng.run_steps(n, want_counts=False)
        left -= n
    t1.record(stream)
    torch.cuda.synchronize()
    # one chain per rank at full size against the C restatement on the same stream (checker only, untimed)
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib as ol
    st_c = eng.stats(0, 1)
    osys = system if args.config == 3 else boxio.System(**{**system.__dict__, "solute": conf[mine[0]]})
    o = ol.OracleChain(osys, b.com, b.hoek)
    o.set_mc_params(start_prod=args.steps // 2)
    o.total_energy()
    o.run_steps_rng(3 * args.launch + args.steps, ol.rng_seed(start + 1), trace=False)
    so = o.stats()
    chain_ok = bool(st_c["nsuc"][0] == so["nsuc"] and st_c["sitepairs"][0] == so["sitepairs"]
                    and abs(st_c["toteng"][0] - so["toteng"]) <= 1e-10 * abs(so["toteng"]))
    okt = torch.tensor([1.0 if chain_ok else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # the job's one collective, through the C ABI (NCCL inside libmcb200.so); torch.distributed only carries the id
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        eng.comm_init(world, rank, idt.cpu().numpy().tobytes(), n_total)
    eng.reduce_stats()   # (a new communicator sets up its connections in its first collective: untimed)
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    c0.record(stream)
    g = eng.reduce_stats()                                                        # the one all-reduce
    per_group = eng.reduce_stats(per_angle) if args.config == 4 else None         # (config 4: per dihedral angle)
    mean_e = eng.gather_chains()["mean_e"]                                        # the one all-gather
    c1.record(stream); torch.cuda.synchronize()
    tot = {k: float(v[0]) for k, v in g.items()}
    w_timed = torch.tensor([float((eng.stats()["sitepairs"] - st_w["sitepairs"]).sum()), float((eng.stats()["step"] - st_w["step"]).sum())],
                           dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(w_timed, op=dist.ReduceOp.SUM)
    if rank == 0:
        sec = ms.item() * 1e-3
        line = {"config": args.config, "n_gpus": world, "n_molecules": b.n, "chains": n_total, "steps_per_chain": args.steps,
                "seconds": sec, "warmup_launches": 3, "site_pairs_per_s": w_timed[0].item() / sec, "trial_moves_per_s": w_timed[1].item() / sec,
                "oracle_check_one_chain_per_rank": bool(okt.item() == 1.0),
                "acceptance": tot["acceptance"], "mean_production_energy": tot["mean_e"], "start_energy_chain0": float(e_start[0]),
                "collectives_ms": c0.elapsed_time(c1)}
        if args.config == 4:
            per = mean_e[: per_angle * len(angles)].reshape(len(angles), per_angle)
            e_angle = per.mean(axis=1)
            line["per_angle_mean_energy"] = [round(float(v), 4) for v in e_angle]
            # the grouped all-reduce gives the same per-angle means from sums (weights: production samples per chain)
            line["per_angle_mean_energy_allreduce"] = [round(float(v), 4) for v in per_group["mean_e"][: len(angles)]]
            line["per_angle_sem"] = [round(float(v), 4) for v in per.std(axis=1) / np.sqrt(per_angle)]
            # window hopping over the gathered energies: one "run" per chain index inside an angle block
            rng = np.random.default_rng(2026)
            old, now, acc = postprocess.window_hopping(per.T.copy(), 300.0, iter(rng.random(4 * per_angle)))
            visits = np.bincount(np.where(acc, now, old), minlength=len(angles))
            line["window_hopping"] = {"runs": int(per_angle), "accepted": int(acc.sum()), "most_visited_angle_deg": float(angles[int(np.argmax(visits))])}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
