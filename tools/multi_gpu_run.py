"""Multi-GPU evidence run (torchrun, NCCL): BASELINE configs 4 and 3 on N GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/multi_gpu_run.py [--scale S] [--columns C] [--out file.json]

Part A (config 4): one CIRC-1 column on the full 1-50000 cm-1 range at 0.01 cm-1 (7 molecules
+ continua + CIA + 3 CFCs), longwave and shortwave, the spectral range split over the ranks
(halo of lines and grid points either side, grtcode_b200/parallel.py) and the partial
broadband integrals summed by one NCCL all-reduce.  Rank 0 then runs the unsplit column and
reports the flux difference.
Part B (config 3 shape): C RFMIP-like columns at 0.1 cm-1 block-partitioned over the ranks, no
collective on the data path, results all-gathered once at the end; rank 0 recomputes every
column alone and compares.
Times are device times (CUDA events), max over ranks.  Rank 0 prints one JSON object.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from grtcode_b200 import api, synth
    from grtcode_b200 import parallel as P
    from grtcode_b200 import workloads as W
    from grtcode_b200.column_engine import ColumnEngine

    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.)
    ap.add_argument("--columns", type=int, default=8)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--data", default="/tmp/grt_bench")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    api.load()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    per_rank = []

    def timed(fn, reps):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(reps):
            out = fn()
        e1.record(stream)
        torch.cuda.synchronize()
        mine = e0.elapsed_time(e1)*1e-3/reps       # this rank's own time (no barrier inside)
        barrier()
        per_rank.clear()
        if world > 1:
            t = torch.tensor([mine], device=f"cuda:{local}", dtype=torch.float64)
            got = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(got, t)
            per_rank.extend(float(g.item()) for g in got)
        else:
            per_rank.append(mine)
        return out, max(per_rank)

    lw2, sw2 = W.circ1_bands(0.01, args.scale)
    if local == 0:
        W.materialise(lw2, args.data, seed=1)
        W.materialise(sw2, args.data, seed=2)
    barrier()
    if local != 0:
        W.materialise(lw2, args.data, seed=1)
        W.materialise(sw2, args.data, seed=2)
    col = W.circ1_column()
    result = {"n_gpus": world, "line_scale": args.scale}

    # ---- Part A: config 4, spectral split ----
    full = [W.Band("lw_full", sw2.w0, sw2.wn, sw2.dw, sw2.lines, sw2.dataset),
            W.Band("sw_full", sw2.w0, sw2.wn, sw2.dw, sw2.lines, sw2.dataset)]
    part = P.SpectralSplitColumn(full, col, local, W.MOLECULES, world, rank)
    part.run(col)                                     # warm-up
    split, secs_split = timed(lambda: part.run(col), args.reps)
    split_ranks = list(per_rank)
    compute = [part.compute_seconds]
    if world > 1:
        t = torch.tensor(compute, device=f"cuda:{local}", dtype=torch.float64)
        got = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(got, t)
        compute = [float(g.item()) for g in got]
    chunk = part.chunks["sw_full"]
    part.close()
    two = P.split_spectral_range(sw2.w0, sw2.wn, sw2.dw, 2)[0]
    halo = two.hi - two.last
    a = {"config": "BASELINE config 4: CIRC-1 column, 1-50000 cm-1 at 0.01 cm-1 (n=4999901), LW + SW, "
                   "7 molecules + continua + 3 CIA + 3 CFCs, spectral range split over the GPUs, "
                   "one NCCL all-reduce of the partial broadband integrals per column",
         "seconds_per_column": secs_split, "seconds_per_rank": split_ranks,
         "compute_seconds_per_rank_before_allreduce": compute,
         "pts_layers_per_s": 2*full[0].n*(len(col.p) - 1)/secs_split,
         "rank0_chunk_points": chunk.n, "halo_points": halo,
         "allreduce_bytes": 2*2*8*len(col.p)}
    if rank == 0:
        whole = ColumnEngine(full, col, local, W.MOLECULES)
        whole.step(col, upload_surface=True)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        ref = whole.step(col, upload_surface=True)
        e1.record(stream)
        torch.cuda.synchronize()
        whole.close()
        a["seconds_per_column_one_gpu"] = e0.elapsed_time(e1)*1e-3
        a["speedup_vs_one_gpu"] = a["seconds_per_column_one_gpu"]/secs_split
        err = {}
        for k in ("lw", "sw"):
            for d, name in ((0, "up"), (1, "down")):
                err[f"{k}_{name}"] = float(np.abs(split[k][d] - ref[k][d]).max())
        a["max_abs_flux_difference_W_m2"] = err
        a["fluxes_W_m2"] = {"lw_up_toa": float(split["lw"][0][0]), "lw_dn_sfc": float(split["lw"][1][-1]),
                            "sw_up_toa": float(split["sw"][0][0]), "sw_dn_sfc": float(split["sw"][1][-1])}
    barrier()
    result["spectral_split"] = a

    # ---- Part B: config 3 shape, columns sharded ----
    lw3, sw3 = W.circ1_bands(0.1, args.scale)
    lw3.dataset, sw3.dataset = lw2.dataset, sw2.dataset
    cfcs = tuple(lw2.dataset.cfc.keys())
    ncols = args.columns

    def column(k: int):
        rng = np.random.default_rng(7000 + k % 100)
        scale = (1., 4., 0.5, 2., 3., 8.)[(k//100) % 6]
        return synth.make_column(61, rng, molecules=W.MOLECULES, cfcs=cfcs, co2_scale=scale,
                                 dt=4.*(k % 3 - 1))

    eng = ColumnEngine((lw3, sw3), column(0), local, W.MOLECULES)
    eng.step(column(0), upload_surface=True)
    mine = P.shard_columns(ncols, world, rank)

    def run_shard(ids):
        rows = []
        for k in ids:
            out = eng.step(column(k), upload_surface=True)
            rows.append(np.stack([out["lw"][0], out["lw"][1], out["sw"][0], out["sw"][1]]))
        return np.stack(rows) if rows else np.zeros((0, 4, 61))

    local_rows, secs_cols = timed(lambda: run_shard(mine), 1)
    gathered = P.gather_columns(local_rows, ncols)
    b = {"config": "BASELINE config 3 shape: 61 levels, LW 1-3250 + SW 1-50000 cm-1 at 0.1 cm-1, "
                   "columns block-partitioned over the GPUs, no collective on the data path",
         "columns": ncols, "seconds": secs_cols, "columns_per_s": ncols/secs_cols,
         "seconds_for_1800_columns": 1800.*secs_cols/ncols}
    if rank == 0:
        alone = run_shard(range(ncols))
        b["max_abs_difference_vs_one_gpu_W_m2"] = float(np.abs(gathered - alone).max())
    eng.close()
    result["column_shard"] = b

    if rank == 0:
        line = json.dumps(result)
        print(line)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as f:
                f.write(line + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
