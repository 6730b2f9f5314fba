#!/usr/bin/env python3
"""Multi-GPU check + measurement of the vertex-range partitioned closure.
Launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
             --master-port P tests/run_dist_closure.py [--vertices V --edges E]
Every rank builds the same seeded graph, uploads only the edges it owns, runs the closure from
the entry points with NCCL allgathers between levels, and rank 0 checks the gathered levels
against the CSR oracle and prints one JSON line."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--vertices", type=int, default=2_000_000)
    ap.add_argument("--edges", type=int, default=50_000_000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--check", default="oracle", choices=["oracle", "single", "none"],
                    help="oracle: CSR oracle on rank 0; single: dmg_closure on the whole graph on rank 0's GPU")
    ap.add_argument("--sync-stepper", action="store_true", help="use the synchronous *_dev primitives")
    ap.add_argument("--fast-gen", action="store_true", help="with-replacement targets (config 5 sizes)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import dmgraph
    from dmgraph import dist as ddist
    from dmgraph import synth
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = dmgraph.Engine(local)
    t_gen = time.perf_counter()
    gen = synth.big_graph_fast if args.fast_gen else synth.big_graph
    sg = gen(args.vertices, args.edges, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
    t_gen = time.perf_counter() - t_gen
    V = sg.n_vertices
    _wr, v_lo, v_hi = ddist.partition(V, rank, world)
    s_own, d_own = ddist.owned_edges(sg.src, sg.dst, v_lo, v_hi)
    t_up = time.perf_counter()
    g = dmgraph.Graph(eng, V, s_own, d_own, [])
    t_up = time.perf_counter() - t_up
    stepper = (ddist.GpuStepper if args.sync_stepper else ddist.GpuStreamStepper)(g, v_lo, v_hi)
    pc = ddist.PartitionedClosure(V, rank, world, stepper, "cuda", dist if world > 1 else None)
    times = []
    for it in range(1 + args.steps):
        pc.stats = {k: 0 for k in pc.stats if k != "level_ms"}
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        level, reached = pc.run(sg.entry, ddist.OUT)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if it:
            times.append(time.perf_counter() - t0)
    edges = torch.tensor([pc.stats["edges"]], dtype=torch.float64, device="cuda")
    tmax = torch.tensor([float(np.median(times))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(edges)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        lv = [torch.empty(pc.wr * 32, dtype=torch.int32, device="cuda") for _ in range(world)]
        mine = torch.full((pc.wr * 32,), -1, dtype=torch.int32, device="cuda")
        mine[: v_hi - v_lo] = level
        dist.all_gather(lv, mine)
        full = torch.cat(lv)[:V].cpu().numpy().view(np.uint32)
    else:
        full = level.cpu().numpy().view(np.uint32)
    if rank == 0:
        ok, single = None, None
        if args.check == "oracle":
            from oracle import oracle as ora
            want, wreached, wte = ora.CsrOracle(V, sg.src, sg.dst, sg.entry).closure(0, sg.entry)
            ok = bool(reached == wreached and np.array_equal(full, want))
        elif args.check == "single":
            # the whole graph on this one GPU: cross-check and the 1-GPU figure for the same graph
            t0 = time.perf_counter()
            gf = dmgraph.Graph(eng, V, sg.src, sg.dst, [])
            t_full = time.perf_counter() - t0
            eng.profile(True)
            want, wreached = gf.closure(dmgraph.OUT, sg.entry)
            eng.profile_reset()
            ts = []
            for _ in range(args.steps):
                t0 = time.perf_counter()
                want, wreached = gf.closure(dmgraph.OUT, sg.entry)
                ts.append(time.perf_counter() - t0)
            st = eng.last_stats()
            kms = sum(v[1] for v in eng.profile_table().values()) / args.steps
            ok = bool(reached == wreached and np.array_equal(full, want))
            single = {"kept_edges": int(gf.info["n_edges"]), "duplicate_edges": int(gf.info["duplicate_edges"]),
                      "upload_s": t_full, "closure_e2e_s": float(np.median(ts)), "closure_kernel_ms": kms,
                      "edges_traversed": int(st["edges_traversed"]), "edges_read": int(st["edges_read"]),
                      "gteps": st["edges_traversed"] / (kms * 1e-3) / 1e9,
                      "device_bytes": int(gf.info["device_bytes"])}
            gf.release()
        print(json.dumps({"check": "partitioned_closure", "n_gpus": world, "vertices": V, "edges": int(sg.src.size),
                          "owned_edges_rank0": int(s_own.size), "reached": int(reached), "levels": pc.stats["levels"],
                          "push_levels": pc.stats["push_levels"], "pull_levels": pc.stats["pull_levels"],
                          "seconds": float(tmax.item()), "edges_read_all_ranks": float(edges.item()),
                          "gteps_read": float(edges.item()) / float(tmax.item()) / 1e9,
                          "exchanged_bytes_per_rank": pc.stats["exchanged_bytes"], "level_ms_rank0": pc.stats.get("level_ms"), "generate_s": t_gen,
                          "upload_s_rank0": t_up, "stepper": "sync" if args.sync_stepper else "stream-ordered", "checked_against": args.check, "matches": ok,
                          "single_gpu": single}), flush=True)
        assert ok is not False
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
