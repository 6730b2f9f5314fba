#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on BASELINE config 2 (the configuration the metric is
quoted on that fits one GPU): 50 project graphs / 100 k classes / 2 M import edges, one batch of
65,536 stack-trace neighbour-resolution queries (mode BOTH, k = 2, sources excluded).

A "step" is one pass of the hot path over one batch.  Three measurements, one JSON line:
  value  queries/s with the batch and the result buffers resident in HBM (dmg_neighbors_batch_dev),
         timed with the CUDA events libdmgraph records on the stream it launches on
  e2e    the same through dmg_neighbors_batch with pinned HOST buffers: query H2D and result D2H
         inside the timed region
  cpu_baseline  the structure-faithful CPU restatement of the Java path (oracle `ref`) on a bounded
         sample of the same batch, all host cores (rank 0, N = 1 only)

  cpu_baseline_csr  the integer-CSR CPU restatement on the WHOLE batch (context: what a CPU does with
         the same data structure the GPU uses)
  parity_checked    after the timed region the resident result of the last timed step is copied to the
         host and a seeded sample of >= 4,096 rows is compared with the oracle, bit for bit
  configs  (N = 1) BASELINE configs 3, 4, 5 -- the HBM-bound ones -- measured by bench_configs.py, each
         with its SURVEY 8d bytes, kernel time, roofline fraction, e2e and its own oracle check

N > 1 (torchrun): the CSR is replicated, every rank runs its own 65,536-query batch (weak
scaling, no data-path collective); value = all queries / max-over-ranks time.

--impl reference times the CPU restatement alone (the reference is Java; no JVM in this image).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
sys.path.insert(0, ROOT)

METRIC = "neighbor_resolution_queries_per_sec"
UNIT = "queries/s"
N_QUERIES = 65536
K_HOPS = 2
WORKLOAD = ("c2: synthetic 50-service ecosystem, 100k classes / 2M import edges, "
            "65536 stack-trace neighbor-resolution queries per GPU, mode BOTH, k=2, sources excluded")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def build_workload(rank):
    from dmgraph import synth
    sg = synth.eco_graph(with_names=True)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=N_QUERIES, seed=synth.SEED_C2_QUERIES + rank)
    return sg, q_off, q_src


def cpu_reference(sg, q_off, q_src, budget_s=15.0, threads=None):
    """Times the string-keyed restatement (`ref`) on a bounded sample; returns (q/s, dict)."""
    from dmgraph import synth
    from oracle import oracle as ora
    threads = threads or ora.hardware_threads()
    name_off, arena = synth.pack_names(sg.names)
    svc = ora.RefService(name_off, arena, sg.src, sg.dst, sg.entry, sg.part_off)
    pilot = max(threads, 8)
    t0 = time.perf_counter()
    svc.khop_batch(2, K_HOPS, True, 0, pilot, q_off, q_src, threads)
    dt = time.perf_counter() - t0
    n = int(min(N_QUERIES - pilot, max(pilot, budget_s / max(dt / pilot, 1e-9))))
    n = max(threads, (n // threads) * threads)
    t0 = time.perf_counter()
    res, _chk = svc.khop_batch(2, K_HOPS, True, pilot, pilot + n, q_off, q_src, threads)
    dt = time.perf_counter() - t0
    qps = n / dt
    return qps, {"value": qps, "unit": UNIT, "cores": threads, "kind": "port",
                 "sample": "%d of %d queries (uniform prefix after a %d-query pilot), %.1f s wall; "
                           "string-keyed C++ restatement of ProjectGraph.resolve/CodeContextService."
                           "resolveNeighbors iterated to k=2; JVM effects not reproduced" % (n, N_QUERIES, pilot, dt),
                 "results_per_query": res / n}


def cpu_csr(sg, q_off, q_src, threads=None):
    """The integer-CSR restatement on the whole batch; returns (dict, oracle handle, result)."""
    from oracle import oracle as ora
    threads = threads or ora.hardware_threads()
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
    t0 = time.perf_counter()
    res = o.khop_batch(2, K_HOPS, 1, q_off, q_src, threads=threads)
    dt = time.perf_counter() - t0
    nq = q_off.size - 1
    return ({"value": nq / dt, "unit": UNIT, "cores": threads, "kind": "port",
             "sample": "all %d queries, %.1f s wall; integer-CSR restatement (same results as the string-keyed one, "
                       "none of its hashing): context, not the baseline of record" % (nq, dt)}, o, res)


def run_reference(args, rank, world):
    if rank != 0:
        return
    sg, q_off, q_src = build_workload(0)
    from oracle import oracle as ora
    threads = ora.hardware_threads()
    from dmgraph import synth
    name_off, arena = synth.pack_names(sg.names)
    svc = ora.RefService(name_off, arena, sg.src, sg.dst, sg.entry, sg.part_off)
    per_step = 16 * threads      # ~5 s per step at ~50 queries/s on 16 threads; shrunk below if the run would not fit
    times, counts = [], []
    pos = 0
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        svc.khop_batch(2, K_HOPS, True, pos, pos + per_step, q_off, q_src, threads)
        dt = time.perf_counter() - t0
        pos = (pos + per_step) % (N_QUERIES - per_step)
        if s >= args.warmup:
            times.append(dt)
            counts.append(per_step)
        if s == 0 and dt * (args.warmup + args.steps) > 240:      # keep the whole run within minutes
            per_step = max(threads, int(per_step * 240 / (dt * (args.warmup + args.steps))))
    qps = sum(counts) / sum(times)
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic", "config": {"workload": WORKLOAD},
            "sample_queries_per_step": per_step,
            "cpu_baseline": {"value": qps, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d queries per step (bounded sample of the 65536-query batch), all host "
                                       "threads; C++ restatement of the Java path, no JVM in this image" % per_step},
            "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--configs", default="c3,c4,c5", help="N=1: extra BASELINE configs to measure ('' = none)")
    ap.add_argument("--c3-names", type=int, default=10_000_000)
    ap.add_argument("--c5-vertices", type=int, default=4_000_000)
    ap.add_argument("--c5-edges", type=int, default=100_000_000)
    ap.add_argument("--parity-rows", type=int, default=4096)
    ap.add_argument("--closure-vertices", type=int, default=-1,
                    help="N>1: partitioned-closure block; -1 = 2.5 M vertices / 62.5 M drawn edges PER GPU (weak "
                         "scaling: N = 8 is BASELINE config 5 at full size, 20 M / 500 M), 0 = skip")
    ap.add_argument("--closure-edges", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import dmgraph   # raises if libdmgraph.so is missing: there is no fallback
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()

    def allmax(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    eng = dmgraph.Engine(local_rank)
    sg, q_off, q_src = build_workload(rank)
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    upload_s = time.perf_counter() - t0
    nq = q_off.size - 1

    # ---- resident buffers -------------------------------------------------------------------
    d_q_off, d_q_src = eng.to_device(q_off), eng.to_device(q_src)
    d_out_off = eng.device_alloc((nq + 1) * 8)
    rc, need = g.neighbors_batch_dev(dmgraph.BOTH, K_HOPS, 1, nq, d_q_off, d_q_src, d_out_off, None, None, 0)
    cap = int(need)
    d_out_ids, d_out_hop = eng.device_alloc(cap * 4), eng.device_alloc(cap)

    def step_dev():
        rc, n = g.neighbors_batch_dev(dmgraph.BOTH, K_HOPS, 1, nq, d_q_off, d_q_src, d_out_off, d_out_ids,
                                      d_out_hop, cap)
        assert rc == 0 and n == cap
        return eng.last_stats()

    sampler = ClockSampler(local_rank)
    sampler.start()                         # sampled from warm-up to the end of the e2e loops (20 ms period)
    for _ in range(args.warmup):
        eng.flush_l2()
        step_dev()
    eng.profile(True)
    eng.profile_reset()
    barrier()
    dev_ms, wall_ms, launches, stats = 0.0, 0.0, 0, None
    for _ in range(args.steps):
        eng.flush_l2()                      # cold L2 at the start of every timed step
        t0 = time.perf_counter()
        stats = step_dev()
        wall_ms += 1e3 * (time.perf_counter() - t0)
        dev_ms += stats["device_ms"]
        launches += stats["kernel_launches"]
    barrier()
    prof = eng.profile_table()
    eng.profile(False)
    dev_ms_max = allmax(dev_ms)
    total_queries = allsum(float(nq * args.steps))
    value = total_queries / (dev_ms_max * 1e-3)
    edges = stats["edges_traversed"]
    gteps = allsum(float(edges * args.steps)) / (dev_ms_max * 1e-3) / 1e9

    # ---- roofline of the dominant kernel ----------------------------------------------------
    peak, peak_src = peaks()
    kname = max(prof, key=lambda k: prof[k][1]) if prof else "khop"
    kl, kms = prof.get(kname, (1, dev_ms))
    n_src_total = int(q_off[-1])
    n_res = int(stats["results"])
    # B_q = sum over expanded u of (8R + 4 deg(u)) + 4 n_sources + 5 n_results, R = 2
    step_bytes = 16 * stats["rows_expanded"] + 4 * stats["edges_traversed"] + 4 * n_src_total + 5 * n_res
    # bytes no implementation can avoid moving once per step: the CSR the step reads (resolve rows:
    # und_ptr + und_idx), the query batch, the packed result (4 B id + 1 B hop per entry + 8 B row offset)
    csr_bytes = 4 * (sg.n_vertices + 1) + 4 * 2 * int(g.info["n_edges"])
    compulsory = csr_bytes + 4 * (nq + 1) + 4 * n_src_total + 5 * n_res + 8 * (nq + 1)
    step_ms = dev_ms / args.steps
    traffic, tsrc = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get("kernels", tj).get(kname)
            tsrc = tj.get("source")
        except Exception:
            traffic = None
    kernel_ms = kms / kl
    kernel_launches_per_step = kl / args.steps
    # bytes of the DOMINANT kernel, not of the step: the emit kernel reads the bit rows (or hop-byte
    # staging rows) and writes the packed result -- it streams through HBM, so its fraction is a real HBM
    # fraction; the traversal kernel is charged the SURVEY 8d definitional bytes and is cache-resident
    if kname == "khop_emit":
        win_cap = (int(sg.part_off[1]) + 127) & ~127          # largest project, rounded up to 128 vertices
        planes = 1 if K_HOPS <= 1 else (2 if K_HOPS <= 3 else 3)   # binary-coded hop planes of a bit row
        row_bytes = planes * (win_cap // 32) * 4
        kernel_bytes = int(row_bytes * nq + 5 * n_res + 16 * nq)   # bit rows + packed rows + (window, offset) per query
        kernel_resident = False
    else:
        kernel_bytes = int(step_bytes)
        kernel_resident = True
    roofline = {"bound": "hbm", "kernel": kname, "achieved": kernel_bytes / (kernel_ms * kernel_launches_per_step * 1e-3) / 1e9,
                "algorithmic_bytes_per_launch": kernel_bytes, "kernel_cache_resident": kernel_resident,
                "peak": peak, "unit": "GB/s", "traffic": traffic, "traffic_source": tsrc, "peak_source": peak_src,
                "algorithmic_bytes_per_step": int(step_bytes), "kernel_ms_per_launch": kernel_ms,
                "kernel_launches_per_step": kernel_launches_per_step,
                "kernel_share_of_step": kms / dev_ms if dev_ms else None,
                "cache_resident": kernel_resident,
                "compulsory_bytes": int(compulsory),
                "compulsory_breakdown": {"csr": int(csr_bytes), "queries": int(4 * (nq + 1) + 4 * n_src_total),
                                         "packed_result": int(5 * n_res + 8 * (nq + 1))},
                "frac_compulsory": compulsory / (step_ms * 1e-3) / 1e9 / peak,
                "whole_step_ms": step_ms,
                "edges_loaded_per_step": int(stats["edges_read"]),
                "note": "frac = the dominant kernel's own algorithmic bytes over its time. For khop_emit that is bit rows "
                        "read + packed (id, hop) rows written: it streams through HBM and frac is an HBM fraction. For "
                        "the traversal kernel it is the SURVEY 8d definitional bytes (every adjacency entry of every "
                        "expanded vertex, per query) over a 25 MB CSR that lives in L2 and is shared by a project's "
                        "queries: then cache_resident is true and frac is NOT an HBM fraction. frac_compulsory = "
                        "bytes the step cannot avoid (CSR once + queries + packed result) over the WHOLE step's time. "
                        "HBM-bound configs: see configs.c3/c4/c5"}
    roofline["frac"] = roofline["achieved"] / peak

    # ---- e2e: host buffers through the reference-facing call --------------------------------
    h_q_off, p1 = eng.pinned(np.uint32, nq + 1)
    h_q_src, p2 = eng.pinned(np.uint32, q_src.size)
    h_off, p3 = eng.pinned(np.uint64, nq + 1)
    h_ids, p4 = eng.pinned(np.uint32, cap)
    h_hop, p5 = eng.pinned(np.uint8, cap)
    h_q_off[:] = q_off
    h_q_src[:] = q_src
    # (1) dmg_neighbors_batch as it ships: on small projects the batch crosses PCIe as bit rows and the library's
    # host threads expand it into the caller's arrays chunk by chunk; (2) the same call with DMG_HOST_PACKED=1:
    # the packed rows of the device-side emit over PCIe
    def e2e_loop():
        ms, h2d_, d2h_ = 0.0, 0, 0
        for it in range(2 + args.steps):
            barrier()
            t0 = time.perf_counter()
            g.neighbors_batch(dmgraph.BOTH, K_HOPS, 1, h_q_off, h_q_src, out=(h_off, h_ids, h_hop))
            dt = 1e3 * (time.perf_counter() - t0)
            if it >= 2:
                ms += dt
                st = eng.last_stats()
                h2d_, d2h_ = st["h2d_bytes"], st["d2h_bytes"]
        assert int(h_off[-1]) == cap
        return ms, h2d_, d2h_
    call_ms, call_h2d, call_d2h = e2e_loop()
    h_call_check = (h_off.copy(), h_ids.copy(), h_hop.copy()) if (rank == 0 and not args.no_cpu_baseline) else None
    os.environ["DMG_HOST_PACKED"] = "1"
    e2e_ms, h2d, d2h = e2e_loop()
    os.environ.pop("DMG_HOST_PACKED", None)
    e2e_ms_max = allmax(e2e_ms)
    call_ms_max = allmax(call_ms)
    e2e_value = total_queries / (e2e_ms_max * 1e-3)

    # the same answer as bit rows (dmg_neighbors_batch_bits: two hop-code planes over the project's
    # vertices, 512 B per query) and their expansion to the same (off, ids, hop) arrays by host threads
    # (dmg_bits_decode): 20x fewer PCIe bytes, the arrays the packed call delivers
    planes, words, _w, _c, _b = g.neighbors_batch_bits(dmgraph.BOTH, K_HOPS, 1, h_q_off[:2], h_q_src[:int(h_q_off[1])])
    h_bits, p6 = eng.pinned(np.uint32, nq * planes * words)
    h_win, p7 = eng.pinned(np.uint32, nq * 2)
    h_cnt, p8 = eng.pinned(np.uint32, nq)
    # (a) the bit-row call alone, (b) the call followed by the host expansion: two loops, because a consumer of
    # (a) does not run the expansion in between (its 650 MB of streaming stores were seen to slow the next call)
    only_ms = 0.0
    for it in range(2 + args.steps):
        barrier()
        t0 = time.perf_counter()
        g.neighbors_batch_bits(dmgraph.BOTH, K_HOPS, 1, h_q_off, h_q_src, out=(h_win.reshape(nq, 2), h_cnt, h_bits))
        if it >= 2:
            only_ms += 1e3 * (time.perf_counter() - t0)
    bits_ms, dec_ms, bits_d2h, bits_dev_ms = 0.0, 0.0, 0, 0.0
    for it in range(2 + args.steps):
        barrier()
        t0 = time.perf_counter()
        g.neighbors_batch_bits(dmgraph.BOTH, K_HOPS, 1, h_q_off, h_q_src, out=(h_win.reshape(nq, 2), h_cnt, h_bits))
        t1 = time.perf_counter()
        _st = eng.last_stats()
        bits_d2h = _st["d2h_bytes"]
        if it >= 2:
            bits_dev_ms += _st["device_ms"]
        h_off[:] = 0
        dmgraph.bits_decode(planes, words, h_win, h_cnt, h_bits, out=(h_off, h_ids, h_hop))
        t2 = time.perf_counter()
        if it >= 2:
            bits_ms += 1e3 * (t1 - t0)
            dec_ms += 1e3 * (t2 - t1)
    assert int(h_off[-1]) == cap
    both_ms_max = allmax(bits_ms + dec_ms)
    only_ms_max = allmax(only_ms)
    bits_ms_max = allmax(bits_ms)
    h_bits_check = (h_off.copy(), h_ids.copy(), h_hop.copy()) if (rank == 0 and not args.no_cpu_baseline) else None
    clocks = sampler.stop()

    # ---- parity of what was just timed: last step's resident result vs the oracle --------------
    parity = None
    csr_line = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import checks
        off_h = eng.to_host(d_out_off, np.uint64, nq + 1)
        ids_h = eng.to_host(d_out_ids, np.uint32, cap)
        hop_h = eng.to_host(d_out_hop, np.uint8, cap)
        if world == 1:
            csr_line, o, res = cpu_csr(sg, q_off, q_src)
            assert np.array_equal(off_h, res[0]) and np.array_equal(ids_h, res[1]) and np.array_equal(hop_h, res[2]), \
                "bench: the timed result differs from the oracle"
            assert all(np.array_equal(x, y) for x, y in zip(h_bits_check, res[:3])), \
                "bench: the decoded bit rows differ from the oracle"
            assert all(np.array_equal(x, y) for x, y in zip(h_call_check, res[:3])), \
                "bench: the rows of the host-pointer call differ from the oracle"
            parity = {"rows": int(nq), "against": "oracle csr, whole batch (ids, hops, offsets bit-exact), both the "
                                                  "resident packed result, the rows delivered by the e2e call (dmg_neighbors_batch) and the "
                                                  "rows of dmg_neighbors_batch_bits + dmg_bits_decode"}
        else:
            from oracle import oracle as ora
            o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
            rows = np.sort(np.random.Generator(np.random.PCG64(7)).choice(nq, min(nq, args.parity_rows), replace=False))
            n_chk = checks.check_khop_rows(o, 2, K_HOPS, 1, q_off, q_src, off_h, ids_h, hop_h, rows=rows)
            parity = {"rows": int(n_chk), "against": "oracle csr, seeded row sample of rank 0's batch"}


    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": WORKLOAD},
            "details": {"l2": "flushed before every timed step (write of 2x L2 size)",
                        "timing": "CUDA events on libdmgraph's launch stream, summed over steps, max over ranks",
                        "results_per_step": n_res, "sources_per_step": n_src_total, "graph_upload_s": upload_s},
            "gteps": gteps, "edges_traversed_per_step": int(edges), "edges_read_per_step": int(stats["edges_read"]),
            "wall_ms_per_step": allmax(wall_ms) / args.steps,
            "roofline": roofline,
            "e2e": {"value": total_queries / (call_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(call_h2d),
                    "d2h_bytes_per_step": int(call_d2h), "ms_per_step": call_ms_max / args.steps,
                    "host_buffers": "pinned",
                    "call": "dmg_neighbors_batch(host q_off/q_src -> host out_off/out_ids/out_hop): the batch crosses "
                            "PCIe as bit rows and the library's host threads expand them into the caller's arrays"},
            "e2e_bits_decode": {"value": total_queries / (both_ms_max * 1e-3), "unit": UNIT,
                    "d2h_bytes_per_step": int(bits_d2h), "ms_per_step": both_ms_max / args.steps,
                    "call": "dmg_neighbors_batch_bits + dmg_bits_decode as two calls (no overlap of copy and expansion)",
                    "transfer_ms_per_step": bits_ms_max / args.steps,
                    "transfer_device_ms_per_step": bits_dev_ms / args.steps,
                    "decode_ms_per_step": allmax(dec_ms) / args.steps, "decode_threads": os.cpu_count()},
            "e2e_packed": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms_max / args.steps,
                           "call": "dmg_neighbors_batch with DMG_HOST_PACKED=1 (packed (id, hop) rows over PCIe)"},
            "e2e_bits_only": {"value": total_queries / (only_ms_max * 1e-3), "unit": UNIT,
                              "ms_per_step": only_ms_max / args.steps, "d2h_bytes_per_step": int(bits_d2h),
                              "call": "dmg_neighbors_batch_bits alone: a host that walks the neighbours reads the "
                                      "planes directly"},
            "gpu_launches": int(launches), "clocks": clocks,
            "parity_checked": parity["rows"] if parity else 0, "parity": parity,
            "kernels": {k: {"launches": int(v[0]), "ms_per_launch": v[1] / v[0]} for k, v in prof.items()}}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        _qps, cb = cpu_reference(sg, q_off, q_src, budget_s=args.cpu_budget)
        line["cpu_baseline"] = cb
        line["cpu_baseline_csr"] = csr_line
    elif rank == 0:
        line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port",
                                "sample": "measured at N=1 only"}
    for p in (p1, p2, p3, p4, p5, p6, p7, p8):
        eng.pinned_free(p)
    g.release()
    for d in (d_q_off, d_q_src, d_out_off, d_out_ids, d_out_hop):
        eng.device_free(d)

    # ---- the HBM-bound configs (N = 1): each block carries its own roofline, e2e and oracle check ----
    if rank == 0 and world == 1 and args.configs:
        import bench_configs as bc
        blocks = {}
        for name in [c for c in args.configs.split(",") if c]:
            t0 = time.perf_counter()
            try:
                if name == "c3":
                    blocks[name] = bc.run_c3(eng, args.c3_names)
                elif name == "c4":
                    blocks[name] = bc.run_c4(eng)
                elif name == "c5":
                    blocks[name] = bc.run_c5(eng, args.c5_vertices, args.c5_edges)
                else:
                    blocks[name] = {"error": "unknown config"}
            except Exception as ex:      # the headline line must survive a failing block, and show it
                blocks[name] = {"error": "%s: %s" % (type(ex).__name__, ex)}
            blocks[name]["block_wall_s"] = time.perf_counter() - t0
        line["configs"] = blocks
    # ---- N > 1: the path's one exchange step, the partitioned closure over all N devices (rank 0 drives
    # them through dmg_dist_* in ONE process, as the single-JVM host would; the other ranks wait) ----------
    if args.closure_vertices < 0:
        args.closure_vertices, args.closure_edges = 2_500_000 * world, 62_500_000 * world
    if world > 1 and args.closure_vertices:
        import torch
        barrier()
        torch.cuda.synchronize()
        # the other ranks wait on the CPU (gloo): an NCCL barrier would park a spinning kernel on every GPU
        # that rank 0 is about to drive (measured: the closure ran 4x slower beside it)
        cpu_group = dist.new_group(backend="gloo")
        if rank == 0:
            import bench_configs as bc
            t0 = time.perf_counter()
            try:
                blk = bc.run_dist_closure(world, args.closure_vertices, args.closure_edges)
            except Exception as ex:
                blk = {"error": "%s: %s" % (type(ex).__name__, ex)}
            blk["block_wall_s"] = time.perf_counter() - t0
            line["closure"] = blk
        dist.barrier(group=cpu_group)
    if rank == 0:
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
