#!/usr/bin/env python
"""bench.py — embedding lookups/s (forward + backward fused with the SGD row update) on N B200s.

Contract (one JSON line on rank 0):
  python bench.py --gpus N --steps K --warmup W            # product arm (libffemb.so, sm_100a)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU tasks on host cores
A "step" is one pass of the hot path over one batch of synthetic input:
  forward gather-pool of every table  ->  [N>1: exchange of pooled rows / output-grads]
  ->  backward sort + segmented reduce + SGD row update.
Default workload = BASELINE.json configs[2] (MLPerf Criteo-Terabyte shape: 26 tables up to 40M
rows x 128, multi-hot 1..20, batch 64K) — the config the north star's roofline and scaling
targets are quoted on; it fits one B200 (96 GB of tables).  --workload 1|2|4|5 selects the
others.  The table set and the global batch are fixed as N grows ("scaling": "strong").  Placement at
N > 1: --placement balanced (the default and the headline: cost-aware, heavy tables split by columns,
sharding.plan_shards — the plan is printed in config.shard_plan and named in config.parallelism) or
--placement modulo, table t whole on GPU t % N like the reference's strategy generator
(examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256).  At N > 1 the whole pipelined step is replayed
from one CUDA graph per index batch (--graph off: eager), and one un-timed parity step runs first
("parity_ok" in the line).  --workload 4 (row-wise, 8 GPUs): route kernel + the owners' padded backward.

value   : device-resident inputs, CUDA events on the launch stream, max over ranks.
e2e     : the same step through the host-side operator mirror with HOST (pinned) buffers:
          indices + output-grads copied H2D, pooled outputs copied D2H inside the timed region.
roofline: the dominant kernel phase (per-phase CUDA events recorded inside the same timed region
          by libffemb's profiling hooks), algorithmic bytes / duration vs MEASURED_PEAKS.json.
cpu_baseline / --impl reference: oracle/_ref (the reference's own embedding.cc, AVX2) where it
          can run the workload, else the oracle port, on a bounded sample, all host cores.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "embedding_lookups_per_sec_fwd_bwd"
UNIT = "lookups/s"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", type=int, default=3, help="BASELINE config number 1..5")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink rows/batch (debug only)")
    ap.add_argument("--pool", type=int, default=16, help="distinct index batches cycled (SURVEY 8d: >= 16)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--placement", default="balanced", choices=["modulo", "balanced"],
                    help="N>1: modulo = the reference's t %% G; balanced = greedy by measured per-lookup cost")
    ap.add_argument("--groups", type=int, default=0,
                    help="N>1, fused exchange: pipeline groups per rank (1 = forward(all), push(all), backward(all)); "
                         "0 = auto: 2 up to 4 GPUs, 1 beyond (measured, profiles/r2_scaling.md: at 8 GPUs a rank's "
                         "share is launch-latency-bound and a second group only doubles the launches)")
    ap.add_argument("--trace", action="store_true",
                    help="N>1: after the timed region, one more step with CUDA events around every piece of the "
                         "pipelined schedule; per-rank timelines go to stderr (tuning aid)")
    ap.add_argument("--no-split", action="store_true",
                    help="N>1: keep every table whole (no column split of heavy tables)")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N>1: fused = pooled rows stored straight into the peer's Concat input by the "
                         "gather-pool kernel (NVLink peer memory); nccl = all_to_all_single + pack/unpack copies")
    ap.add_argument("--backward", default="sorted", choices=["sorted", "atomic"],
                    help="atomic = dense-grad atomicAdd comparison (not the product path)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="run the backward's index sort in stream order instead of on a side stream "
                         "under the forward pass / exchange (it depends on the indices only)")
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"],
                    help="N=1: replay the step from a CUDA graph (StepGraph); auto = the launch-bound small "
                         "configs (<= 2M lookups per step: cfg1, cfg2), where it is worth 1.6-1.9x")
    ap.add_argument("--l2-persist", type=int, default=-1,
                    help="experiment: cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, MB) before the run "
                         "(-1: leave the device as it is)")
    ap.add_argument("--aggr", type=int, default=0, help="override the workload's aggregation (21 SUM, 22 AVG literal, 122 MEAN): experiments")
    ap.add_argument("--rowwise", default="routed", choices=["routed", "legacy"],
                    help="cfg4: routed = bucketise by owner + remote stores + owners reduce their own lookups (SURVEY 8e); "
                         "legacy = round 1's two all-gathers + whole-batch backward on every owner")
    ap.add_argument("--rows-cap", type=int, default=0, help="cap rows per table (profiling under ncu only)")
    ap.add_argument("--cpu-batch", type=int, default=65536,
                    help="batch of the CPU sample (capped at the workload's batch)")
    ap.add_argument("--cpu-rows-cap", type=int, default=2_000_000)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
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
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1])), pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# workload on the device
# ---------------------------------------------------------------------------------------------
def placement(wl, world, kind):
    """table -> rank (sharding.place_tables).  modulo = the reference's strategy generator (t % G,
    examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256); balanced = greedy by measured cost."""
    pkg = importlib.import_module("dlrm-flexflow_b200")
    return pkg.sharding.place_tables(wl.rows, wl.bags, wl.dim, world, kind)


def zipf_indices(torch, gen, rows, shape, alpha, table_id, dev):
    """Zipf(alpha) ranks via inverse CDF (exact head of 2^20 ranks, continuous tail), then a fixed
    affine rank->row bijection so hot rows are not address-adjacent (synth.gen_indices)."""
    import math
    n = shape[0] * shape[1]
    head = min(rows, 1 << 20)
    w = 1.0 / torch.arange(1, head + 1, dtype=torch.float64, device=dev).pow(alpha)
    cdf = torch.cumsum(w, 0)
    tail = 0.0
    if rows > head:
        tail = ((head + 0.5) ** (1 - alpha) - (rows + 0.5) ** (1 - alpha)) / (alpha - 1)
    total = cdf[-1].item() + tail
    u = torch.rand(n, dtype=torch.float64, device=dev, generator=gen) * total
    r = torch.searchsorted(cdf, u, right=True)
    if tail > 0:
        m = u >= cdf[-1]
        x = ((head + 0.5) ** (1 - alpha) - (u[m] - cdf[-1]) * (alpha - 1)).pow(1.0 / (1 - alpha)) - 0.5
        r[m] = x.long().clamp(head, rows - 1)
    r = r.clamp_(max=rows - 1)
    a = (1234567 + 2 * table_id * 7919) | 1
    while math.gcd(a, rows) != 1:
        a += 2
    a %= rows
    return ((r * a + 977 * (table_id + 1)) % rows).view(shape).contiguous()


class DeviceWorkload:
    """This rank's parts (sharding.Part: whole tables, or column ranges of split ones) with their
    synthetic inputs.  Index batches are seeded per TABLE, so the ranks that share a split table see the
    same indices."""

    def __init__(self, torch, pkg, wl, parts, dev, pool, seed, num_groups=None):
        self.torch, self.wl, self.parts, self.dev = torch, wl, parts, dev
        self.owned = [p.table for p in parts]
        self.widths = [p.width for p in parts]
        g = torch.Generator(device=dev)
        self.W, self.idx, self.outgrad, self.out, self.uniq = [], [], [], [], []
        for p in parts:
            t = p.table
            R, L = wl.rows[t], wl.bags[t]
            a = (1.0 / R) ** 0.5  # examples/cpp/DLRM/dlrm.cc:44-45
            g.manual_seed(seed + t + 1000 * p.col0)
            self.W.append(torch.empty((R, p.width), dtype=torch.float32, device=dev).uniform_(-a, a, generator=g))
            g.manual_seed(seed + t)
            batches = []
            for k in range(pool):
                if wl.dist == "zipf":
                    batches.append(zipf_indices(torch, g, R, (wl.batch, L), wl.zipf_alpha, t, dev))
                else:
                    batches.append(torch.randint(0, R, (wl.batch, L), dtype=torch.int64, device=dev, generator=g))
            self.idx.append(batches)
            self.uniq.append([int(torch.unique(b).numel()) for b in batches])
            self.outgrad.append(torch.empty((wl.batch, p.width), dtype=torch.float32, device=dev).uniform_(-1, 1, generator=g))
            self.out.append(torch.empty((wl.batch, p.width), dtype=torch.float32, device=dev))
        self.pool = pool
        # one collection per pipeline group (a group's parts share one width)
        # (the number of groups is the PLAN's: a rank may hold parts of fewer groups than its peers)
        self.K = num_groups if num_groups is not None else 1 + max((p.group for p in parts), default=0)
        self.group_of = [[i for i, p in enumerate(parts) if p.group == k] for k in range(self.K)]
        self.colls = [pkg.EmbeddingCollection([self.W[i] for i in ids], [wl.bags[parts[i].table] for i in ids], wl.aggr)
                      if ids else None for ids in self.group_of]
        self.coll = self.colls[0] if self.K == 1 else None

    def lookups(self):
        """lookups this rank walks (a column-split table's lookups are walked by every rank of the split)"""
        return sum(self.wl.batch * self.wl.bags[t] for t in self.owned)

    def frac(self, i):
        return self.widths[i] / float(self.wl.dim)

    def bytes_per_step(self):
        """SURVEY §8d algorithmic bytes for the owned tables, averaged over the pool:
        fwd = 8N + 4DN + 4D*T*B ; bwd = 8N + 4D*T*B + 8D*U."""
        B = self.wl.batch
        fwd = bwd = U = 0.0
        for i, t in enumerate(self.owned):
            w, n = self.widths[i], B * self.wl.bags[t]
            u = sum(self.uniq[i]) / len(self.uniq[i])
            fwd += 8 * n + 4 * w * n + 4 * w * B
            bwd += 8 * n + 4 * w * B + 8 * w * u
            U += u * self.frac(i)
        return fwd, bwd, U

    def compulsory_bytes_per_step(self):
        """The cache-aware count: §8d credits every lookup with a row read from HBM (4*D*N), which
        holds only for tables far larger than L2 with no repeats.  Compulsory traffic reads each
        DISTINCT touched row once per step (4*D*U; a table that fits L2 can at most be credited once
        per step), and the backward then finds it again only if it is still resident, so its
        read-modify-write stays at 8*D*U:  fwd = 8N + 4D*U + 4D*T*B ; bwd = 8N + 4D*T*B + 8D*U."""
        B = self.wl.batch
        fwd = bwd = 0.0
        for i, t in enumerate(self.owned):
            w, n = self.widths[i], B * self.wl.bags[t]
            u = sum(self.uniq[i]) / len(self.uniq[i])
            fwd += 8 * n + 4 * w * u + 4 * w * B
            bwd += 8 * n + 4 * w * B + 8 * w * u
        return fwd, bwd


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(reference_arm(args, max(world, args.gpus))), flush=True)
        return 0

    import torch
    pkg = importlib.import_module("dlrm-flexflow_b200")
    pkg.lib()  # fails loudly when libffemb.so is missing: there is no fallback
    assert torch.cuda.is_available(), "bench.py needs a CUDA device"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    if args.l2_persist >= 0:
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        rc = rt.cudaDeviceSetLimit(6, ctypes.c_size_t(args.l2_persist << 20))   # cudaLimitPersistingL2CacheSize
        got = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(got), 6)
        print(f"[l2-persist] cudaDeviceSetLimit rc={rc}, limit now {got.value >> 20} MB, "
              f"max {torch.cuda.get_device_properties(dev).L2_cache_size >> 20} MB L2", file=sys.stderr)
    wl = pkg.synth.config(args.workload, args.scale)
    if args.aggr:
        wl.aggr = args.aggr
        wl.name += f"-aggr{args.aggr}"
    if args.rows_cap:
        wl.rows = [min(r, args.rows_cap) for r in wl.rows]
        wl.name += f"-rowscap{args.rows_cap}"
    if args.workload == 4 and world > 1:
        return run_rowwise(args, torch, pkg, dist, wl, rank, world, dev)
    # shard plan: whole tables by t % N (the reference's generator) or by cost, heavy tables split by
    # columns, this rank's parts dealt into pipeline groups (sharding.plan_shards); the NCCL comparison
    # form moves whole tables only
    if args.groups <= 0:
        args.groups = 2 if world <= 4 else 1
    fused = world > 1 and args.exchange == "fused"
    plan = pkg.sharding.plan_shards(wl.rows, wl.bags, wl.dim, world, args.placement,
                                    split=fused and not args.no_split, groups=args.groups if fused else 1)
    own = [next(p.rank for p in plan if p.table == t) for t in range(wl.num_tables)]
    mine = sorted([p for p in plan if p.rank == rank], key=lambda p: (p.group, p.table, p.col0))
    owned = [p.table for p in mine]
    dw = DeviceWorkload(torch, pkg, wl, mine, dev, args.pool, seed=1234, num_groups=pkg.sharding.num_groups(plan))
    ex = fx = sstep = None
    if world > 1 and args.exchange == "nccl":
        ex = pkg.exchange.TableWiseExchange(wl.num_tables, own, wl.batch, wl.dim, rank, world, dev)
    elif world > 1:
        fx = pkg.exchange.FusedTableWiseExchange(wl.num_tables, own, wl.batch, wl.dim, rank, world, dev, parts=plan)
        g = torch.Generator(device=dev).manual_seed(4321 + rank)
        fx.dx.uniform_(-1, 1, generator=g)
        dw.outgrad = fx.outgrads  # the owners' output-grad buffers are the peers' push targets
        sstep = pkg.exchange.ShardedStep(fx, dw.colls, dev)
    stream = torch.cuda.current_stream()

    wgrads = None
    if args.backward == "atomic":  # the measured comparison: dense grad + atomics + dense update
        wgrads = [torch.zeros_like(w) for w in dw.W]

    overlap = not args.no_overlap and args.backward == "sorted" and (dw.coll is not None or fx is not None)
    side = torch.cuda.Stream(device=dev) if overlap and fx is None else None

    def presort(idx):
        """keygen + radix sort on the side stream: needs the indices only, so it runs under the
        forward pass and the exchange; the main stream joins before the segmented sum."""
        if overlap and fx is None:
            side.wait_stream(stream)  # the previous step's reduction is done with the sort buffers
            dw.coll.backward_prepare(idx, stream=side)

    def step(k):
        p = k % dw.pool
        idx = [b[p] for b in dw.idx]
        presort(idx)
        if fx is not None:
            # gather-pool kernels that store into the peers' Concat inputs, output-grad push over NVLink and
            # the owners' fused backward, pipelined over the plan's groups on three streams (ShardedStep)
            sstep.run([[idx[i] for i in ids] for ids in dw.group_of], wl.lr, overlap=overlap)
            return
        if dw.coll is not None:
            dw.coll.forward(idx, dw.out)
        if ex is not None:
            ex.forward(dw.out)            # pooled rows -> interaction input of each sample's rank
            ex.backward(dw.outgrad)       # output-grads back to the table owners
        if dw.coll is None:
            return
        if wgrads is None:
            if overlap:
                stream.wait_stream(side)
            dw.coll.backward_sgd(idx, dw.outgrad, wl.lr, prepared=overlap)
        else:
            for t, w in enumerate(dw.W):
                wgrads[t].zero_()
                pkg.Embedding.backward_kernel(idx[t].data_ptr(), dw.outgrad[t].data_ptr(), wgrads[t].data_ptr(),
                                              wl.bags[owned[t]], wl.dim, wl.batch, wl.aggr)
                w.add_(wgrads[t], alpha=-wl.lr)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    use_graph = (args.graph == "on" or (args.graph == "auto" and wl.lookups <= 2_000_000)) and \
        world == 1 and args.backward == "sorted"
    # N > 1, fused exchange: a rank's share of a step is ~50 launches of 5-100 us on three streams; the
    # whole pipelined step (flag kernels included) is replayed from one CUDA graph per index batch
    use_sgraph = fx is not None and overlap and args.graph != "off" and dw.pool % 2 == 0
    per_step_launches, eager_phases = 0, None
    parity = None
    if fx is not None:
        parity = sharded_parity_step(torch, pkg, dist, wl, dw, fx, sstep, plan, rank, world, dev)
    if use_sgraph:
        for k in range(args.warmup):
            step(k)
        barrier()
        pkg.profile_enable(True)
        pkg.profile_collect()
        l0 = pkg.launch_count()
        for k in range(args.steps):
            step(k)
        barrier()
        per_step_launches = (pkg.launch_count() - l0) / args.steps
        eager_phases = pkg.profile_collect()
        pkg.profile_enable(False)
        sgraphs = [sstep.capture([[dw.idx[i][p] for i in ids] for ids in dw.group_of], wl.lr, parity=p % 2)
                   for p in range(dw.pool)]
        barrier()

        def step(k):  # noqa: F811
            sgraphs[k % dw.pool].replay()
        use_graph = True
    elif use_graph:
        # phases and the launch count come from an eager profiled pass of the same kernels; the timed
        # region then replays one captured graph per index batch of the pool
        for k in range(args.warmup):
            step(k)
        barrier()
        pkg.profile_enable(True)
        pkg.profile_collect()
        l0 = pkg.launch_count()
        for k in range(args.steps):
            step(k)
        barrier()
        per_step_launches = (pkg.launch_count() - l0) / args.steps
        eager_phases = pkg.profile_collect()
        pkg.profile_enable(False)
        graphs = [pkg.StepGraph(dw.coll, wl.batch, wl.lr, idx=[b[p] for b in dw.idx], outs=dw.out,
                                out_grads=dw.outgrad) for p in range(dw.pool)]
        overlap = False

        def step(k):  # noqa: F811
            graphs[k % dw.pool].graph.replay()

    for k in range(args.warmup):
        step(k)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    pkg.profile_enable(True)
    pkg.profile_collect()
    l0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.profiler.start()  # ncu --profile-from-start off sees exactly the timed region
    e0.record(stream)
    for k in range(args.steps):
        step(args.warmup + k)
    e1.record(stream)
    barrier()
    torch.cuda.profiler.stop()
    ms_total = e0.elapsed_time(e1)
    launches = pkg.launch_count() - l0
    phase_ms, phase_launch = pkg.profile_collect()
    pkg.profile_enable(False)
    if use_graph:
        launches = int(round(per_step_launches * args.steps))  # kernels inside the replayed graphs
        phase_ms, phase_launch = eager_phases
    clocks = sampler.stop() if sampler else None
    if dist is not None:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = t.item()
    ms_step = ms_total / args.steps
    total_lookups = wl.lookups
    if args.trace and sstep is not None:
        for rep_ in range(2):
            barrier()
            sstep.trace = []
            kk = args.warmup + args.steps + rep_
            sstep.run([[dw.idx[i][kk % dw.pool] for i in ids] for ids in dw.group_of], wl.lr, overlap=overlap)
            torch.cuda.synchronize()
            tl = sstep.trace_ms()
            sstep.trace = None
        for r in range(world):
            barrier()
            if r == rank:
                print(f"[trace rank {rank}] " + " ".join(f"{l}:{a:.3f}-{b:.3f}" for l, a, b in tl), file=sys.stderr, flush=True)

    # ---- e2e: host buffers, copies inside the timed region ----------------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(torch, pkg, args, wl, dw, ex, fx, dist, dev, barrier, total_lookups)

    # ---- roofline of the dominant kernel phase ----------------------------------------------
    peak, peak_src = peaks()
    fwd_b, bwd_b, U = dw.bytes_per_step()
    kernel_phases = {k: v / args.steps for k, v in phase_ms.items() if v > 0}
    D, B = wl.dim, wl.batch
    n_own, T_own = dw.lookups(), len(owned)
    # tables whose gradient fits on chip take the bag-major "small" kernels, the rest the sorted path
    path = pkg.lib().ffemb_backward_path
    small = [i for i, t in enumerate(owned) if path(wl.rows[t], wl.bags[t], dw.widths[i], B)]
    n_s = sum(B * wl.bags[owned[i]] for i in small)
    umean = lambda i: sum(dw.uniq[i]) / len(dw.uniq[i])
    sorted_ids = [i for i in range(T_own) if i not in small]
    alg = {"forward": fwd_b, "atomic": bwd_b, "sort": 0, "reduce_up": 0,
           "reduce1": sum(4 * dw.widths[i] * B + 8 * dw.widths[i] * umean(i) + 8 * B * wl.bags[owned[i]] for i in sorted_ids),
           "exchange": (8 * D * wl.num_tables * fx.local_batch) if fx is not None else 0,
           "keygen": 8 * (n_own - n_s),
           "small": sum(8 * B * wl.bags[owned[i]] + 4 * dw.widths[i] * B + 8 * dw.widths[i] * umean(i) for i in small)}
    # dominant kernel = the longest phase that moves algorithmic bytes (sort / keygen / upper levels are
    # overhead phases: no bytes are credited to them, and at N > 1 the sort's span is stretched by the
    # exchange it runs under)
    # ("small" is excluded too: those kernels are forked onto a side stream and run UNDER the sort / level 1,
    # so their event span is not busy time)
    cand = {k: v for k, v in kernel_phases.items() if alg.get(k, 0) > 0 and k in ("forward", "reduce1", "atomic")}
    # cache-aware (compulsory) bytes: the forward reads each DISTINCT row once; level 1 / the atomic
    # comparison are already counted that way by §8d (one read-modify-write per distinct row)
    cfwd_b, cbwd_b = dw.compulsory_bytes_per_step()
    comp = dict(alg)
    comp["forward"] = cfwd_b
    # dominant kernel = the longest such phase whose credited bytes DRAM could actually have delivered
    # (compulsory count within the 1.2 x peak sanity bound); a phase beyond it ran out of cache and says
    # nothing about HBM
    order = sorted(cand, key=cand.get, reverse=True)
    sane = [k for k in order if comp.get(k, 0) / (cand[k] * 1e-3) / 1e9 <= 1.2 * peak]
    dom = sane[0] if sane else (order[0] if order else (max(kernel_phases, key=kernel_phases.get) if kernel_phases else "forward"))
    dom_ms = kernel_phases.get(dom, ms_step)
    achieved = alg.get(dom, 0) / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    achieved_c = comp.get(dom, 0) / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    # §8d's figure is the headline unless on-chip hits inflate it beyond what DRAM can deliver (the
    # B200_PROFILING.md sanity bound: a good kernel may read a little above 1.0, far above means the bytes
    # credited were not moved); then the compulsory count is the figure and the line says so
    basis = "survey-8d algorithmic bytes"
    frac = achieved / peak
    if frac > 1.2:
        achieved, frac = achieved_c, achieved_c / peak
        basis = "compulsory bytes (the 8d credit exceeded 1.2 x peak: rows served from L2/L1, not HBM)"
    traffic = load_traffic(dom) if world == 1 and not args.rows_cap and args.workload == 3 else None
    step_dram = load_traffic("step_total") if world == 1 and not args.rows_cap and args.workload == 3 else None
    gbs = lambda b: b / (ms_step * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(frac, 4), "frac_basis": basis,
            "frac_8d": round(alg.get(dom, 0) / (dom_ms * 1e-3) / 1e9 / peak, 4) if dom_ms > 0 else None,
            "frac_compulsory": round(achieved_c / peak, 4),
            "traffic": traffic,
            "peak_source": peak_src, "ms_per_launch": round(dom_ms, 4),
            "algorithmic_bytes_per_launch": int(alg.get(dom, 0)),
            "compulsory_bytes": int(comp.get(dom, 0)),
            "phases_ms_per_step": {k: round(v, 4) for k, v in kernel_phases.items()},
            "step_algorithmic_bytes": int(fwd_b + bwd_b),
            "step_compulsory_bytes": int(cfwd_b + cbwd_b),
            "step_achieved_gbs": round(gbs(fwd_b + bwd_b), 1),
            "step_frac_of_peak": round(gbs(fwd_b + bwd_b) / peak, 4),
            "step_frac_of_8000": round(gbs(fwd_b + bwd_b) / 8000.0, 4),
            "step_frac_compulsory": round(gbs(cfwd_b + cbwd_b) / peak, 4),
            "step_frac_compulsory_of_8000": round(gbs(cfwd_b + cbwd_b) / 8000.0, 4),
            # DRAM bytes of one whole step as ncu counted them (sum over every kernel of the step, the
            # radix-sort passes included; profiles/traffic.json, captured at this size)
            "step_dram_bytes": step_dram,
            "step_frac_dram": round(gbs(step_dram) / peak, 4) if step_dram else None}
    if roof["frac"] > 1.2:  # no phase within the bound (everything on chip): the line says so, frac is capped
        roof["frac_uncapped"], roof["frac"] = roof["frac"], 1.2
        roof["frac_basis"] += "; CAPPED at the 1.2 sanity bound: this launch ran out of cache, not HBM"

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0
    line = {
        "metric": METRIC, "value": total_lookups / (ms_step * 1e-3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name, "tables": wl.num_tables, "dim": wl.dim, "batch": wl.batch,
                   "bags": wl.bags, "rows_total": int(sum(wl.rows)), "lookups_per_step": total_lookups,
                   "aggr": wl.aggr, "index_dist": wl.dist, "lr": wl.lr,
                   "parallelism": ("single" if world == 1 else
                                   f"table-wise x{world} ({args.placement} placement"
                                   + (f", heavy tables split by columns, {fx.K} pipeline groups per rank" if fx is not None else "")
                                   + f"), exchange={args.exchange}"),
                   "table_owner": own if world > 1 else None,
                   "shard_plan": [[p.table, p.col0, p.width, p.rank, p.group] for p in plan] if world > 1 else None,
                   "backward": args.backward, "sort_overlapped_on_side_stream": bool(overlap),
                   "cuda_graph": bool(use_graph),
                   "cache": f"inputs larger than L2: {args.pool} index batches cycled, "
                            f"{sum(wl.rows) * wl.dim * 4 / 1e9:.1f} GB of tables, random rows",
                   "unique_rows_per_step": int(U)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
    }
    if parity is not None:
        line["parity_ok"] = bool(parity["ok"])
        line["parity"] = parity
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args, wl, steps=3, warmup=1)
        ref_gpu = reference_gpu_leg(torch, wl, dw, args)
        if ref_gpu is not None:
            line["reference_on_b200"] = ref_gpu
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def rowwise_parity_step(torch, dist, wl, rc, idx0, dx, out, rank, world, dev):
    """One un-timed step of the row-wise path held against a recomputation (the check may use
    collectives and torch reads of peer memory; the product path uses neither):
      forward : pooled rows bit-exact against the owners' shard rows added in j order;
      route   : the owner's slot arrays == (global index - first row of the shard, -1 elsewhere) and the
                routed output-grad rows == the senders' dx rows, bit for bit;
      update  : touched rows of this shard == old - lr * (sum of their gradients) within 2e-6 relative
                (fp32 summation order / FMA contraction), a sample of untouched rows bit-identical."""
    T, D, ext, lr = wl.num_tables, wl.dim, rc.ext, wl.lr
    n = rc.hi - rc.lo
    res = {"forward_bit_exact": True, "routed_slots_bit_exact": True, "routed_rows_bit_exact": True,
           "updated_rows_ok": True, "untouched_rows_bit_identical": True, "worst_rel_err": 0.0}
    full_idx = []
    for t in range(T):
        pad = torch.full((ext, wl.bags[t]), -1, dtype=torch.int64, device=dev)
        pad[:n] = idx0[t]
        full = torch.empty((world * ext, wl.bags[t]), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(full, pad)
        full_idx.append(full)
    padg = torch.zeros((ext, T * D), device=dev)
    padg[:n] = dx
    dx_full = torch.empty((world * ext, T * D), device=dev)
    dist.all_gather_into_tensor(dx_full, padg)
    outs = rc.forward(idx0, out)
    torch.cuda.synchronize()
    for t in range(T):
        L, rps = wl.bags[t], rc.rps[t]
        bufs = [rc.handles[t].get_buffer(g, (rps + 1, D), torch.float32) for g in range(world)]
        want = torch.zeros((n, D), device=dev)
        for j in range(L):
            own = idx0[t][:, j] // rps
            loc = idx0[t][:, j] - own * rps
            row = torch.empty((n, D), device=dev)
            for g in range(world):
                m = own == g
                row[m] = bufs[g][loc[m]]
            want = want + row
            if wl.aggr == 22:
                want = want / float(L)
        if wl.aggr == 122:
            want = want * (1.0 / L)
        res["forward_bit_exact"] &= bool(torch.equal(outs[t], want))
    # rows of this shard the step will touch, and their values before it
    olds, keys, rnd_rows, rnd_old = [], [], [], []
    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    for t in range(T):
        rps, full = rc.rps[t], full_idx[t]
        mine = (full >= rank * rps) & (full < (rank + 1) * rps)
        u, inv = torch.unique(full[mine] - rank * rps, return_inverse=True)
        keys.append((mine, u, inv))
        olds.append(rc.shards[t][u].clone())
        r = torch.randint(0, max(1, rc.local_rows[t]), (4096,), device=dev, generator=gen)
        r = r[~torch.isin(r, u)]
        rnd_rows.append(r)
        rnd_old.append(rc.shards[t][r].clone())
    dist.barrier()
    rc.backward_sgd(idx0, dx, lr)
    torch.cuda.synchronize()
    dist.barrier()
    for t in range(T):
        L, rps, full = wl.bags[t], rc.rps[t], full_idx[t]
        mine, u, inv = keys[t]
        if not rc.legacy:
            want_slots = torch.where(mine, full - rank * rps, torch.full_like(full, -1))
            res["routed_slots_bit_exact"] &= bool(torch.equal(rc.slot_views[t], want_slots))
            bags_here = mine.any(dim=1)
            res["routed_rows_bit_exact"] &= bool(torch.equal(rc.grow_views[t][bags_here],
                                                             dx_full[bags_here, t * D:(t + 1) * D]))
        g = dx_full[:, t * D:(t + 1) * D]
        if wl.aggr != 21:
            g = g / float(L)
        bag_of = torch.arange(full.shape[0], device=dev).unsqueeze(1).expand_as(full)[mine]
        acc = torch.zeros((u.numel(), D), device=dev).index_add_(0, inv, g[bag_of])
        want = olds[t] - lr * acc
        got = rc.shards[t][u]
        err = ((got - want).abs() / want.abs().clamp_min(1e-3)).max().item() if u.numel() else 0.0
        res["worst_rel_err"] = max(res["worst_rel_err"], err)
        res["updated_rows_ok"] &= err <= 2e-6
        res["untouched_rows_bit_identical"] &= bool(torch.equal(rc.shards[t][rnd_rows[t]], rnd_old[t]))
    flags = [res[k] for k in res if k != "worst_rel_err"]
    okt = torch.tensor([1 if all(flags) else 0], device=dev)
    errt = torch.tensor([res["worst_rel_err"]], device=dev, dtype=torch.float64)
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    dist.all_reduce(errt, op=dist.ReduceOp.MAX)
    res["ok"] = bool(okt.item())
    res["worst_rel_err"] = errt.item()
    res["what"] = ("one un-timed step: forward vs the owners' shard rows (bit-exact), routed slots / output-grad rows "
                   "(bit-exact), touched rows vs old - lr*sum(g) (<= 2e-6 relative), untouched rows bit-identical; "
                   "flags of rank 0, `ok` = all ranks")
    return res


def run_rowwise(args, torch, pkg, dist, wl, rank, world, dev):
    """BASELINE config 4: tables too large for one GPU, sharded ROW-WISE over the ranks
    (exchange.RowWiseCollection).  Every rank pools its own samples (rows read from the owners' shards
    over NVLink inside the kernel); backward = route (bucketise by owner, remote stores of local row
    numbers + output-grad rows), flag, the owners' sorted backward over their own lookups, flag.
    --rowwise legacy: the round-1 form (two all-gathers, every owner walks the whole batch)."""
    lo, hi = pkg.sharding.batch_block(wl.batch, world, rank)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    T, D = wl.num_tables, wl.dim
    legacy = args.rowwise == "legacy"
    rc = pkg.exchange.RowWiseCollection(wl.rows, D, wl.bags, wl.batch, rank, world, dev, wl.aggr, legacy=legacy)
    for t in range(T):
        a = (1.0 / wl.rows[t]) ** 0.5
        rc.shards[t][:rc.local_rows[t]].uniform_(-a, a, generator=g)
    # this rank's samples: sample-major index batches [B_h][sum L] (the data loader's layout) and the
    # per-table [B_h][L] tensors cut out of them once (ffemb_load_sparse_inputs)
    idx_sm, idx = [], []
    for _ in range(args.pool):
        cols = [torch.randint(0, wl.rows[t], (hi - lo, wl.bags[t]), dtype=torch.int64, device=dev, generator=g)
                for t in range(T)]
        sm = torch.cat(cols, dim=1).contiguous()
        idx_sm.append(sm)
        idx.append(pkg.load_sparse_inputs(sm, wl.bags, hi - lo))
    dx = torch.empty((hi - lo, T * D), device=dev).uniform_(-1, 1, generator=g)
    out = [torch.empty((hi - lo, D), device=dev) for _ in range(T)]
    stream = torch.cuda.current_stream()
    torch.cuda.synchronize()
    dist.barrier()
    parity = None
    if not legacy:
        parity = rowwise_parity_step(torch, dist, wl, rc, idx[0], dx, out, rank, world, dev)

    def step(k):
        rc.forward(idx[k % args.pool], out)
        rc.backward_sgd(idx_sm[k % args.pool] if legacy else idx[k % args.pool], dx, wl.lr)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for k in range(args.warmup):
        step(k)
    barrier()
    use_graph = not legacy and args.graph != "off"
    if use_graph:   # the step is ~25 launches of a few microseconds: replay it from a CUDA graph per index batch
        graphs = []
        for p in range(args.pool):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, capture_error_mode="thread_local"):
                step(p)
            graphs.append(gr)
        barrier()
        eager_step = step

        def step(k):  # noqa: F811
            graphs[k % args.pool].replay()
        for k in range(args.warmup):
            step(k)
        barrier()
    sampler = ClockSampler(dev.index) if rank == 0 else None
    if sampler:
        sampler.start()
    pkg.profile_collect()
    l0 = pkg.launch_count()
    if use_graph:   # launches per step from one eager step
        eager_step(0)
        barrier()
        per_step = pkg.launch_count() - l0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    l0 = pkg.launch_count()
    e0.record(stream)
    for k in range(args.steps):
        step(args.warmup + k)
    e1.record(stream)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / args.steps
    launches = per_step * args.steps if use_graph else pkg.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    peak, peak_src = peaks()
    n, D = wl.lookups, wl.dim
    alg = 16 * n + 4 * D * n + 8 * D * wl.num_tables * wl.batch + 8 * D * n  # U ~ N (uniform over 100M rows)
    # NVLink bytes this rank sends per step: forward = the rows it READS from foreign shards arrive inbound
    # (counted at the reader); backward route = one gradient row per (bag, foreign owner) + 8 B per lookup and peer
    lookups_rank = (hi - lo) * sum(wl.bags)
    far = (world - 1) / world
    nvl = {"forward_rows_read_remote": int(lookups_rank * far * D * 4),
           "route_rows_sent": int(lookups_rank * far * D * 4) if not legacy else None,
           "route_slots_sent": int(lookups_rank * (world - 1) * 8) if not legacy else None,
           "allgather_received": int((wl.batch - (hi - lo)) * (T * D * 4 + sum(wl.bags) * 8)) if legacy else None}
    if rank == 0:
        line = {
            "metric": METRIC, "value": n / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.name, "tables": wl.num_tables, "dim": D, "batch": wl.batch, "bags": wl.bags,
                       "rows_total": int(sum(wl.rows)), "table_bytes_per_gpu": int(sum(sh.numel() * 4 for sh in rc.shards)),
                       "lookups_per_step": n, "parallelism": f"row-wise x{world}: remote gathers over NVLink in the "
                       "forward kernels; backward = " + (
                           "ONE all-gather (NCCL) of the sample-major indices and ONE of the interaction-input "
                           "gradients, then one local deterministic backward over the whole batch (round-1 form)"
                           if legacy else
                           "route kernel (bucketise by owner; local row numbers + output-grad rows stored into the "
                           "owners' buffers over NVLink), epoch flag, the owners' sorted backward + SGD over their own "
                           "lookups only, epoch flag; no collective"),
                       "cuda_graph": bool(use_graph),
                       "nvlink_bytes_per_rank_per_step": nvl,
                       "cache": "tables far larger than L2, random rows"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "step (launch/latency-bound at 64K lookups per table)",
                         "achieved": round(alg / (ms_step * 1e-3) / 1e9, 1), "peak": peak * world, "unit": "GB/s",
                         "frac": round(alg / (ms_step * 1e-3) / 1e9 / (peak * world), 4), "traffic": None,
                         "peak_source": peak_src}}
        if parity is not None:
            line["parity_ok"] = bool(parity["ok"])
            line["parity"] = parity
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()
    return 0


def reference_gpu_leg(torch, wl, dw, args):
    """BASELINE.md 3.2, "reference on B200": the reference's own GPU code for this path, compiled
    unmodified for sm_100a (oracle/_ref/libffref_gpu.so), timed on the SAME device-resident inputs as
    the product arm — per table and step: ZeroInitializer::init_task on the dense weight-grad
    (Op::zero_grad), Embedding::forward_kernel, Embedding::backward_kernel (atomicAdd scatter) and
    SGDOptimizer::ps_update_task (sgd_update over all R*D elements).  A reported baseline like
    cpu_baseline (same rule: never the thing shipped); only for workloads whose tables stay below the
    reference's int cliffs (R*D < 2^31: cfg1, cfg2) and fit a second dense buffer."""
    try:
        from oracle import oracle as O
        if not O.ref_gpu_available() or any(r * wl.dim >= (1 << 31) for r in wl.rows):
            return None
        if sum(wl.rows) * wl.dim * 4 > 40e9:
            return None
        L = O.ref_gpu()
        steps, warmup = max(3, min(args.steps, 20)), 3
        Ws = [w.clone() for w in dw.W]
        Wg = [torch.empty_like(w) for w in dw.W]
        stream = torch.cuda.current_stream()

        def step(k):
            p = k % dw.pool
            for t in range(len(Ws)):
                R = int(Ws[t].shape[0])
                L.ffref_gpu_zero_task(Wg[t].data_ptr(), R, wl.dim)
                L.ffref_gpu_forward_kernel(dw.idx[t][p].data_ptr(), dw.out[t].data_ptr(), Ws[t].data_ptr(),
                                           wl.bags[t], wl.dim, wl.batch, wl.aggr, None)
                L.ffref_gpu_backward_kernel(dw.idx[t][p].data_ptr(), dw.outgrad[t].data_ptr(), Wg[t].data_ptr(),
                                            wl.bags[t], wl.dim, wl.batch, wl.aggr, None)
                L.ffref_gpu_sgd_task(Wg[t].data_ptr(), Ws[t].data_ptr(), R, wl.dim, float(wl.lr), 0.0)

        # the reference's tasks run on the legacy default stream (built with DISABLE_LEGION_CUDA_HIJACK)
        torch.cuda.synchronize()
        with torch.cuda.stream(torch.cuda.default_stream()):
            for k in range(warmup):
                step(k)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(torch.cuda.default_stream())
            for k in range(steps):
                step(warmup + k)
            e1.record(torch.cuda.default_stream())
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        dense = 16 * sum(wl.rows) * wl.dim  # zero-fill 4 + sgd_update 12 bytes per element
        return {"value": wl.lookups / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": steps,
                "kind": "reference", "dense_bytes_per_step": int(dense),
                "what": "oracle/_ref/libffref_gpu.so: the reference's embedding.cu / initializer_kernel.cu / "
                        "optimizer_kernel.cu compiled unmodified for sm_100a; per table zero-fill + forward_kernel + "
                        "backward_kernel (atomicAdd) + ps_update_task (sgd_update over every row), same inputs, "
                        "device-resident (the reference keeps outputs in zero-copy host memory: not charged here)"}
    except Exception as e:
        return {"value": None, "unit": UNIT, "what": f"failed: {e!r}"}


def sharded_parity_step(torch, pkg, dist, wl, dw, fx, sstep, plan, rank, world, dev):
    """One UN-TIMED step before warm-up whose results are checked (the flag travels in the bench line, so
    the driver's scaling runs carry their own parity evidence):
      * interaction input: every column block of x on every rank BIT-EXACT against the pooled rows the
        part's owner computes with the plain (non-exchange) forward kernel — broadcast by the owner;
      * routed output-grads of one sampled part per rank BIT-EXACT against the ranks' dx blocks
        (all-gathered);
      * the sampled part's table after the step against a single-rank recomputation (a one-table call on
        a copy, the routed reference gradient): rows with at most one lookup bit-identical, every row
        within 1e-6 * (|W0| + lr * count) (|g| <= 1; the summation tree's tiles depend on which tables
        are batched together, the addends do not).
    torch.distributed (NCCL) moves the reference data: plumbing for the check only."""
    try:
        B, D, lr = wl.batch, wl.dim, wl.lr
        lo, hi = fx.lo, fx.hi
        idx0 = [b[0] for b in dw.idx]
        # plain forward of this rank's parts
        ref_out = {}
        for ids, coll in zip(dw.group_of, dw.colls):
            if coll is not None:
                outs = coll.forward([idx0[i] for i in ids])
                for i, o in zip(ids, outs):
                    ref_out[(dw.parts[i].table, dw.parts[i].col0)] = o
        # sampled part: most lookups among those whose copies fit comfortably
        free = torch.cuda.mem_get_info(dev)[0]
        cands = [i for i in range(len(dw.parts)) if 2.5 * dw.W[i].numel() * 4 < free]
        samp = max(cands, key=lambda i: (wl.bags[dw.parts[i].table] * min(wl.rows[dw.parts[i].table], 10**7), -i)) if cands else None
        W0 = dw.W[samp].clone() if samp is not None else None
        sstep.run([[idx0[i] for i in ids] for ids in dw.group_of], lr, overlap=True)
        torch.cuda.synchronize()
        dist.barrier()
        ok = torch.ones(3, dtype=torch.int32, device=dev)
        x = fx.x[:hi - lo]
        for p in sorted(plan, key=lambda p: (p.table, p.col0)):
            buf = ref_out[(p.table, p.col0)] if p.rank == rank else torch.empty((B, p.width), device=dev)
            dist.broadcast(buf, src=p.rank)
            c0 = p.table * D + p.col0
            if not torch.equal(x[:, c0:c0 + p.width], buf[lo:hi]):
                ok[0] = 0
        # routed output-grads + table update of each rank's sampled part
        for r in range(world):
            meta = torch.tensor([dw.parts[samp].table, dw.parts[samp].col0, dw.parts[samp].width] if (r == rank and samp is not None)
                                else [-1, 0, 0], dtype=torch.int64, device=dev)
            dist.broadcast(meta, src=r)
            t, c0, w = [int(v) for v in meta.tolist()]
            if t < 0:
                continue
            blk = torch.zeros((fx.ext, w), device=dev)
            blk[:hi - lo].copy_(fx.dx[:, t * D + c0:t * D + c0 + w])
            full = torch.empty((world * fx.ext, w), device=dev)
            dist.all_gather_into_tensor(full, blk)
            if r != rank:
                continue
            og = full[:B].contiguous()
            k = dw.parts[samp].group
            j = dw.group_of[k].index(samp)
            if not torch.equal(fx.outgrads_by_group[k][j], og):
                ok[1] = 0
            L = wl.bags[t]
            pkg.EmbeddingCollection([W0], [L], wl.aggr).backward_sgd([idx0[samp]], [og], lr)
            torch.cuda.synchronize()
            cnt = torch.bincount(idx0[samp].reshape(-1), minlength=W0.shape[0])
            single = cnt <= 1
            if not torch.equal(dw.W[samp][single], W0[single]):
                ok[2] = 0
            # W0 now holds the recomputed table; the bound uses its magnitude (the update is tiny against it)
            bound = 1e-6 * (W0.abs() + lr * cnt[:, None].float())
            if not bool(((dw.W[samp] - W0).abs() <= bound).all()):
                ok[2] = 0
            del W0, full, og
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        o = [int(v) for v in ok.tolist()]
        return {"ok": all(o), "interaction_input_bit_exact": bool(o[0]), "routed_outgrads_bit_exact": bool(o[1]),
                "sampled_tables_match_single_rank_recompute": bool(o[2]),
                "what": "one un-timed step before warm-up: x vs the owners' plain forward (all parts, all ranks), "
                        "routed output-grads and updated table of one sampled part per rank vs a single-rank recomputation"}
    except Exception as e:  # the check must never take the bench line with it
        return {"ok": False, "what": f"parity step failed: {e!r}"}


def load_traffic(phase):
    """dram bytes per launch of the dominant kernel from the committed ncu summary, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p)).get(phase)
    except Exception:
        return None


def run_e2e(torch, pkg, args, wl, dw, ex, fx, dist, dev, barrier, total_lookups):
    """The same step through the public host-buffer API (pkg.HostStep): per step H2D of that step's
    indices and output-grads from pinned host memory, D2H of the pooled outputs (N>1: of this rank's
    interaction input x; its dx is the uploaded gradient); a step ends when every copy has landed."""
    if ex is not None:
        return None  # the NCCL comparison form has no host-buffer API
    have = len(dw.owned) > 0
    hs = pkg.HostStep(dw.colls if fx is not None else dw.coll, wl.batch, exchange=fx)
    if fx is None:
        # two pinned host buffer sets laid out like HostStep's device slabs (one copy per kind and step);
        # consecutive steps are in flight together (HostStep.submit / wait)
        sets = [hs.host_buffers() for _ in range(2)]
        for k, (hi, hg, ho) in enumerate(sets):
            for t in range(len(hi)):
                hi[t].copy_(dw.idx[t][k % len(dw.idx[t])])
                hg[t].copy_(dw.outgrad[t])
        torch.cuda.synchronize()
        h_idx = [[sets[0][0][t], sets[1][0][t]] for t in range(len(sets[0][0]))]
        h_g = [sets[0][1], sets[1][1]]
        h_out = [sets[0][2], sets[1][2]]
        h2d = sum(x[0].numel() * 8 for x in h_idx) + sum(x.numel() * 4 for x in h_g[0])
        d2h = sum(x.numel() * 4 for x in h_out[0])
    else:
        h_idx = [[b.cpu().pin_memory() for b in bs[:2]] for bs in dw.idx] if have else []
        h_g = [fx.dx.cpu().pin_memory()] * 2
        h_out = [torch.empty(fx.x.shape, dtype=fx.x.dtype).pin_memory()] * 2
        h2d = sum(x[0].numel() * 8 for x in h_idx) + h_g[0].numel() * 4
        d2h = h_out[0].numel() * 4
    stream = torch.cuda.current_stream()

    def submit(k):
        """Every step uploads ITS indices and output-grads and downloads ITS pooled outputs; the
        host side only waits for step k-1 before it reuses that step's host buffers for k+1."""
        hs.wait(k - 2)
        return hs.submit([x[k % 2] for x in h_idx], h_g[k % 2], h_out[k % 2], wl.lr)

    steps = max(3, min(args.steps, 200 if wl.lookups <= 2_000_000 else 10))
    for k in range(4):   # warm-up; with hs.use_graph the second use of each buffer set captures its graphs
        submit(k)
    hs.wait()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for k in range(steps):
        submit(4 + k)
    hs.join(stream)   # the last download and the last row update are inside the timed region
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms, h2d, d2h], dtype=torch.float64, device=dev)
        dist.all_reduce(t[:1], op=dist.ReduceOp.MAX)
        dist.all_reduce(t[1:], op=dist.ReduceOp.SUM)
        ms, h2d, d2h = t[0].item(), t[1].item(), t[2].item()
    return {"value": total_lookups / (ms / steps * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
            "steps": steps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "cuda_graph": bool(hs.use_graph),
            "api": "HostStep.submit/wait: 3 streams (H2D / compute / D2H), pinned host buffers, device "
                   f"buffers x{hs.depth} so consecutive steps' copies run back to back; host buffers laid out like "
                   "the device slabs: one copy per kind and step",
            "h2d_gbs": round(h2d / (ms / steps * 1e-3) / 1e9, 1), "d2h_gbs": round(d2h / (ms / steps * 1e-3) / 1e9, 1)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU embedding tasks on the box's host cores
# ---------------------------------------------------------------------------------------------
class CpuSample:
    """Bounded sample of `wl` in host memory: every table (rows capped), batch B, two index batches."""

    def __init__(self, wl, B, cap):
        import numpy as np
        self.wl, self.B, self.cap = wl, B, cap
        rng = np.random.default_rng(1234)
        self.tabs = []
        for t in range(wl.num_tables):
            R = min(wl.rows[t], cap)
            W = np.empty((R, wl.dim), np.float32)
            rng.random(out=W, dtype=np.float32)
            W *= 2 * (1.0 / R) ** 0.5
            idx = [rng.integers(0, R, (B, wl.bags[t]), dtype=np.int64) for _ in range(2)]
            g = rng.random((B, wl.dim), dtype=np.float32) * 2 - 1
            self.tabs.append((R, W, idx, g, np.zeros((R, wl.dim), np.float32)))
        self.lookups = B * sum(wl.bags)


def cpu_arm(wl, sample, steps, warmup, P, full_step=False):
    """One timing of the CPU path over `sample`.
    Forward = the reference's AVX2 EmbeddingLookup (oracle/_ref, src/ops/embedding.cc:23-319)
    when built, else the oracle port; backward = the reference's backward_task_cpu for bag 1,
    the oracle port of embed_backward (src/ops/embedding.cu:192-217) for multi-hot bags, which
    the reference's CPU code cannot do (src/ops/embedding.cc:408).  One task per table, tables
    spread over P host threads (Legion -ll:cpu P; dlrm_strategy_hetero.cc:28-36).
    full_step: each table's task additionally zero-fills the dense weight-grad before the backward
    (Op::zero_grad, src/runtime/model.cc:466-490) and runs sgd_update over all R*D elements after it
    (src/runtime/optimizer_kernel.cu:23-41) — RESTATED on the CPU (oracle port): the reference has no
    CPU optimizer (SGD_UPD_* are GPU-only, src/runtime/model.cc:2990-3022)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as O
    O.lib()
    have_ref = O.ref_cpu_available(ndebug=True)
    kind = "reference" if have_ref and all(l == 1 for l in wl.bags) else "port"

    def run_table(t, p):
        R, W, idx, g, Wg = sample.tabs[t]
        L = wl.bags[t]
        if have_ref:
            O.ref_cpu_lookup(idx[p], W, normalize=(wl.aggr != 21), ndebug=True)
        else:
            O.forward(idx[p], W, 122 if wl.aggr != 21 else 21, check_range=False)
        if full_step:
            Wg.fill(0.0)
        if have_ref and L == 1:
            O.ref_cpu_backward_task(idx[p], g, Wg, ndebug=True)
        else:
            O.backward_dense(idx[p], g, R, wl.aggr, Wgrad=Wg)
        if full_step:
            O.lib().oracle_sgd_update(W.size, wl.lr, 0.0, 0.0, 0, Wg.reshape(-1), None, W.reshape(-1), 1)

    if kind == "reference" and not full_step:
        # the reference's own task bodies (forward_task_cpu + backward_task_cpu) driven from C++ worker
        # threads, table t on thread t % P (oracle/ref_glue_cpu.cc): no Python between the tasks.  Built
        # with the reference's flags; asserts stay on (its Release flags keep them) unless a table has more
        # rows than the data_size = 1,000,000 the CPU kernel asserts against (src/ops/embedding.cc:71-73,404)
        import numpy as np
        nd = max(tb[0] for tb in sample.tabs) > 1_000_000
        T = wl.num_tables
        idxb = [np.ascontiguousarray(np.stack(tb[2])) for tb in sample.tabs]
        outs = [np.empty((sample.B, wl.dim), np.float32) for _ in range(T)]
        Ws, gs, Wgs = [tb[1] for tb in sample.tabs], [tb[3] for tb in sample.tabs], [tb[4] for tb in sample.tabs]
        if warmup:
            O.ref_cpu_run_tables(idxb, outs, Ws, gs, Wgs, P, warmup, ndebug=nd)
        dt = O.ref_cpu_run_tables(idxb, outs, Ws, gs, Wgs, P, steps, ndebug=nd)
        return {"value": sample.lookups / (dt / steps), "ms_per_step": dt / steps * 1e3, "kind": kind,
                "have_ref": True, "driver": "C++ worker threads (ffref_cpu_run_tables)"}
    order = sorted(range(wl.num_tables), key=lambda t: -wl.bags[t])  # longest tasks first
    with ThreadPoolExecutor(P) as pool:
        def one_step(k):
            list(pool.map(lambda t: run_table(t, k % 2), order))
        for k in range(warmup):
            one_step(k)
        t0 = time.perf_counter()
        for k in range(steps):
            one_step(warmup + k)
        dt = time.perf_counter() - t0
    return {"value": sample.lookups / (dt / steps), "ms_per_step": dt / steps * 1e3, "kind": kind,
            "have_ref": have_ref}


def cpu_report(args, wl, steps, warmup):
    """cpu_baseline object: task bodies at P = all usable host cores (the headline
    `value`), the same at P = 1 and P = 4 (the reference scripts' -ll:cpu 4, run_random.sh:9), and the
    restated full step (dense zero-fill + sgd_update) at P = all."""
    cores = os.cpu_count() or 1
    P = min(cores, wl.num_tables)
    B = min(args.cpu_batch, wl.batch)
    sample = CpuSample(wl, B, args.cpu_rows_cap)
    main = cpu_arm(wl, sample, steps, warmup, P)
    side = max(1, min(steps, 2))
    by = {str(P): round(main["value"], 1)}
    for p in (1, 4):
        if p < P:
            # bounded: one pass over a quarter of the sample's batch would change the cache behaviour;
            # instead fewer steps of the same sample
            by[str(p)] = round(cpu_arm(wl, sample, 1, 0, p)["value"], 1)
    full = cpu_arm(wl, sample, side, 1 if warmup else 0, P, full_step=True)
    have_ref, kind = main["have_ref"], main["kind"]
    return {"value": main["value"], "unit": UNIT, "cores": P, "kind": kind,
            "ms_per_step": main["ms_per_step"], "host_cores": cores, "steps": steps, "warmup": warmup,
            "by_threads": by,
            "full_step": {"value": round(full["value"], 1), "ms_per_step": round(full["ms_per_step"], 3),
                          "cores": P, "what": "task bodies + RESTATED dense zero-fill and sgd_update over all rows "
                          "(oracle port; the reference has no CPU optimizer)"},
            "sample": f"{wl.name}: all {wl.num_tables} tables, batch {B} of {wl.batch}, rows capped at "
                      f"{args.cpu_rows_cap}; fwd={'reference AVX2 EmbeddingLookup (oracle/_ref)' if have_ref else 'oracle port'}, "
                      f"bwd={'reference backward_task_cpu' if kind == 'reference' else 'oracle port of embed_backward (multi-hot)'}"
                      f"; `value` = task bodies only (no dense zero-fill / sgd_update), see full_step"}


def cpu_baseline(args, wl, steps, warmup):
    try:
        return cpu_report(args, wl, steps, warmup)
    except Exception as e:  # the baseline is a reported figure, never a reason to lose the line
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e!r}"}


def reference_arm(args, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores, K timed steps
    after W warm-up steps as asked; each step is a bounded sample of the workload (stated in `sample`)."""
    pkg = importlib.import_module("dlrm-flexflow_b200")
    wl = pkg.synth.config(args.workload, args.scale)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    r = cpu_report(args, wl, steps, warmup)
    return {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.name, "tables": wl.num_tables, "dim": wl.dim, "batch": wl.batch,
                       "bags": wl.bags, "aggr": wl.aggr, "index_dist": wl.dist},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "by_threads",
                                               "full_step", "host_cores")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


if __name__ == "__main__":
    sys.exit(main())
