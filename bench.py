#!/usr/bin/env python
"""Headline benchmark: contraction throughput on BASELINE.json configs[1]
(bristlecone_70 depth 1-32-1 with the shipped ordering).  The job is a FIXED SUBSET of
its 65536 cut-slices -- 64 slices selected through cut value lists exactly as SURVEY 8d
prescribes, the same 64 whatever the GPU count (STRONG scaling: the slices are dealt
over the ranks, one all-reduce of the amplitude table per step).

  python bench.py --gpus N --steps K --warmup W            (our arm)
  python bench.py --impl reference --gpus N --steps K --warmup W   (CPU reference arm)

A step = one pass of the schedule over the job's slices.  `value` is measured with the
network resident in HBM (gate tensors uploaded, grid tensors flattened) and counts the
job's algorithmic flops ONCE (work replicated on several ranks is not counted twice);
`e2e` is the same metric through qflex_b200_evaluate (the EvaluateCircuit C-ABI call)
with HOST text buffers: parse + plan + host->device upload + flatten + contract +
device->host amplitudes inside the timed region.

The reference arm and the cpu_baseline leg run the UNMODIFIED reference (oracle/_ref,
built by oracle/Makefile) on the box's host cores; they load nothing of the product.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
CONFIG = os.path.join(ROOT, "config")

WORKLOADS = {
    # name -> files, cut value lists ("all": the same job at every world size = strong
    # scaling; per-world lists = weak scaling), the bounded sample the reference is timed on
    # BASELINE configs[1], the headline: 64 of the 65536 slices
    "bristlecone_70_1-32-1": dict(
        circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt", scaling="strong",
        cuts={"all": [[0, 1, 2, 3], [0, 1, 2, 3], [0, 1], [0, 1]]},
        reference_sample=[[0], [0], [0], [0]], golden="bristlecone_70_1-32-1_1slice",
        sweep_proxy="bristlecone_70_1-24-1", memory=64 << 30, total_slices=65536),
    # the round-1 form of the headline: 8 slices per GPU (weak scaling)
    "bristlecone_70_1-32-1_weak": dict(
        circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt",
        cuts={1: [[0, 1], [0, 1], [0, 1], [0]], 2: [[0, 1], [0, 1], [0, 1], [0, 1]],
              4: [[0, 1], [0, 1], [0, 1], [0, 1, 2, 3]],
              8: [[0, 1], [0, 1], [0, 1, 2, 3], [0, 1, 2, 3]]},
        reference_sample=[[0], [0], [0], [0]], golden="bristlecone_70_1-32-1_1slice",
        sweep_proxy="bristlecone_70_1-24-1", memory=64 << 30, total_slices=65536),
    # BASELINE configs[0]: the whole circuit the reference's CI runs (2 slices x 4 amplitudes)
    "bristlecone_48_1-24-1": dict(
        circuit="bristlecone_48_1-24-1_0.txt", ordering="bristlecone_48.txt",
        grid="bristlecone_48.txt", scaling="strong", cuts={"all": None},
        reference_sample="all", golden="bristlecone_48_1-24-1", memory=8 << 30, total_slices=2),
    "bristlecone_70_1-24-1": dict(
        circuit="bristlecone_70_1-24-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt", scaling="strong",
        cuts={"all": [[0, 1, 2, 3], [0, 1, 2, 3], [0, 1], [0, 1]]},
        reference_sample=[[0, 1], [0, 1], [0, 1], [0]], memory=8 << 30, total_slices=4096),
    # BASELINE configs[4] stand-in: Sycamore-53 m=10, a low-fidelity batch of 64 of its
    # 16384 slices, strong scaling
    "sycamore_53_10": dict(
        circuit="sycamore_53_10_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1, 2, 3], [0, 1, 2, 3], [0, 1, 2, 3]]},
        reference_sample=[[0], [0], [0]], memory=100 << 30, total_slices=16384),
    # Sycamore-53 m=12 (29.6 GB arena here, 7.1e14 flop per slice; the reference cannot
    # hold its patches in this host's memory), 8 of its slices, strong scaling
    "sycamore_53_12": dict(
        circuit="sycamore_53_12_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1], [0, 1], [0, 1]]},
        reference_sample=None, memory=100 << 30, total_slices=None),
    # BASELINE configs[2] as named: Sycamore-53 m=14 with config/ordering/sycamore_53_m14.txt
    # (shipped ordering + six bond cuts from tools/augment_ordering.py; 2^58 slices, 81 GB
    # arena).  8 slices x 64 output bitstrings, strong scaling.
    "sycamore_53_14": dict(
        circuit="sycamore_53_14_0.txt", ordering="sycamore_53_m14.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1]] * 3 + [[0]] * 6 + [[0, 1]] * 6},
        reference_sample=None, memory=160 << 30, total_slices=2.0 ** 58),
    # Sycamore-53 m=16: shipped ordering + 22 bond cuts, cut-aware flatten; 64 slices x
    # 64 bitstrings, strong scaling
    "sycamore_53_16": dict(
        circuit="sycamore_53_16_0.txt", ordering="sycamore_53_m16.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1, 2, 3]] * 3 + [[0]] * 22 + [[0, 1]] * 6},
        reference_sample=None, memory=64 << 30, total_slices=2.0 ** 200),
    # BASELINE configs[4] as named: Sycamore-53 m=20 (+36 bond cuts, cut-aware flatten,
    # 120 GB arena).  8 of the slices x 64 output bitstrings, strong scaling.
    "sycamore_53_20": dict(
        circuit="sycamore_53_20_0.txt", ordering="sycamore_53_m20.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1]] * 3 + [[0]] * 36 + [[0, 1]] * 6},
        reference_sample=None, memory=170 << 30, total_slices=2.0 ** 390),
    # BASELINE configs[2] at m=8: 64 slices x a batch of 64 output bitstrings, strong scaling
    "sycamore_53_8_batch64": dict(
        circuit="sycamore_53_8_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        scaling="strong",
        cuts={"all": [list(range(16)), [0, 1, 2, 3], [0]] + [[0, 1]] * 6},
        reference_sample=[[0], [0], [0]] + [[0, 1]] * 6, memory=8 << 30, total_slices=4096),
    # BASELINE configs[3]: 7x7 1-32-1 with the reference's ExampleOrdering schedule,
    # 64 of its 256 slices x 64 output bitstrings, strong scaling
    "rectangular_7x7_1-32-1": dict(
        circuit="rectangular_7x7_1-32-1_0.txt", ordering="rectangular_7x7_v2.txt",
        grid="rectangular_7x7.txt", scaling="strong",
        cuts={"all": [list(range(16)), [0, 1, 2, 3]] + [[0, 1]] * 6},
        reference_sample=[[0], [0]] + [[0, 1]] * 6, memory=16 << 30, total_slices=256),
    # BASELINE configs[3] as named: 7x7 depth 1-40-1 (two bond cuts of 32 values = 1024
    # slices; 64 GB arena).  64 output bitstrings per slice; weak scaling.
    "rectangular_7x7_1-40-1": dict(
        circuit="rectangular_7x7_1-40-1_0.txt", ordering="rectangular_7x7_v2.txt",
        grid="rectangular_7x7.txt",
        cuts={1: [[0], [0]] + [[0, 1]] * 6, 2: [[0, 1], [0]] + [[0, 1]] * 6,
              4: [[0, 1], [0, 1]] + [[0, 1]] * 6, 8: [[0, 1, 2, 3], [0, 1]] + [[0, 1]] * 6},
        reference_sample=None, memory=160 << 30, total_slices=1024),
}
FLOPS_CACHE = os.path.join(ROOT, "bench_reference_flops.json")


def workload_cuts(w, world):
    return w["cuts"]["all"] if "all" in w["cuts"] else w["cuts"][world]


def restrict_cuts(ordering_text, values_per_cut):
    """Value list of the k-th `cut` line := values_per_cut[k] (same rewrite as
    qflex_b200.restrict_cuts; repeated here so that the reference arm imports nothing of
    the product)."""
    import re
    out, k = [], 0
    for line in ordering_text.splitlines():
        if line.strip().startswith("cut"):
            if k < len(values_per_cut) and values_per_cut[k] is not None:
                vals = ",".join(str(v) for v in values_per_cut[k])
                line = re.sub(r"\([0-9,\s]*\)", "(" + vals + ")", line, count=1)
            k += 1
        out.append(line)
    return "\n".join(out) + "\n"


def load_texts(w, cut_values):
    def read(kind, name):
        with open(os.path.join(CONFIG, kind, name)) as f:
            return f.read()
    o = read("ordering", w["ordering"])
    if cut_values not in (None, "all"):
        o = restrict_cuts(o, cut_values)
    return read("circuits", w["circuit"]), o, read("grid", w["grid"])


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "sm_max_mhz": 1965.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------- reference (CPU) legs
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_run(wname, cut_values, omp, blas, wait="passive", timeout=3000):
    """Times the UNMODIFIED reference (oracle/_ref) on `cut_values` of workload `wname`
    in a child process (thread settings must be in the environment before OpenBLAS and
    libgomp load; the child imports oracle/ref_run.py only).  Returns
    (seconds of its EvaluateCircuit, n_slices, [[re, im]...])."""
    code = (
        "import sys, json; sys.path.insert(0, %r)\n"
        "import bench\n"
        "from oracle import ref_run\n"
        "w = bench.WORKLOADS[%r]\n"
        "c, o, g = bench.load_texts(w, %r)\n"
        "ref_run.reference_set_threads(%d)\n"
        "res, secs = ref_run.reference_evaluate(c, o, g, memory=%d)\n"
        "print(json.dumps({'secs': secs, 'amps': [[a.real, a.imag] for _, a in res]}))\n"
    ) % (ROOT, wname, cut_values, blas, WORKLOADS[wname]["memory"])
    env = dict(os.environ, OMP_NUM_THREADS=str(omp), OPENBLAS_NUM_THREADS=str(blas),
               OMP_WAIT_POLICY=wait)
    p = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError("reference run failed: " + p.stderr[-2000:])
    out = json.loads(p.stdout.strip().splitlines()[-1])
    n = 1
    if cut_values not in (None, "all"):
        for v in cut_values:
            n *= len(v)
    else:
        n = WORKLOADS[wname].get("total_slices") or 1
    return out["secs"], n, out["amps"]


def sweep_reference_threads(wname, budget_s=75.0):
    """(OMP_NUM_THREADS, OPENBLAS_NUM_THREADS) in {1, n/2, n}^2 x
    OMP_WAIT_POLICY in {active, passive}; the reference collapses under oversubscription
    (229 s vs 0.97 s in the survey's probe), so the sweep runs on a cheap proxy of the
    workload (its small sibling circuit, or itself when cheap), every run under a
    timeout, the whole sweep under a time budget.  Returns (best setting, table)."""
    w = WORKLOADS[wname]
    proxy = w.get("sweep_proxy", wname)
    sample = WORKLOADS[proxy]["reference_sample"]
    n = host_cores()
    counts = sorted({1, max(1, n // 2), n})
    # most promising first, so that a short budget still sees them
    order = sorted(((o, b, wt) for o in counts for b in counts for wt in ("passive", "active")),
                   key=lambda s: (-min(s[0], n // 2 or 1) - s[1], s[2] != "passive"))
    table, best, t0, limit = [], None, time.time(), 20.0
    for omp, blas, wait in order:
        if time.time() - t0 > budget_s:
            break
        try:
            secs, _, _ = cpu_reference_run(proxy, sample, omp, blas, wait, timeout=limit)
        except subprocess.TimeoutExpired:
            table.append(dict(omp=omp, blas=blas, wait=wait, seconds=None, note="> %.0f s" % limit))
            continue
        except Exception as e:
            table.append(dict(omp=omp, blas=blas, wait=wait, seconds=None, note=str(e)[:80]))
            continue
        table.append(dict(omp=omp, blas=blas, wait=wait, seconds=secs))
        if best is None or secs < best[3]:
            best = (omp, blas, wait, secs)
            limit = max(3.0, min(limit, 4.0 * secs + 2.0))  # python start-up included
    if best is None:
        best = (min(n, 16), min(n, 32), "passive", None)
    return best[:3], dict(proxy=proxy, proxy_sample=sample, runs=table, host_cores=n)


def pick_reference_threads(wname, budget_s=60.0):
    """Thread setting for the reference on THIS workload's sample: the proxy sweep ranks
    the settings, then the two best of it and the two all-cores settings (a large slice
    favours more BLAS threads than the small proxy does) are timed on the sample itself.
    Returns (setting, seconds on the sample, n_slices, amplitudes, sweep record)."""
    w = WORKLOADS[wname]
    _, sweep = sweep_reference_threads(wname, budget_s)
    n = sweep["host_cores"]
    ranked = sorted((r for r in sweep["runs"] if r["seconds"] is not None), key=lambda r: r["seconds"])
    cands = [(r["omp"], r["blas"], r["wait"]) for r in ranked[:2]]
    for extra in ((n, n, "passive"), (max(1, n // 2), n, "passive")):
        if extra not in cands:
            cands.append(extra)
    best, finals = None, []
    for omp, blas, wait in cands:
        limit = 3000 if best is None else max(30.0, 2.0 * best[1])
        try:
            secs, n_slices, amps = cpu_reference_run(wname, w["reference_sample"], omp, blas, wait,
                                                     timeout=limit)
        except Exception as e:
            finals.append(dict(omp=omp, blas=blas, wait=wait, seconds=None, note=str(e)[:60]))
            continue
        finals.append(dict(omp=omp, blas=blas, wait=wait, seconds=secs))
        if best is None or secs < best[1]:
            best = ((omp, blas, wait), secs, n_slices, amps)
    if best is None:
        raise RuntimeError("the reference did not run with any thread setting")
    sweep["on_the_sample"] = finals
    return best + (sweep,)


def reference_flops(wname, sample):
    """Algorithmic flops (sum 8*M*N*K) of the REFERENCE's schedule on the
    slices `sample` selects -- projected leaves, slice by slice.  Read from
    bench_reference_flops.json (written by tools/make_reference_flops.py with the
    planner on the CPU-emulated ABI) so that this arm loads nothing of the product."""
    with open(FLOPS_CACHE) as f:
        table = json.load(f)
    return table[wname + "|" + json.dumps(sample)]


def reference_line(args, wname, w, seconds, n_slices, threads, cores, extra=None):
    ms = 1e3 * sum(seconds) / len(seconds)
    val = reference_flops(wname, w["reference_sample"]) / (ms * 1e-3) / 1e12
    omp, blas, wait = threads
    sample = ("%s slice(s) of %s per step (cut values %s), reference EvaluateCircuit, "
              "OMP=%d OPENBLAS=%d %s wait, %d host cores"
              % (n_slices, wname, w["reference_sample"], omp, blas, wait, cores))
    line = {
        "impl": "reference", "metric": "contraction_tflops", "value": val, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": w.get("scaling", "weak"), "vs_baseline": None,
        "dtype": "complex64",
        "data": "shipped circuit (config/circuits/%s), deterministic" % w["circuit"],
        "config": {"workload": wname, "slices_per_step": n_slices,
                   "note": "unmodified reference qflex (oracle/_ref) on host cores, OpenBLAS + OpenMP"},
        "slices_per_s": n_slices / (ms * 1e-3),
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": max(omp, blas),
                         "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if extra:
        line["cpu_baseline"].update(extra)
    return line


def run_reference_arm(args, wname, w):
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    if w["reference_sample"] is None:
        print(json.dumps({"impl": "reference", "unavailable":
                          "%s: the reference needs more host memory and time than a bounded "
                          "sample allows (see the workload's comment in bench.py)" % wname}), flush=True)
        return 0
    threads, _, n_slices, _, sweep = pick_reference_threads(wname)  # (also warms the page cache)
    times = []
    for i in range(args.warmup + args.steps):
        secs, n_slices, _ = cpu_reference_run(wname, w["reference_sample"], *threads)
        if i >= args.warmup:
            times.append(secs)
    print(json.dumps(reference_line(args, wname, w, times, n_slices, threads, sweep["host_cores"],
                                    {"thread_sweep": sweep})), flush=True)
    return 0


# ---------------------------------------------------------------- our arm
def gate_upload_bytes(c):
    """Bytes of gate tensors copied host->device per EvaluateCircuit call."""
    import re
    n1 = len(re.findall(r"^\s*\d+\s+(h|t|x_1_2|y_1_2|hz_1_2|rz\S*)\s", c, flags=re.M))
    n2 = len(re.findall(r"^\s*\d+\s+(cz|cx)\s", c, flags=re.M))
    nf = len(re.findall(r"^\s*\d+\s+fsim", c, flags=re.M))
    nq = int(c.split()[0])
    return 8 * (4 * n1 + 2 * 8 * n2 + 2 * 16 * nf + 2 * 2 * nq)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="bristlecone_70_1-32-1", choices=list(WORKLOADS))
    ap.add_argument("--precision", default="auto", choices=["auto", "ffma", "bf16x3"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference_arm(args, args.workload, w)
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps

    import numpy as np
    import torch
    import torch.distributed as dist
    import qflex_b200
    from qflex_b200 import sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torchrun)" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    reducer = sharded.AmplitudeReducer(dev if world > 1 else None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cuts = workload_cuts(w, world)
    c, o, g = load_texts(w, cuts)
    lib = qflex_b200.default_library()
    prep = lib.prepare(c, o, g, memory_limit=w["memory"], device=local_rank,
                       precision=args.precision, world=world)

    # ---- resident: W warm-up + K timed steps (no per-launch profiling in here)
    for _ in range(args.warmup):
        amps = reducer.allreduce(prep.contract(rank, world))
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    dev_ms, flops_done, launches, leaves, tc_launches = 0.0, 0.0, 0, 0, 0
    for _ in range(args.steps):
        part = prep.contract(rank, world)
        amps = reducer.allreduce(part)
        st = prep.stats
        dev_ms += st["device_ms_contract"]
        flops_done += st["flops"]
        launches += int(st["launches"])
        leaves += int(st["leaves"])
        tc_launches += int(st["launches_gemm_tc"])
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    wall = max_over_ranks(wall)
    dev_ms_max = max_over_ranks(dev_ms)
    executed_flops = sum_over_ranks(flops_done)   # replicated work counted per rank
    total_leaves = sum_over_ranks(leaves)
    total_launches = int(sum_over_ranks(launches))
    job_flops = prep.stats["flops_job"]            # the job once, whatever the rank count
    ms_per_step = 1e3 * wall / args.steps
    value = job_flops * args.steps / wall / 1e12
    arena_bytes = prep.stats["arena_bytes"]
    plan_info = {k: int(prep.stats[k]) for k in ("fused_bits", "open_outputs", "lanes")}
    n_amps = len(prep.states)

    # ---- one more pass with per-launch CUDA events: shape classes for the roofline
    prof = {}
    prep.profile(True)
    for _ in range(max(1, min(args.steps, 2))):
        prep.contract(rank, world)
        for r in prep.profile_rows():
            key = (int(r["kind"]), int(r["m_bits"]), int(r["n_bits"]), int(r["k_bits"]))
            acc = prof.setdefault(key, dict(launches=0, ms=0.0, flops=0.0, bytes=0.0))
            for f in ("launches", "ms", "flops", "bytes"):
                acc[f] += r[f]
    prep.profile(False)
    prep.close()

    # ---- e2e: EvaluateCircuit through the C ABI from host text buffers
    e2e_steps = 5
    lib.evaluate(c, o, g, memory_limit=w["memory"], device=local_rank, rank=rank, world=world,
                 precision=args.precision)  # warm-up (allocator, module load)
    barrier()
    e2e_sampler = ClockSampler(local_rank)
    if rank == 0:
        e2e_sampler.start()
    e2e_calls_ms = []
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        t_call = time.perf_counter()
        res, st = lib.evaluate(c, o, g, memory_limit=w["memory"], device=local_rank, rank=rank,
                               world=world, precision=args.precision)
        full = reducer.allreduce(np.array([a for _, a in res]))
        e2e_calls_ms.append(round(1e3 * (time.perf_counter() - t_call), 1))
    barrier()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e_clocks = e2e_sampler.stop() if rank == 0 else None
    e2e_value = st["flops_job"] * e2e_steps / e2e_wall / 1e12
    # the same call with the plan memory off (every call parses AND plans from scratch;
    # above, calls after the first recall the plan of the identical previous call)
    os.environ["QFLEX_PLAN_MEMORY"] = "0"
    barrier()
    t0 = time.perf_counter()
    for _ in range(2):
        res2, _ = lib.evaluate(c, o, g, memory_limit=w["memory"], device=local_rank, rank=rank,
                               world=world, precision=args.precision)
        reducer.allreduce(np.array([a for _, a in res2]))
    barrier()
    e2e_cold_ms = 1e3 * max_over_ranks(time.perf_counter() - t0) / 2
    del os.environ["QFLEX_PLAN_MEMORY"]
    h2d = gate_upload_bytes(c) + len(c) + len(o) + len(g)
    parity = float(np.linalg.norm(full - amps) / max(np.linalg.norm(amps), 1e-300))

    # ---- parity of THIS workload against the reference: the bounded sample the CPU arm
    # runs (one slice), through the same C-ABI call
    parity_ref, sample_amps = {}, None
    if rank == 0 and w.get("reference_sample") is not None:
        sc, so, sg = load_texts(w, w["reference_sample"])
        sres, _ = lib.evaluate(sc, so, sg, memory_limit=w["memory"], device=local_rank,
                               precision=args.precision)
        sample_amps = np.array([a for _, a in sres])
        gpath = os.path.join(ROOT, "tests", "golden", "amplitudes.json")
        if w.get("golden") and os.path.exists(gpath):
            with open(gpath) as f:
                gold = json.load(f)[w["golden"]]
            want = np.array(gold["re"]) + 1j * np.array(gold["im"])
            parity_ref["parity_vs_reference_golden"] = float(
                np.linalg.norm(sample_amps - want) / np.linalg.norm(want))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel class (by summed device time)
    peaks, peak_src = measured_peaks()
    dom_key = max(prof, key=lambda k: prof[k]["ms"])
    dom = prof[dom_key]
    per_launch_ms = dom["ms"] / dom["launches"]
    ai = dom["flops"] / dom["bytes"]
    sm_max = peaks.get("sm_max_mhz", 1965.0)
    ffma_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12  # TFLOP/s fp32 FFMA
    ridge = ffma_peak * 1e12 / (peaks["hbm_gbs"] * 1e9)
    tc_class = (args.precision != "ffma" and min(dom_key[1], dom_key[2]) >= 5 and
                max(dom_key[1], dom_key[2]) >= 7 and dom_key[3] >= 5 and tc_launches > 0)
    if ai < ridge:
        roof = {"bound": "hbm", "achieved": dom["bytes"] / dom["ms"] / 1e6,
                "peak": peaks["hbm_gbs"], "unit": "GB/s"}
    elif tc_class:
        # tcgen05 split-bf16 path: every algorithmic flop is executed as 3 bf16
        # tensor-core flops (a_hi*b_hi + a_hi*b_mid + a_mid*b_hi); `achieved` counts
        # the executed bf16 flops, `peak` is the measured sustained cuBLAS bf16 rate.
        # The class time includes its gather+split pre-pass kernels.
        roof = {"bound": "tensor", "achieved": 3.0 * dom["flops"] / dom["ms"] / 1e9,
                "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "algorithmic_tflops": dom["flops"] / dom["ms"] / 1e9}
    else:
        roof = {"bound": "ffma", "achieved": dom["flops"] / dom["ms"] / 1e9, "peak": ffma_peak,
                "unit": "TFLOP/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["traffic"] = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel class,
        # NOT measured in this run: read from the committed `ncu --set full` capture that
        # profiles/ncu_traffic.json names
        with open(tpath) as f:
            t = json.load(f).get("%d,%d,%d" % dom_key[1:])
        if t:
            roof["traffic"] = t["dram_bytes_per_launch"]
            roof["traffic_source"] = "from committed ncu capture: " + t["source"]
            roof["algorithmic_bytes_per_launch"] = dom["bytes"] / dom["launches"]
    roof["kernel"] = "contract m=%d n=%d k=%d bits" % dom_key[1:]
    roof["launches"] = int(dom["launches"])
    roof["ms_per_launch"] = per_launch_ms
    roof["share_of_device_time"] = dom["ms"] / sum(v["ms"] for v in prof.values())
    roof["arithmetic_intensity"] = ai
    roof["timing"] = "CUDA events per launch on the context stream, in a separate profiled pass"
    roof["peak_source"] = ("%s: hbm_gbs %.1f, bf16 sustained %.1f (MEASURED_PEAKS.json); FFMA peak "
                           "derived = 148 SM x 128 lanes x 2 x %.0f MHz"
                           % (peak_src, peaks["hbm_gbs"], peaks["bf16_tflops_sustained"], sm_max))

    # whole-step roofline: sum of per-class t_min over the measured class times
    def t_min_ms(key, v):
        tc = (args.precision != "ffma" and min(key[1], key[2]) >= 4 and key[3] >= 4 and
              max(key[1], key[2]) >= 7)
        pipe = peaks["bf16_tflops_sustained"] / 3.0 if tc else ffma_peak
        return max(v["flops"] / (pipe * 1e9), v["bytes"] / (peaks["hbm_gbs"] * 1e6))
    step_frac = (sum(t_min_ms(k, v) for k, v in prof.items()) /
                 max(1e-9, sum(v["ms"] for v in prof.values())))

    slices_per_step = total_leaves / args.steps / max(1, n_amps)
    line = {
        "metric": "contraction_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": w.get("scaling", "weak"), "vs_baseline": None,
        "dtype": "complex64",
        "data": "shipped circuit (config/circuits/%s), deterministic; synthetic = none needed" % w["circuit"],
        "config": {"workload": args.workload, "slices_per_step": int(round(slices_per_step)),
                   "slices_per_gpu": slices_per_step / world, "amplitudes_per_slice": n_amps,
                   "cut_values": cuts, "arena_gb": arena_bytes / 2 ** 30,
                   "precision": args.precision, "plan": plan_info,
                   "l2": "working set per contraction (>= 2 GB patches) exceeds the 126 MB L2",
                   "collective": "one NCCL all-reduce(sum) of %d complex128 per step" % n_amps},
        # `value` counts the job's flops once; executed = what the ranks ran in all
        # (steps above the dealt cut level are replicated on every rank)
        "executed_tflops": executed_flops / wall / 1e12,
        "replicated_share": max(0.0, executed_flops / (job_flops * args.steps) - 1.0),
        "slices_per_s": total_leaves / max(1, n_amps) / wall,
        "leaves_per_s": total_leaves / wall,
        # amplitudes/s of the FULL circuit at this rate: every output bitstring needs all
        # of the circuit's bond-cut slices
        "amplitudes_per_s_full_circuit": (total_leaves / wall / float(w["total_slices"])
                                          if w.get("total_slices") else None),
        "device_ms_per_step": dev_ms_max / args.steps,
        "e2e": dict({"value": e2e_value, "unit": "TFLOP/s", "h2d_bytes_per_step": h2d,
                     "d2h_bytes_per_step": 16 * n_amps, "steps": e2e_steps,
                     "ms_per_step": 1e3 * e2e_wall / e2e_steps,
                     "ms_per_call_rank0": e2e_calls_ms, "clocks": e2e_clocks,
                     "plan_memory": "on: calls after the first recall the plan of the identical "
                                    "previous call (parse, upload, flatten, contract still run)",
                     "ms_per_step_planning_every_call": e2e_cold_ms,
                     "parity_vs_resident_run": parity}, **parity_ref),
        "gpu_launches": total_launches,
        "clocks": clocks, "roofline": roof, "step_roofline_frac": step_frac,
        "kernel_classes": [dict(kind=k[0], m_bits=k[1], n_bits=k[2], k_bits=k[3],
                                launches=int(v["launches"]), ms=v["ms"],
                                tflops=v["flops"] / v["ms"] / 1e9 if v["ms"] else None,
                                gbs=v["bytes"] / v["ms"] / 1e6 if v["ms"] else None,
                                frac_of_roof=t_min_ms(k, v) / v["ms"] if v["ms"] else None)
                           for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:10]],
    }
    if world == 1 and not args.no_cpu_baseline:
        if w["reference_sample"] is None:
            line["cpu_baseline"] = {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                                    "sample": "not run: no bounded sample of this workload fits the host"}
        else:
            try:
                threads, secs, n_slices, ref_amps, sweep = pick_reference_threads(args.workload, 45.0)
                ref_line = reference_line(args, args.workload, w, [secs], n_slices, threads,
                                          sweep["host_cores"], {"seconds": secs, "thread_sweep": sweep})
                line["cpu_baseline"] = ref_line["cpu_baseline"]
                want = np.array([complex(a, b) for a, b in ref_amps])
                line["e2e"]["parity_vs_reference"] = float(
                    np.linalg.norm(sample_amps - want) / np.linalg.norm(want))
            except Exception as e:  # the baseline is reported, never a gate
                line["cpu_baseline"] = {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                                        "sample": "failed: %s" % str(e)[:200]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
