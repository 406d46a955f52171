#!/usr/bin/env python
"""Headline benchmark: contraction throughput on BASELINE.json configs[1]
(bristlecone_70 depth 1-32-1 with the shipped ordering; a FIXED SUBSET of its
65536 cut-slices, 8 slices per GPU, selected through cut value lists exactly as
SURVEY 8d prescribes, identically for this engine and for the reference).

  python bench.py --gpus N --steps K --warmup W            (our arm)
  python bench.py --impl reference --gpus N --steps K --warmup W   (CPU reference arm)

A step = one pass of the schedule over this rank's slices.  `value` is measured
with the network resident in HBM (gate tensors uploaded, grid tensors
flattened); `e2e` is the same metric through qflex_b200_evaluate (the
EvaluateCircuit C-ABI call) with HOST text buffers: parse + plan + host->device
upload + flatten + contract + device->host amplitudes inside the timed region.
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

SLICES_PER_GPU = 8
WORKLOADS = {
    # name -> (circuit, ordering, grid, cut value lists per world size)
    "bristlecone_70_1-32-1": dict(
        circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt",
        cuts={1: [[0, 1], [0, 1], [0, 1], [0]], 2: [[0, 1], [0, 1], [0, 1], [0, 1]],
              4: [[0, 1], [0, 1], [0, 1], [0, 1, 2, 3]],
              8: [[0, 1], [0, 1], [0, 1, 2, 3], [0, 1, 2, 3]]},
        reference_sample=[[0], [0], [0], [0]], memory=64 << 30, total_slices=65536),
    "bristlecone_70_1-24-1": dict(
        circuit="bristlecone_70_1-24-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt",
        cuts={1: [[0, 1], [0, 1], [0, 1], [0]], 2: [[0, 1], [0, 1], [0, 1], [0, 1]],
              4: [[0, 1], [0, 1], [0, 1], [0, 1, 2, 3]],
              8: [[0, 1], [0, 1], [0, 1, 2, 3], [0, 1, 2, 3]]},
        reference_sample=[[0, 1], [0, 1], [0, 1], [0]], memory=8 << 30, total_slices=4096),
    # BASELINE configs[4] stand-in (m=20 is not representable): Sycamore-53
    # m=10, a low-fidelity batch of 64 of its 16384 slices, STRONG scaling (the same
    # 64 slices whatever the GPU count)
    "sycamore_53_10": dict(
        circuit="sycamore_53_10_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1, 2, 3], [0, 1, 2, 3], [0, 1, 2, 3]]},
        reference_sample=[[0], [0], [0]], memory=100 << 30, total_slices=16384),
    # one step up from the m=10 stand-in: Sycamore-53 m=12 (29.6 GB arena here, 7.1e14
    # flop per slice; the reference cannot hold its patches in this host's memory), 8
    # of its slices, strong scaling
    "sycamore_53_12": dict(
        circuit="sycamore_53_12_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1], [0, 1], [0, 1]]},
        reference_sample=None, memory=100 << 30, total_slices=None),
    # BASELINE configs[2] as named: Sycamore-53 m=14.  The shipped 3-cut ordering needs
    # a 2^44-element patch; config/ordering/sycamore_53_m14.txt = that ordering plus the
    # six bond cuts tools/augment_ordering.py adds for a 2^30 limit (largest tensor
    # 2^28, 2.45e14 flop per slice, 2^58 slices, 81 GB arena: the flattened grid
    # tensors).  8 slices x 64 output bitstrings, strong scaling.
    "sycamore_53_14": dict(
        circuit="sycamore_53_14_0.txt", ordering="sycamore_53_m14.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1]] * 3 + [[0]] * 6 + [[0, 1]] * 6},
        reference_sample=None, memory=160 << 30, total_slices=2.0 ** 58),
    # Sycamore-53 m=16: shipped ordering + 22 bond cuts (largest tensor 2^24, 2^200
    # slices, 4.6 GB), cut-aware flatten; 64 slices x 64 bitstrings, strong scaling
    "sycamore_53_16": dict(
        circuit="sycamore_53_16_0.txt", ordering="sycamore_53_m16.txt", grid="sycamore_53.txt",
        scaling="strong", cuts={"all": [[0, 1, 2, 3]] * 3 + [[0]] * 22 + [[0, 1]] * 6},
        reference_sample=None, memory=64 << 30, total_slices=2.0 ** 200),
    # BASELINE configs[4] as named: Sycamore-53 m=20, heavily sliced.  Its grid tensors
    # (four bonds of 2^10) are not representable; config/ordering/sycamore_53_m20.txt =
    # the shipped ordering + 36 bond cuts (tools/augment_ordering.py --max-log2 30
    # --extra 14: largest tensor 2^30, 2^390 slices) and the planner's cut-aware
    # flatten projects them in the gate tensors (120 GB arena).  8 of the slices x 64
    # output bitstrings, strong scaling.
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
    # BASELINE configs[3] as named: 7x7 depth 1-40-1 (bond dimension 32; two bond cuts
    # of 32 values each = 1024 slices).  64.2 GB arena here (the reference's buffer
    # scheme needs 72 GB and ~7e14 flop per slice: it is not timed on the host).
    # 64 output bitstrings per slice; slices grow with the GPU count (weak scaling).
    "rectangular_7x7_1-40-1": dict(
        circuit="rectangular_7x7_1-40-1_0.txt", ordering="rectangular_7x7_v2.txt",
        grid="rectangular_7x7.txt",
        cuts={1: [[0], [0]] + [[0, 1]] * 6, 2: [[0, 1], [0]] + [[0, 1]] * 6,
              4: [[0, 1], [0, 1]] + [[0, 1]] * 6, 8: [[0, 1, 2, 3], [0, 1]] + [[0, 1]] * 6},
        reference_sample=None, memory=160 << 30, total_slices=1024),
}


def workload_cuts(w, world):
    return w["cuts"].get("all") or w["cuts"][world]


def load_texts(w, cut_values):
    from tests import support
    return support.case_texts(dict(circuit=w["circuit"], ordering=w["ordering"], grid=w["grid"],
                                   cut_values=cut_values))


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


def cpu_reference_run(w, cut_values, threads):
    """Times the UNMODIFIED reference (oracle/_ref) on `cut_values` in a child
    process (thread env must be set before OpenBLAS/libgomp load).  Returns
    (seconds, n_slices) of its EvaluateCircuit."""
    code = (
        "import sys, json; sys.path.insert(0, %r)\n"
        "from tests import support\n"
        "import bench\n"
        "w = bench.WORKLOADS[%r]\n"
        "c, o, g = bench.load_texts(w, %r)\n"
        "support.reference_set_threads(%d)\n"
        "res, secs = support.reference_evaluate(c, o, g, memory=%d)\n"
        "print(json.dumps({'secs': secs, 'amps': [[a.real, a.imag] for _, a in res]}))\n"
    ) % (ROOT, [k for k, v in WORKLOADS.items() if v is w][0], cut_values, threads[1], w["memory"])
    env = dict(os.environ, OMP_NUM_THREADS=str(threads[0]), OPENBLAS_NUM_THREADS=str(threads[1]),
               OMP_WAIT_POLICY="passive")
    p = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=3000)
    if p.returncode != 0:
        raise RuntimeError("reference run failed: " + p.stderr[-2000:])
    out = json.loads(p.stdout.strip().splitlines()[-1])
    n = 1
    for v in cut_values:
        n *= len(v)
    return out["secs"], n, out["amps"]


def host_threads():
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        pass
    # OpenMP for the reorder passes, OpenBLAS threads for cgemm; passive waiting
    # avoids the oversubscription collapse recorded in SURVEY section 0.
    return cores, (min(cores, 16), min(cores, 32))


def run_reference_arm(args, w, wname):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores, threads = host_threads()
    sample = w["reference_sample"]
    if sample is None:
        print(json.dumps({"impl": "reference", "unavailable":
                          "%s: the reference needs more host memory and time than a bounded "
                          "sample allows (see the workload's comment in bench.py)" % wname}), flush=True)
        return 0
    times = []
    for i in range(args.warmup + args.steps):
        secs, n_slices, _ = cpu_reference_run(w, sample, threads)
        if i >= args.warmup:
            times.append(secs)
    ms = 1e3 * sum(times) / len(times)
    val = plan_flops(w, sample) / (ms * 1e-3) / 1e12
    line = {
        "impl": "reference", "metric": "contraction_tflops", "value": val, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": w.get("scaling", "weak"), "vs_baseline": None,
        "dtype": "complex64",
        "data": "shipped circuit (config/circuits/%s), deterministic" % w["circuit"],
        "config": {"workload": wname, "slices_per_step": n_slices,
                   "note": "reference qflex EvaluateCircuit on host cores, OpenBLAS + OpenMP"},
        "slices_per_s": n_slices / (ms * 1e-3),
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": max(threads), "kind": "reference",
                         "sample": "%d slice(s) of %s per step, OMP=%d OPENBLAS=%d passive wait, %d host cores"
                                   % (n_slices, wname, threads[0], threads[1], cores)},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


_FLOPS_CACHE = {}


def plan_flops(w, cut_values):
    """Algorithmic flops of one pass of the REFERENCE's schedule over the slices
    `cut_values` selects = sum over the schedule of 8*M*N*K x (how often the step runs: the product of the value
    counts of the cuts that precede it).  Planner's symbolic replay; the plan is only
    described, not contracted."""
    import re
    key = (w["circuit"], json.dumps(cut_values))
    if key not in _FLOPS_CACHE:
        import qflex_b200
        c, o, g = load_texts(w, cut_values)
        fan = []
        for line in o.splitlines():
            m = re.match(r"\s*cut\s*\(([0-9,\s]*)\)", line)
            if m:
                vals = [v for v in m.group(1).split(",") if v.strip()]
                if not vals:
                    raise ValueError("bench workloads must give every cut an explicit value list")
                fan.append(len(vals))
        # the reference contracts projected leaves (one per output bitstring): count
        # the flops of THAT plan, not of the open-output plan the engine itself runs
        saved = os.environ.get("QFLEX_OPEN_OUTPUTS")
        os.environ["QFLEX_OPEN_OUTPUTS"] = "0"
        try:
            try:
                p = qflex_b200.default_library().prepare(c, o, g, memory_limit=w["memory"])
            except qflex_b200.QflexError:
                from tests import support
                p = support.emulated_host_library().prepare(c, o, g, memory_limit=w["memory"])
        finally:
            if saved is None:
                del os.environ["QFLEX_OPEN_OUTPUTS"]
            else:
                os.environ["QFLEX_OPEN_OUTPUTS"] = saved
        total = 0.0
        for r in p.describe():
            if r["phase"] == "grid" and r["kind"] in ("contract", "dot"):
                times = 1
                for f in fan[: r["level"]]:
                    times *= f
                total += times * 8.0 * 2.0 ** (r["m_bits"] + r["n_bits"] + r["k_bits"])
        p.close()
        _FLOPS_CACHE[key] = total
    return _FLOPS_CACHE[key]


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
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference_arm(args, w, args.workload)

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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allreduce_amps(a):
        # one NCCL sum of 2*n_amps doubles (qflex_b200/sharded.py)
        return sharded.allreduce_amplitudes(a, device=dev)

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

    # ---- resident: W warm-up + K timed steps
    for _ in range(args.warmup):
        amps = allreduce_amps(prep.contract(rank, world))
    sampler = ClockSampler(local_rank)
    prep.profile(True)
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    dev_ms, flops_done, launches, leaves, prof, tc_launches = 0.0, 0.0, 0, 0, {}, 0
    for _ in range(args.steps):
        part = prep.contract(rank, world)
        amps = allreduce_amps(part)
        st = prep.stats
        dev_ms += st["device_ms_contract"]
        flops_done += st["flops"]
        launches += int(st["launches"])
        leaves += int(st["leaves"])
        tc_launches += int(st["launches_gemm_tc"])
        for r in prep.profile_rows():
            key = (int(r["kind"]), int(r["m_bits"]), int(r["n_bits"]), int(r["k_bits"]))
            acc = prof.setdefault(key, dict(launches=0, ms=0.0, flops=0.0, bytes=0.0))
            for f in ("launches", "ms", "flops", "bytes"):
                acc[f] += r[f]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    prep.profile(False)
    wall = max_over_ranks(wall)
    total_flops = sum_over_ranks(flops_done)
    total_leaves = sum_over_ranks(leaves)
    total_launches = int(sum_over_ranks(launches))
    ms_per_step = 1e3 * wall / args.steps
    value = total_flops / wall / 1e12
    arena_bytes = prep.stats["arena_bytes"]
    n_amps = len(prep.states)
    prep.close()

    # ---- e2e: EvaluateCircuit through the C ABI from host text buffers
    h2d = 0
    e2e_steps = max(1, min(args.steps, 2))
    lib.evaluate(c, o, g, memory_limit=w["memory"], device=local_rank, rank=rank, world=world,
                 precision=args.precision)  # warm-up (allocator, module load)
    barrier()
    t0 = time.perf_counter()
    e2e_flops = 0.0
    for _ in range(e2e_steps):
        res, st = lib.evaluate(c, o, g, memory_limit=w["memory"], device=local_rank, rank=rank,
                               world=world, precision=args.precision)
        part = np.array([a for _, a in res])
        full = allreduce_amps(part)
        e2e_flops += st["flops"]
    barrier()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e_value = sum_over_ranks(e2e_flops) / e2e_wall / 1e12
    h2d = gate_upload_bytes(lib, c, o, g, w)
    parity = float(np.linalg.norm(full - amps) / max(np.linalg.norm(amps), 1e-300))

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
        # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel class
        # from one `ncu --set full` capture (profiles/ncu_traffic.json names the file)
        with open(tpath) as f:
            t = json.load(f).get("%d,%d,%d" % dom_key[1:])
        if t:
            roof["traffic"] = t["dram_bytes_per_launch"]
            roof["traffic_source"] = t["source"]
            roof["algorithmic_bytes_per_launch"] = dom["bytes"] / dom["launches"]
    roof["kernel"] = "contract m=%d n=%d k=%d bits" % dom_key[1:]
    roof["launches"] = int(dom["launches"])
    roof["ms_per_launch"] = per_launch_ms
    roof["share_of_device_time"] = dom["ms"] / sum(v["ms"] for v in prof.values())
    roof["arithmetic_intensity"] = ai
    roof["peak_source"] = ("%s: hbm_gbs %.1f (MEASURED_PEAKS.json); FFMA peak derived = "
                           "148 SM x 128 lanes x 2 x %.0f MHz" % (peak_src, peaks["hbm_gbs"], sm_max))

    line = {
        "metric": "contraction_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": w.get("scaling", "weak"), "vs_baseline": None,
        "dtype": "complex64",
        "data": "shipped circuit (config/circuits/%s), deterministic; synthetic = none needed" % w["circuit"],
        "config": {"workload": args.workload,
                   "slices_per_gpu": total_leaves / args.steps / world / max(1, n_amps),
                   "slices_per_step": int(total_leaves / args.steps), "cut_values": cuts,
                   "arena_gb": arena_bytes / 2 ** 30, "precision": args.precision,
                   "l2": "working set per contraction (>= 2 GB patches) exceeds the 126 MB L2",
                   "collective": "one NCCL all-reduce(sum) of %d complex128 per step" % n_amps},
        # leaves = (bond-cut slices) x (output bitstrings): one leaf = one scalar result
        "slices_per_s": total_leaves / wall,
        # amplitudes/s of the FULL circuit at this rate: every output bitstring needs all
        # of the circuit's bond-cut slices
        "amplitudes_per_s_full_circuit": (total_leaves / wall / float(w["total_slices"])
                                          if w.get("total_slices") else None),
        "device_ms_per_step_rank0": dev_ms / args.steps,
        "e2e": {"value": e2e_value, "unit": "TFLOP/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 16 * n_amps, "steps": e2e_steps,
                "ms_per_step": 1e3 * e2e_wall / e2e_steps,
                "parity_vs_resident_run": parity},
        "gpu_launches": total_launches,
        "clocks": clocks, "roofline": roof,
        "kernel_classes": [dict(kind=k[0], m_bits=k[1], n_bits=k[2], k_bits=k[3],
                                launches=int(v["launches"]), ms=v["ms"],
                                tflops=v["flops"] / v["ms"] / 1e9 if v["ms"] else None,
                                gbs=v["bytes"] / v["ms"] / 1e6 if v["ms"] else None)
                           for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:8]],
    }
    if world == 1 and not args.no_cpu_baseline and w["reference_sample"] is None:
        line["cpu_baseline"] = {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                                "sample": "not run: no bounded sample of this workload fits the host"}
    elif world == 1 and not args.no_cpu_baseline:
        try:
            cores, threads = host_threads()
            secs, n_slices, ref_amps = cpu_reference_run(w, w["reference_sample"], threads)
            fl = plan_flops(w, w["reference_sample"])
            line["cpu_baseline"] = {
                "value": fl / secs / 1e12, "unit": "TFLOP/s", "cores": max(threads),
                "kind": "reference", "seconds": secs,
                "sample": "%d slice(s) of %s (cut values %s), reference EvaluateCircuit, "
                          "OMP=%d OPENBLAS=%d passive wait, %d host cores"
                          % (n_slices, args.workload, w["reference_sample"], threads[0], threads[1], cores)}
        except Exception as e:  # the baseline is reported, never a gate
            line["cpu_baseline"] = {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                                    "sample": "failed: %s" % str(e)[:200]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def gate_upload_bytes(lib, c, o, g, w):
    """Bytes of gate tensors copied host->device per EvaluateCircuit call."""
    import re
    n1 = len(re.findall(r"^\s*\d+\s+(h|t|x_1_2|y_1_2|hz_1_2|rz\S*)\s", c, flags=re.M))
    n2 = len(re.findall(r"^\s*\d+\s+(cz|cx)\s", c, flags=re.M))
    nf = len(re.findall(r"^\s*\d+\s+fsim", c, flags=re.M))
    nq = int(c.split()[0])
    return 8 * (4 * n1 + 2 * 8 * n2 + 2 * 16 * nf + 2 * 2 * nq)


if __name__ == "__main__":
    sys.exit(main())
