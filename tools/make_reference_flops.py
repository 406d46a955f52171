"""Writes bench_reference_flops.json: the algorithmic flops (sum of 8*M*N*K)
of the REFERENCE's schedule -- projected leaves, slice by slice -- on the bounded
sample bench.py times the reference on, per workload.  Planner's symbolic replay on
the CPU-emulated ABI (nothing is contracted); bench.py's reference arm only reads the
JSON, so that it loads nothing of the product.

    python tools/make_reference_flops.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tests import support  # noqa: E402


def main():
    os.environ["QFLEX_OPEN_OUTPUTS"] = "0"   # the reference contracts one leaf per bitstring
    os.environ["QFLEX_FUSE_BITS"] = "0"      # ... and one slice at a time
    lib = support.emulated_host_library()
    out = {}
    for name, w in bench.WORKLOADS.items():
        sample = w["reference_sample"]
        if sample is None:
            continue
        c, o, g = bench.load_texts(w, sample)
        p = lib.prepare(c, o, g, memory_limit=w["memory"])
        flops = p.plan_stats()["flops_job"]
        p.close()
        out[name + "|" + json.dumps(sample)] = flops
        print("%-28s %-40s %.4e flop" % (name, json.dumps(sample), flops))
    with open(bench.FLOPS_CACHE, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
