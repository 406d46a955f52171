"""Add bond cuts to a contraction ordering until every intermediate tensor fits.

SURVEY 8f-4 (the reference's python/ordering heuristics need cirq; only the Schmidt
ranks of the two-qubit gates matter here: 2 for cz/cx, 4 for fsim, as in the gate
tables of src/read_circuit.cpp:296-544).  The ordering's expand/merge sequence is
replayed on the grid graph: a patch is a set of sites, its tensor has one index per
bond that leaves the set, of dimension prod(rank of the gates on that bond).  While
the largest intermediate exceeds 2^max_log2 elements, the bond whose cut shrinks the
total work (slices x flops per slice) least is cut: `cut () a b` lines are
put at the top of the ordering (every site still unused there, which is what
Tensor::project requires, src/tensor.cpp:280-285).  Each cut multiplies the number of
slices by the bond dimension and leaves the amplitude (the sum over slices) unchanged.

  python tools/augment_ordering.py CIRCUIT GRID ORDERING --max-log2 30 -o NEW_ORDERING
"""
import argparse
import math
import re
import sys


def parse_grid(text):
    rows = [[int(x) for x in line.split()] for line in text.splitlines() if line.strip()]
    J = len(rows[0])
    active = {i * J + j for i, r in enumerate(rows) for j, v in enumerate(r) if v}
    return len(rows), J, active


def bond_bits(circuit_text):
    """log2 of the bond dimension between every pair of interacting qubits."""
    bits = {}
    for line in circuit_text.splitlines()[1:]:
        m = re.match(r"\s*\d+\s+([a-z_0-9]+)(\(.*\))?\s+(\d+)\s+(\d+)\s*$", line)
        if not m:
            continue
        gate, a, b = m.group(1), int(m.group(3)), int(m.group(4))
        rank_bits = 2 if gate.startswith("fs") else 1
        key = (min(a, b), max(a, b))
        bits[key] = bits.get(key, 0) + rank_bits
    return bits


def parse_ordering(text):
    ops = []
    for line in text.splitlines():
        t = line.split("#")[0].split()
        if not t:
            continue
        if t[0] == "cut":
            qs = [int(x) for x in line.split(")")[1].split()]
            ops.append(("cut", tuple(sorted(qs))))
        elif t[0] == "expand":
            ops.append(("expand", t[1], int(t[2])))
        elif t[0] == "merge":
            ops.append(("merge", t[1], t[2]))
    return ops


def replay(ops, bits, cut):
    """[(log2 size of the result, log2 flops/8, frozenset of sites)] per contraction."""
    patches, out = {}, []

    def boundary(sites):
        b = 0
        for (x, y), w in bits.items():
            if (x, y) not in cut and ((x in sites) != (y in sites)):
                b += w
        return b

    def between(s1, s2):
        b = 0
        for (x, y), w in bits.items():
            if (x, y) not in cut and ((x in s1 and y in s2) or (x in s2 and y in s1)):
                b += w
        return b

    for op in ops:
        if op[0] == "expand":
            new = frozenset([op[2]])
            cur = patches.get(op[1])
            if cur is None:
                patches[op[1]] = new
                out.append((boundary(new), 0, new))
                continue
            merged = cur | new
            out.append((boundary(new), 0, new))  # the grid tensor itself (cut-aware flatten)
            out.append((boundary(merged), boundary(merged) + between(cur, new), merged))
            patches[op[1]] = merged
        elif op[0] == "merge":
            src, dst = patches[op[1]], patches.get(op[2])
            if dst is None:
                patches[op[2]] = src
                continue
            merged = src | dst
            out.append((boundary(merged), boundary(merged) + between(src, dst), merged))
            patches[op[2]] = merged
    return out


def cost(steps):
    return max(s[0] for s in steps), sum(2.0 ** s[1] for s in steps)


def augment(circuit_text, grid_text, ordering_text, max_log2, extra=0):
    bits = bond_bits(circuit_text)
    ops = parse_ordering(ordering_text)
    cut = {op[1] for op in ops if op[0] == "cut" and len(op[1]) == 2}
    added = []
    while True:
        steps = replay(ops, bits, cut)
        worst, _ = cost(steps)
        if worst <= max_log2:
            if extra <= 0:
                break
            extra -= 1  # keep going: `extra` more cuts on the largest tensor (arena size)
        big = max(steps, key=lambda s: s[0])[2]
        cands = [k for k in bits if k not in cut and ((k[0] in big) != (k[1] in big))]
        if not cands:
            raise SystemExit("no bond left to cut")
        # total work over all slices first (a cut of w bits costs 2^w more slices),
        # then the largest tensor

        def key(k):
            worst_k, work_k = cost(replay(ops, bits, cut | {k}))
            return (round(bits[k] + math.log2(work_k), 6), worst_k)

        best = min(cands, key=key)
        cut.add(best)
        added.append(best)
    steps = replay(ops, bits, cut)
    info = dict(max_log2=cost(steps)[0], flops_per_slice=8.0 * cost(steps)[1],
                log2_slices=sum(bits[k] for k in cut), added=added)
    lines = ordering_text.splitlines()
    first = next((i for i, l in enumerate(lines) if l.split("#")[0].strip()), len(lines))
    new = lines[:first] + ["cut () %d %d" % k for k in added] + lines[first:]
    return "\n".join(new) + "\n", info


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("circuit")
    ap.add_argument("grid")
    ap.add_argument("ordering")
    ap.add_argument("--max-log2", type=int, default=30)
    ap.add_argument("--extra", type=int, default=0, help="more cuts after the limit is met")
    ap.add_argument("-o", "--output")
    a = ap.parse_args()
    text, info = augment(open(a.circuit).read(), open(a.grid).read(), open(a.ordering).read(),
                         a.max_log2, a.extra)
    print("largest tensor 2^%d elements, %.3g flop per slice, 2^%d slices, cuts added: %s"
          % (info["max_log2"], info["flops_per_slice"], info["log2_slices"],
             " ".join("%d-%d" % k for k in info["added"]) or "none"), file=sys.stderr)
    if a.output:
        with open(a.output, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)


if __name__ == "__main__":
    main()
