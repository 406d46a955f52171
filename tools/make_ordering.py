"""Contraction ordering (expand / merge / cut) from a circuit and a grid alone.

SURVEY 8f-4.  The reference's generator (qflexcirq/ordering/order_circuit_simulation.py:
206-335: create_ordering_data, get_steps_for_graph, match_fidelity, circuit_to_ordering)
works on cirq objects and needs cirq only for the Schmidt ranks of the two-qubit gates;
this tool works on the qFlex text formats directly (rank 2 for cz/cx, 4 for fsim, as in
the gate tables of src/read_circuit.cpp:296-544) and adds what the large configurations
need: a MAX-TENSOR constraint (cuts are added until every intermediate fits
2^max_log2 elements) and a flop/slice report.

Method.  The qubits of the grid are the nodes of a weighted graph; the weight of an
edge is log2 of the bond dimension between the two world lines (sum over the gates on
that pair).  A schedule is a sequence of edge contractions; it is grown greedily by the
reference's three-level criterion -- (size of the result / size of its two inputs,
few neighbours per result element, multiplication count) -- and written out as qFlex
commands: contracting two fresh qubits opens a patch, a patch and a fresh qubit is an
`expand`, two patches a `merge`.  Cuts are chosen greedily over the bonds: with a size
limit, the bond whose removal brings the largest intermediate closest to the limit
(ties: least total work = slices x multiplications); without one, the reference's rule
(a cut must lower the cost of one slice).  `fidelity` < 1 keeps only a prefix of each
cut's values so that the kept fraction is the smallest one >= fidelity.

  python tools/make_ordering.py CIRCUIT GRID [--max-log2 28] [--max-cuts 2]
         [--fidelity 1.0] [--outputs 5,6,14] [-o ORDERING]
"""
import argparse
import itertools
import math
import re
import sys


def parse_grid(text):
    rows = [[int(x) for x in line.split()] for line in text.splitlines() if line.strip()]
    width = len(rows[0])
    return len(rows), width, sorted(i * width + j for i, r in enumerate(rows)
                                    for j, v in enumerate(r) if v)


def bond_bits(circuit_text):
    """{(a, b): log2 bond dimension} with a < b, from the two-qubit gates."""
    bits = {}
    for line in circuit_text.splitlines()[1:]:
        m = re.match(r"\s*\d+\s+([a-z_0-9]+)(\(.*\))?\s+(\d+)\s+(\d+)\s*$", line)
        if m:
            a, b = sorted((int(m.group(3)), int(m.group(4))))
            bits[(a, b)] = bits.get((a, b), 0) + (2 if m.group(1).startswith("fs") else 1)
    return bits


class PatchGraph:
    """Nodes are sets of qubits, `link[u][v]` the bits of the bonds between two nodes."""

    def __init__(self, qubits, bits, removed=()):
        self.members = {q: frozenset([q]) for q in qubits}
        self.link = {q: {} for q in qubits}
        for (a, b), w in bits.items():
            if (a, b) in removed or a not in self.link or b not in self.link:
                continue
            self.link[a][b] = self.link[a].get(b, 0) + w
            self.link[b][a] = self.link[b].get(a, 0) + w

    def rank(self, u):
        return sum(self.link[u].values())

    def score(self, u, v):
        """The reference's (heuristic, estimate, time) triple for contracting u-v
        (order_circuit_simulation.py:88-108), in log2 where that is exact."""
        shared = self.link[u][v]
        ru, rv = self.rank(u), self.rank(v)
        result = ru + rv - 2 * shared
        heuristic = 2.0 ** result / (2.0 ** ru + 2.0 ** rv)
        neighbours = len((set(self.link[u]) | set(self.link[v])) - {u, v})
        estimate = neighbours / 2.0 ** result if neighbours else 0.0
        return heuristic, estimate, ru + rv - shared

    def contract(self, u, v):
        """v is absorbed by u."""
        for x, w in self.link.pop(v).items():
            del self.link[x][v]
            if x != u:
                self.link[u][x] = self.link[u].get(x, 0) + w
                self.link[x][u] = self.link[u][x]
        self.members[u] = self.members[u] | self.members.pop(v)


def greedy_schedule(qubits, bits, removed=()):
    """Returns (steps, log2-sum of multiplications, largest intermediate in log2);
    steps = [(qubits of one side, qubits of the other)] in contraction order."""
    g = PatchGraph(qubits, bits, removed)
    steps, work, largest = [], 0.0, 0
    for q in qubits:
        largest = max(largest, g.rank(q))
    while True:
        best = None
        for u in sorted(g.link):
            for v in sorted(g.link[u]):
                if u < v:
                    key = g.score(u, v) + (u, v)
                    if best is None or key < best:
                        best = key
        if best is None:
            break
        u, v = best[3], best[4]
        work += 2.0 ** best[2]
        steps.append((g.members[u], g.members[v]))
        g.contract(u, v)
        largest = max(largest, g.rank(u))
    return steps, work, largest


def choose_cuts(qubits, bits, max_cuts, max_log2):
    """Greedy cut selection.  Returns ([(a, b)], steps, work per slice, largest)."""
    cuts = []
    steps, work, largest = greedy_schedule(qubits, bits)
    while True:
        over = max_log2 is not None and largest > max_log2
        if not over and (max_log2 is not None or len(cuts) >= max_cuts):
            break
        if len(cuts) >= len(bits):
            break
        slices = 2.0 ** sum(bits[c] for c in cuts)
        best = None
        for bond in sorted(bits):
            if bond in cuts:
                continue
            s2, w2, l2 = greedy_schedule(qubits, bits, removed=set(cuts) | {bond})
            if max_log2 is not None:
                key = (max(0, l2 - max_log2), slices * 2.0 ** bits[bond] * w2)
            else:
                key = (w2,)
            if best is None or key < best[0]:
                best = (key, bond, s2, w2, l2)
        if best is None:
            break
        if max_log2 is None and not best[3] < work:
            break  # the reference's rule: a cut must make one slice cheaper
        cuts.append(best[1])
        steps, work, largest = best[2], best[3], best[4]
    return cuts, steps, work, largest


def widths_for_fidelity(dims, fidelity):
    """Per-cut number of kept values: the product of kept fractions is the smallest one
    that is still >= fidelity (role of match_fidelity, order_circuit_simulation.py:
    279-318).  Returns (fraction, widths)."""
    if fidelity >= 1.0 or not dims:
        return 1.0, list(dims)
    ranges = [range(min(d, math.ceil(d * fidelity)), d + 1) for d in dims]
    total = float(math.prod(dims))
    best = (1.0, list(dims))
    for combo in itertools.product(*ranges):
        frac = math.prod(combo) / total
        if fidelity <= frac < best[0]:
            best = (frac, list(combo))
    return best


def emit(steps, qubits):
    """Edge contractions -> expand / merge lines.  Patch names are A, B, ... AA, AB ..."""
    def name(n):
        s = ""
        n += 1
        while n:
            n, r = divmod(n - 1, 26)
            s = chr(ord("A") + r) + s
        return s
    patch_of, lines, n_patches = {}, [], 0
    for left, right in steps:
        pl = {patch_of[q] for q in left if q in patch_of}
        pr = {patch_of[q] for q in right if q in patch_of}
        assert len(pl) <= 1 and len(pr) <= 1
        if not pl and not pr:
            p = name(n_patches)
            n_patches += 1
            for q in sorted(left | right):
                lines.append("expand %s %d" % (p, q))
                patch_of[q] = p
        elif pl and pr:
            src, dst = sorted((pl.pop(), pr.pop()))
            lines.append("merge %s %s" % (src, dst))
            for q in patch_of:
                if patch_of[q] == src:
                    patch_of[q] = dst
        else:
            p = (pl or pr).pop()
            for q in sorted((left | right) - set(patch_of)):
                lines.append("expand %s %d" % (p, q))
                patch_of[q] = p
    # world lines without any two-qubit gate still have to be contracted
    idle = [q for q in qubits if q not in patch_of]
    if idle:
        p = patch_of[next(iter(patch_of))] if patch_of else name(0)
        for q in idle:
            lines.append("expand %s %d" % (p, q))
            patch_of[q] = p
    # cuts may have separated the graph: the pieces are multiplied at the end (the
    # schedule's result is its LAST patch, src/contraction_utils.cpp:340-352)
    pieces = sorted(set(patch_of.values()))
    for src in pieces[:-1]:
        lines.append("merge %s %s" % (src, pieces[-1]))
    return lines


def make_ordering(circuit_text, grid_text, max_log2=None, max_cuts=2, fidelity=1.0, outputs=()):
    """Returns (ordering text, info dict)."""
    _, _, qubits = parse_grid(grid_text)
    bits = {k: v for k, v in bond_bits(circuit_text).items()
            if k[0] in set(qubits) and k[1] in set(qubits)}
    cuts, steps, work, largest = choose_cuts(qubits, bits, max_cuts, max_log2)
    dims = [2 ** bits[c] for c in cuts]
    frac, widths = widths_for_fidelity(dims, fidelity)
    out = ["# generated by tools/make_ordering.py: %d bond cut(s), largest tensor 2^%d, "
           "%.3e multiplications per slice, %d slice(s)%s"
           % (len(cuts), largest, work, math.prod(widths),
              "" if frac >= 1.0 else ", kept fraction %.4f" % frac)]
    for (a, b), d, wdt in zip(cuts, dims, widths):
        vals = "" if wdt >= d else ",".join(str(v) for v in range(wdt))
        out.append("cut (%s) %d %d" % (vals, a, b))
    for q in outputs:  # qubit id, or (qubit id, values) for a restricted terminal cut
        q, vals = (q, ()) if isinstance(q, int) else q
        out.append("cut (%s) %d" % (",".join(str(v) for v in vals), q))
    out += emit(steps, qubits)
    info = dict(cuts=cuts, cut_dims=dims, widths=widths, fidelity=frac, max_log2=largest,
                flops_per_slice=8.0 * work, slices=math.prod(widths))
    return "\n".join(out) + "\n", info


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("circuit")
    ap.add_argument("grid")
    ap.add_argument("--max-log2", type=int, default=None,
                    help="largest intermediate tensor, log2 of its element count")
    ap.add_argument("--max-cuts", type=int, default=2,
                    help="without --max-log2: number of cuts to attempt (reference -x)")
    ap.add_argument("--fidelity", type=float, default=1.0)
    ap.add_argument("--outputs", default="", help="comma-separated qubit ids with a terminal cut")
    ap.add_argument("-o", "--output-file", default=None)
    args = ap.parse_args()
    outputs = [int(x) for x in args.outputs.split(",") if x.strip()]
    text, info = make_ordering(open(args.circuit).read(), open(args.grid).read(),
                               args.max_log2, args.max_cuts, args.fidelity, outputs)
    if args.output_file:
        with open(args.output_file, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)
    print("cuts %s, largest tensor 2^%d, %.3e flop per slice x %d slices"
          % (info["cuts"], info["max_log2"], info["flops_per_slice"], info["slices"]), file=sys.stderr)


if __name__ == "__main__":
    main()
