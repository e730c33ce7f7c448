"""Per-rank, per-phase accounting of cap_potrf_1d timelines (CAPITAL_B200_PROFILE=2 + CAPITAL_B200_PROFILE_FILE=prefix writes one
file per rank).    python tools/timeline_summary.py profiles/r02_timeline_n8/tl_n8.rank*.txt

For every rank: busy milliseconds per phase, and for the chain (csolve + cdiag + leaf on the panel stream) how much of it ran
while the bulk update (U2) of the same rank was running, i.e. was hidden.  Then, over all ranks, the critical path of the
diagonal chain: when each panel's X became available (end of its leaf on its owner) and the gap to the previous panel."""
import re
import sys


def parse(path):
    ev = []
    for line in open(path):
        m = re.match(r"\s*panel\s+(\d+) \(owner (\d+)\):(.*)", line)
        if not m:
            continue
        p, own = int(m.group(1)), int(m.group(2))
        for name, a, b in re.findall(r"(\w+) ([\d.]+)\.\.([\d.]+)", m.group(3)):
            ev.append((p, own, name, float(a), float(b)))
    return ev


def overlap(a0, a1, spans):
    return sum(max(0.0, min(a1, b1) - max(a0, b0)) for b0, b1 in spans)


def main():
    files = sorted(sys.argv[1:], key=lambda f: int(re.search(r"rank(\d+)", f).group(1)))
    leaf_end = {}
    print("%-5s %8s | %s | %s" % ("rank", "span ms", "busy ms per phase", "chain ms (hidden under this rank's U2)"))
    for f in files:
        rank = int(re.search(r"rank(\d+)", f).group(1))
        ev = parse(f)
        if not ev:
            continue
        span = max(e[4] for e in ev) - min(e[3] for e in ev)
        busy = {}
        for p, own, name, a, b in ev:
            busy[name] = busy.get(name, 0.0) + (b - a)
            if name == "leaf":
                leaf_end[p] = (b, rank, min(e[3] for e in ev))
        u2 = [(a, b) for p, own, name, a, b in ev if name == "U2"]
        chain = [(a, b) for p, own, name, a, b in ev if name in ("csolve", "cdiag", "leaf")]
        ch = sum(b - a for a, b in chain)
        hid = sum(overlap(a, b, u2) for a, b in chain)
        order = ["U2", "U1", "solve", "xchg", "leaf", "csolve", "cdiag"]
        print("%-5d %8.2f | %s | %.2f (%.2f)" % (rank, span, "  ".join("%s %.2f" % (k, busy.get(k, 0.0)) for k in order), ch, hid))
    # chain cadence: every rank's clock starts at its own first stamp, so only per-owner differences of consecutive panels owned by
    # the same rank are exact; the cadence below uses each leaf's end on its owner's clock and is good to the skew of the starts (< 1 ms)
    if leaf_end:
        ks = sorted(leaf_end)
        gaps = [(k, leaf_end[k][0] - leaf_end[k - 1][0]) for k in ks if k - 1 in leaf_end]
        third = max(1, len(gaps) // 3)
        for name, part in (("first third", gaps[:third]), ("middle third", gaps[third:2 * third]), ("last third", gaps[2 * third:])):
            if part:
                print("chain cadence, %s of the panels: %.2f ms per panel (panels %d..%d)" % (name, sum(g for _, g in part) / len(part), part[0][0], part[-1][0]))
        print("last leaf ends at %.2f ms" % max(v[0] for v in leaf_end.values()))


if __name__ == "__main__":
    main()
