#!/usr/bin/env python
"""Summarises an ncu launch list (`--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`)
of `bench.py`: per-kernel time share and DRAM traffic, and the traffic of one train wave / one aggregate launch.

    python scripts/ncu_traffic.py launches.csv NODES [--update profiles/traffic.json] [--md out.md]
"""
import collections, csv, json, os, sys


def main():
    path, nodes = sys.argv[1], int(sys.argv[2])
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10 and r[0] != "ID"]
    launches = collections.OrderedDict()
    for r in rows:
        d = launches.setdefault(int(r[0]), {"name": r[4], "grid": r[8].replace(" ", "")})
        try:
            d[r[12]] = float(r[14].replace(",", ""))
        except ValueError:
            pass
    L = [v for v in launches.values() if "gpu__time_duration.sum" in v]

    def short(n):
        n = n.replace("void ", "").replace("dls::", "").replace("gnl::", "")
        return n.split("(")[0]

    is_agg = lambda v: "agg_" in v["name"]
    ours = [v for v in L if "dls::" in v["name"] or "gnl::" in v["name"] or "fem::" in v["name"]]
    # train waves = runs of our launches between aggregate launches
    waves, cur = [], []
    for v in ours:
        if is_agg(v):
            if cur:
                waves.append(cur)
            cur = []
        else:
            cur.append(v)
    if cur:
        waves.append(cur)
    full = [w for w in waves if len(w) == max(len(x) for x in waves)]
    tr = lambda w: sum(v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0) for v in w)
    train_traffic = sum(tr(w) for w in full) / len(full)
    aggs = [v for v in ours if is_agg(v)]
    agg_traffic = sum(v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0) for v in aggs) / max(1, len(aggs))
    by = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for v in ours:
        b = by[short(v["name"]) + " " + v["grid"]]          # train and test waves launch the same kernels on different grids
        b[0] += 1; b[1] += v["gpu__time_duration.sum"]
        b[2] += v.get("dram__bytes_read.sum", 0); b[3] += v.get("dram__bytes_write.sum", 0)
    tot = sum(b[1] for b in by.values())
    lines = ["| kernel | launches | us / launch | share % | DRAM read MB / launch | DRAM write MB / launch |", "|---|---|---|---|---|---|"]
    for k, b in sorted(by.items(), key=lambda kv: -kv[1][1]):
        lines.append("| %s | %d | %.1f | %.1f | %.1f | %.1f |" % (k, b[0], b[1] / b[0] / 1e3, 100 * b[1] / tot, b[2] / b[0] / 1e6, b[3] / b[0] / 1e6))
    out = "\n".join(lines)
    print(out)
    print("\ntrain wave: %d launches, DRAM traffic %.1f MB per wave (%d full waves); aggregate: %.1f MB per launch (%d launches)"
          % (len(full[0]), train_traffic / 1e6, len(full), agg_traffic / 1e6, len(aggs)))
    if "--update" in sys.argv:
        tj = sys.argv[sys.argv.index("--update") + 1]
        cur = json.load(open(tj)) if os.path.exists(tj) else {}
        src = os.path.basename(path)
        cur["train_gnlenet_%d" % nodes] = {"bytes": int(train_traffic), "source": "profiles/" + src,
                                            "what": "dram__bytes_read.sum + dram__bytes_write.sum over the %d launches of one train wave" % len(full[0])}
        cur["agg_gnlenet_%d" % nodes] = {"bytes": int(agg_traffic), "source": "profiles/" + src,
                                          "what": "dram__bytes_read.sum + dram__bytes_write.sum of one aggregate launch"}
        json.dump(cur, open(tj, "w"), indent=1, sort_keys=True)
    if "--md" in sys.argv:
        open(sys.argv[sys.argv.index("--md") + 1], "w").write(out + "\n")


if __name__ == "__main__":
    main()
