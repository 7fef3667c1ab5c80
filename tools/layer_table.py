"""Debug tool: per-layer time / TFLOP/s table from an ncu launch list (gpu__time_duration.sum) of one forward and
the plan dump of opb200_net_describe. usage: layer_table.py launches.csv plan.txt batch H W [top]"""
import csv
import io
import sys


def main():
    launches, plan, batch, H, W = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    top = int(sys.argv[6]) if len(sys.argv) > 6 else 100
    txt = [l for l in open(launches).read().splitlines() if not l.startswith("==")]
    rows = list(csv.DictReader(io.StringIO("\n".join(txt))))
    ops = [l.split() for l in open(plan).read().splitlines() if l.split() and l.split()[0] in ("conv", "pool", "convert")]
    n = len(ops)
    first = max(i for i, r in enumerate(rows) if ("firstConv" in r["Kernel Name"] or "inputPatch" in r["Kernel Name"])
                and i + n <= len(rows))
    rows = rows[first:first + n]                       # the last forward
    h, w = H, W
    table, total = [], 0.0
    for op, r in zip(ops, rows):
        ns = float(r["Metric Value"])
        kv = dict(t.split("=", 1) for t in op[2:] if "=" in t)
        kind = op[0]
        flops = 0.0
        if kind == "pool":
            desc = "pool %dx%d" % (h, w)
            h, w = (h + 1) // 2, (w + 1) // 2
        else:
            cin, cout, k = int(kv["cin"]), int(kv["cout"]), int(kv["k"])
            flops = 2.0 * batch * h * w * cin * cout * k * k
            desc = "%s %d->%d k%d @%dx%d%s" % (op[1], cin, cout, k, h, w, "+pool" if kv.get("pool") == "1" else "")
            if kv.get("pool") == "1":
                h, w = (h + 1) // 2, (w + 1) // 2
        assert ("Pool" in r["Kernel Name"]) == (kind == "pool"), (op, r["Kernel Name"])
        table.append((ns, flops, desc, r["Kernel Name"][:28]))
        total += ns
    conv_ns = sum(t[0] for t in table if t[1] > 0)
    conv_fl = sum(t[1] for t in table)
    print("launches %d, total %.3f ms, conv %.3f ms, %.1f GFLOP/frame, %.0f TFLOP/s over conv time" % (
        n, total * 1e-6, conv_ns * 1e-6, conv_fl / batch * 1e-9, conv_fl / conv_ns * 1e-3))
    # group identical shapes
    groups = {}
    for ns, fl, desc, kn in table:
        key = desc.split(" ", 1)[1] if fl > 0 else desc
        g = groups.setdefault(key, [0, 0.0, 0.0])
        g[0] += 1
        g[1] += ns
        g[2] += fl
    for key, (cnt, ns, fl) in sorted(groups.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%-28s x%-3d %8.1f us (%4.1f%%)  %6.0f TF/s  %7.1f us each" % (
            key, cnt, ns * 1e-3, 100 * ns / total, fl / ns * 1e-3 if fl else 0, ns * 1e-3 / cnt))


if __name__ == "__main__":
    main()
