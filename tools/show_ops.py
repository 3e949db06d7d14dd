"""Per-op table of a bench.py --dump-ops file: python tools/show_ops.py ops.json [B]"""
import json, sys
ops = json.load(open(sys.argv[1]))
tot = sum(o["ms"] for o in ops)
print("%-24s %8s %8s %7s" % ("op", "us", "TFLOP/s", "share"))
for o in ops:
    tf = o["flops"] / (o["ms"] * 1e-3) / 1e12 if o["flops"] else 0
    print("%-24s %8.1f %8.0f %6.1f%%" % (o["name"], o["ms"] * 1e3, tf, 100 * o["ms"] / tot))
print("total %.1f us, conv %.1f us" % (tot * 1e3, 1e3 * sum(o["ms"] for o in ops if o["kind"] == "conv")))
