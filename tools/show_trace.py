"""Print conv in-kernel timelines written by bench.py --trace (microseconds from kernel entry)."""
import json
import sys

tr = json.load(open(sys.argv[1]))
clk = 1.965e3
keys = ("setup_done", "first_A_issued", "last_load_issued", "first_operands", "last_mma_commit", "first_acc_ready",
        "last_store_issued", "stores_drained", "exit")
pro = mma = epi = wall = 0
print("name grid smemKB eager_us | cta0: " + " ".join(keys) + " wall")
for t in tr:
    a = t["cta0"]
    pro += a["first_operands"] / clk
    mma += (a["last_mma_commit"] - a["first_operands"]) / clk
    epi += (a["exit"] - a["last_mma_commit"]) / clk
    wall += a["wall_ns"] / 1000
    print("%-20s %4d %4d %6.1f | %s w=%.2f" % (t["name"], t["grid"], t["smem"] // 1024, t["ms_eager"] * 1000,
          " ".join("%6.2f" % (a[k] / clk) if a[k] is not None else "     -" for k in keys), a["wall_ns"] / 1000))
print("sum: prologue %.0f  mma-phase %.0f  epilogue-tail %.0f  wall %.0f us" % (pro, mma, epi, wall))
