"""Per-tile stamps of CTA 0 from bench.py --trace: python tools/show_tiles.py trace.json name [name...]"""
import json, sys
tr = {t["name"]: t for t in json.load(open(sys.argv[1]))}
for name in sys.argv[2:]:
    t = tr.get(name)
    if not t:
        print(name, "not traced"); continue
    a = t["cta0"]
    print("== %s grid %d smem %d KB eager %.1f us | setup %d firstA %d first_ops %d last_mma %d last_store %d exit %d (cycles)" % (
        name, t["grid"], t["smem"] // 1024, t["ms_eager"] * 1e3, a["setup_done"], a["first_A_issued"], a["first_operands"],
        a["last_mma_commit"], a["last_store_issued"] or -1, a["exit"]))
    for k in ("tile_prod", "tile_mma", "tile_acc", "tile_stage_free", "tile_staged", "tile_store"):
        v = t.get(k, [])
        print("  %-16s %s" % (k, " ".join("%6d" % x for x in v[:14])))
    m = t.get("tile_mma", [])
    if len(m) > 3:
        d = [m[i + 1] - m[i] for i in range(len(m) - 1)]
        print("  mma-commit interval per round: median %d cycles" % sorted(d)[len(d) // 2])
