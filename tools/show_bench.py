"""Print the headline numbers of bench JSON lines: python tools/show_bench.py gpurun_out/x*.json"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print("%s: dev %.0f fps (%.4f ms/step)  e2e %.0f  p50 %.3f ms" % (f.split("/")[-1], d["value"], d["ms_per_step"],
              d["e2e"]["value"], d["latency_ms_p50_sync_batch"] or 0))
        print("  roofline: in-pipeline %.1f TF/s = %.3f of sustained, %.3f of burst; isolated %.1f TF/s (%.3f), conv ms %.3f of eager %.3f" % (
              r["achieved"], r["frac"], r.get("frac_of_burst", 0), r["isolated"]["achieved"], r["isolated"]["frac"],
              r["isolated"]["kernel_ms_per_step"], d["kernels"]["eager_step_ms"]))
        for k, v in d["e2e"].get("variants", {}).items():
            print("  e2e variant %-24s %.0f fps  %.3f ms/step  h2d %d" % (k, v["value"], v["ms_per_step"], v["h2d_bytes_per_step"]))
        k = d["kernels"]
        print("  preprocess %.1f us (%.2f of HBM)  decode+nms %.1f us %s  other %.1f us" % (
              1e3 * k["preprocess"]["ms"], k["preprocess"]["frac_hbm"], 1e3 * k["decode_nms"]["ms"],
              k["decode_nms"].get("parts_ms"), 1e3 * k["other_layers_ms"]))
        print("  slowest convs: " + ", ".join("%s %.1fus %.0fTF" % (c["name"], 1e3 * c["ms"], c["TFLOP/s"]) for c in k["slowest_convs"]))
        print("  parity:", d.get("parity_check"), " clocks:", d.get("clocks"))
    except Exception as e:
        print(f, "ERR", repr(e))
