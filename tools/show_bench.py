"""Print the headline numbers of bench JSON lines: python tools/show_bench.py gpurun_out/x*.json"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
        print("%-44s dev %6.0f  e2e %6.0f  ms/step %.3f  p50 %.3f  conv-frac %.4f  conv-ms %.3f  other %.3f  post %.3f" % (
            f.split("/")[-1], d["value"], d["e2e"]["value"], d["ms_per_step"], d["latency_ms_p50_sync_batch"] or 0,
            d["roofline"]["frac"], d["roofline"].get("kernel_ms_per_step") or d["roofline"]["isolated"]["kernel_ms_per_step"], d["kernels"]["other_layers_ms"],
            d["kernels"]["decode_nms"]["ms"]))
    except Exception as e:
        print(f, "ERR", e)
