"""One-line summaries of bench.py output files:  python tools/bench_summary.py profiles/r02_bench_v3.json ..."""
import json, sys
for path in sys.argv[1:]:
    try:
        d = json.loads(open(path).read().strip().splitlines()[-1])
    except Exception as e:
        print(path, "unreadable", e); continue
    r = d.get("roofline", {})
    print(f"{path}: mode={d['config']['numeric_mode']} value={d['value']:.0f} e2e={d['e2e']['value']:.0f} ms/step={d['ms_per_step']:.2f} "
          f"conv_ms={r.get('conv_ms_per_pass', 0):.3f} convTF={r.get('achieved', 0):.1f} stem={r.get('stem_ms', 0):.3f} proj={r.get('proj_ms', 0):.3f}")
    print("   per-layer TF:", [round(x) for x in r.get("per_layer_tflops", [])])
    for name, o in d.get("modes", {}).items():
        print(f"   mode {name}: value={o['value']:.0f} e2e={o['e2e']:.0f} keypoint delta vs fp32-equivalent {o.get('keypoint_delta_vs_fp32_mm_max')} mm")
    if "keypoint_delta_vs_fp32_mm_max" in d: print("   this mode keypoint delta vs fp32-equiv (mm):", d["keypoint_delta_vs_fp32_mm_max"])
