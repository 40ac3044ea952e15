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
    for o in d.get("other_modes", []): print("   other:", {k: (round(v, 5) if isinstance(v, float) else v) for k, v in o.items()})
    if "keypoint_delta_vs_fp32_mm_max" in d: print("   this mode keypoint delta vs fp32-equiv (mm):", d["keypoint_delta_vs_fp32_mm_max"])
