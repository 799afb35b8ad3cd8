#!/usr/bin/env python
"""print the interesting parts of bench.py JSON lines: python tools/show_bench.py FILE..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e); continue
    print(f"== {f}: {d['config']['workload']}  {d['ms_per_step']:.2f} ms/step  value {d['value']:.4g}  e2e {d['e2e']['value']:.4g} ({d['e2e'].get('ms_per_step', 0):.1f} ms)")
    print("  counts", d.get("counts")); print("  stages", d.get("stages_ms_rank0")); print("  stages e2e", d.get("stages_e2e_ms_rank0")); print("  path", d.get("path_roofline")); print("  clocks", d.get("clocks"))
    for k in d.get("kernels", [])[:12]:
        print(f"   {k['kernel']:20s} n={k['launches']:3d} ms={k['ms']:9.3f} algoGB/s={k['algo_gbs']:8.1f} share={k['share']:.3f}")
    if d.get("cpu_baseline"): print("  cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], d["cpu_baseline"]["kind"], "|", d["cpu_baseline"].get("parity"))
    if d.get("digest"): print("  digest", d["digest"].get("check"), d["digest"].get("entries_xor"))
    print("  roofline", d.get("roofline"))
    if d.get("e2e_file"): print("  file", d["e2e_file"])
