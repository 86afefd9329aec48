import json, sys
for f in sys.argv[1:]:
    d=json.loads(open(f).read().strip().splitlines()[-1])
    e=d.get('e2e') or {}
    print(f, "ms/step %.1f  value %.0f GF"%(d['ms_per_step'],d['value']), ("e2e %.0f GF (%.1f ms: h2d %.1f fac %.1f d2h %.1f)"%(e['value'],e['ms_per_step'],e['ms_h2d'],e['ms_factor'],e['ms_d2h'])) if e else "")
    print("  roofline", {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['roofline'].items() if k in('achieved','peak','frac','share_of_step','avg_launch_ms')}, "whole", {k:round(v,3) for k,v in d['whole_step'].items()}, "clocks", d['clocks'])
    for k,v in d['kernels'].items(): print("   %-12s ms %8.2f launches %5d gflop %9.1f  -> %.1f TF"%(k,v['ms'],v['launches'],v['gflop'], v['gflop']/max(v['ms'],1e-9)))
    if d.get('cpu_baseline'): print("  cpu", d['cpu_baseline'])
