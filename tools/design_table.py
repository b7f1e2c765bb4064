"""Rows of DESIGN.md section 7's round table from the bench lines of one round:
    python tools/design_table.py profiles/bench_r2_final_{1,2,4,8}gpu.json"""
import json, sys
L = {}
for f in sys.argv[1:]:
    d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    L[d['n_gpus']] = d
def M(x): return f"{x / 1e6:.1f} M"
def T(x): return f"{x / 1e12:.2f}e12"
one = L[1]
def c2(d): return f"{M(d['value'])} env-steps/s ({1e3 * d['ms_per_step']:.1f} µs/step), e2e {M(d['e2e']['value'])}"
def c3(d): x = d['c3']; return f"{M(x['value'])} ({x['ms_per_step']:.3f} ms/step), e2e {M(x['e2e']['value'])}"
def c5(d): x = d['c5']; return f"{T(x['value'])} pair-evals/s ({x['ms_per_step']:.3f} ms)"
rows = [("C2 Horizontal 4096 × 20 per GPU (weak)", c2, lambda d: f"{d['roofline']['frac']:.3f} ({d['roofline'].get('frac_at_16384_envs', 0):.3f} at 16 384 envs)"),
        ("C3 Sector 65 536 × 50 total, MVP (strong)", c3, lambda d: f"{d['c3']['roofline']['frac']:.3f}"),
        ("C5 giant 16 × 10 000, peer mode (strong)", c5, lambda d: f"{d['c5']['roofline']['frac']:.2f} (ordered entries; every unordered pair once)")]
for name, fn, fr in rows:
    print(f"| {name} | {fn(one)} | {fr(one)} | " + " | ".join(fn(L[n]) if n in L else "—" for n in (2, 4, 8)) + " |")
print("scaling vs N = 1 (device-timed value):")
for key, get in (("c2", lambda d: d['value']), ("c2 e2e", lambda d: d['e2e']['value']), ("c2 e2e 2-in-flight", lambda d: d['e2e'].get('two_steps_in_flight', {}).get('value', 0)),
                 ("c3", lambda d: d['c3']['value']), ("c3 e2e", lambda d: d['c3']['e2e']['value']), ("c5", lambda d: d['c5']['value'])):
    base = get(one)
    print(f"  {key:20s} " + "  ".join(f"N={n}: {get(L[n]) / base:.2f}x (eff {get(L[n]) / base / n:.2f})" for n in sorted(L) if n > 1))
if 'v' in one:
    print("v", M(one['v']['value']), one['v']['ms_per_step'], "c1", one.get('c1', {}).get('value'), "detect", one.get('numpy_detect', {}).get('by_aircraft'))
    print("cpu", one.get('cpu_baseline', {}).get('value'), one.get('cpu_baseline', {}).get('cores'), one.get('cpu_baseline_c_openmp', {}).get('value'))
