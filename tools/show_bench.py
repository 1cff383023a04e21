#!/usr/bin/env python
"""Pretty-print the interesting parts of a bench.py JSON line."""
import json
import sys

j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(f"{j['metric']}: {j['value']:.0f} {j['unit']}  {j['ms_per_step']:.4f} ms/step  n_gpus={j['n_gpus']}  launches={j['gpu_launches']}")
r = j["roofline"]
print(f"roofline: {r['kernel']} achieved {r['achieved']:.0f} / {r['peak']:.0f} = {r['frac']:.3f} (of 8 TB/s: {r['frac_of_spec_8TBs']:.3f}) traffic={r.get('traffic')}")
for k, v in r.get("per_shape", {}).items():
    ind = f"   | independent groups: {v['independent_us_per_matrix']:6.2f} us {v['independent_GBps']:6.0f} GB/s frac {v['independent_frac']:.3f}" if "independent_frac" in v else ""
    print(f"   {k:18s} {v['us_per_matrix']:7.2f} us  {v['GBps']:6.0f} GB/s  frac {v['frac']:.3f}{ind}")
for p in r.get("prefill", []):
    print(f"   prefill {p['type']} N={p['N']:6d} K={p['K']:6d} T={p['T']:5d}: {p['TFLOPS']:6.0f} TFLOPS  {p['frac_of_measured_sustained']:.2f} of sustained  {p['frac_of_spec_2250']:.2f} of 2.25PF  {p['sm_mhz']} MHz {p['power_w_max']} W {p['reasons']}")
for m in r.get("mixmul", []):
    print(f"   mixmul tokens={m['tokens']:4d}: {m['ms'] * 1e3:7.1f} us  {m['GBps']:6.0f} GB/s  {m['TFLOPS']:6.1f} TFLOPS (experts touched {m['experts_touched']})")
if r.get("quantizer"):
    q = r["quantizer"]
    print(f"   quantizer {q['rows']}x{q['k']}: {q['ms'] * 1e3:.2f} us {q['GBps']:.0f} GB/s frac {q['frac']:.2f}")
e = j.get("e2e")
if e:
    print(f"e2e: {e['value']:.0f} {e['unit']} {e['ms_per_step']:.3f} ms/step  h2d {e['h2d_bytes_per_step']} d2h {e['d2h_bytes_per_step']}  {e.get('kernel_launches_per_step')} launches, {e.get('nodes_in_decode_plan')} plan nodes; {e['api'][:90]}")
    wn = e.get("with_neighbours")
    if wn:
        print("   graph with the matmuls' neighbours:", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in wn.items() if k != "graph"})
    s = e.get("llamafile_sgemm_shim")
    if s:
        print(f"   llamafile_sgemm shim: {s['value']:.0f} GB/s {s['ms_per_step']:.3f} ms/step")
c = j.get("cpu_baseline")
if c:
    print(f"cpu_baseline: {c['value']} {c['unit']} cores={c['cores']} kind={c['kind']}  {c['sample'][:110]}")
print("clocks:", {k: v for k, v in (j.get("clocks") or {}).items() if k != "whole_run"}, " parity_checked:", j.get("parity_checked"))
