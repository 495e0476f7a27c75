#!/usr/bin/env python
"""print the figures of bench.py JSON lines that matter when comparing runs: tools/show_bench.py FILE.json ..."""
import json, sys
def show(tag, x):
    r = x.get("roofline") or {}
    ht = x.get("halo_traffic") or {}
    c = x.get("clocks") or {}
    print("  %-16s %8.1f Gpt/s  ms/step %.4f  kernel_ms %.4f  frac %.3f  step-sweep %s us  sm %s %s" % (
        tag, x["value"] / 1e9, x["ms_per_step"], r.get("kernel_ms", 0), r.get("frac", 0),
        ("%.1f" % ht["step_minus_sweep_us"]) if ht.get("step_minus_sweep_us") is not None else "-",
        c.get("sm_mhz"), ",".join(c.get("reasons", []))))
for f in sys.argv[1:]:
    lines = [l for l in open(f).read().splitlines() if l.startswith("{")]
    if not lines:
        print(f, "no JSON line"); continue
    d = json.loads(lines[-1])
    print(f, "n_gpus", d.get("n_gpus"), d.get("scaling"))
    show("headline", d)
    if d.get("one_step"): show("one_step", d["one_step"])
    s = d.get("sustained")
    if s: print("  sustained        %8.1f Gpt/s  kernel_ms %.4f frac %.3f sm %s" % (s["value"] / 1e9, s["kernel_ms"], s["roofline_frac"], s["clocks"].get("sm_mhz")))
    for k, v in (d.get("strong") or {}).items():
        if isinstance(v, dict) and "roofline" in v: show("strong " + k, v)
    c3 = d.get("config3") or {}
    for k, v in c3.items():
        if isinstance(v, dict) and "roofline" in v: show("config3 " + k, v)
    e = d.get("e2e")
    if e and e.get("value"): print("  e2e %.3f Gpt/s frac_of_pcie %s" % (e["value"] / 1e9, e.get("frac_of_pcie")))
    p = d.get("parity")
    if p: print("  parity cases %s checks %s mismatches %s" % (p.get("cases"), p.get("checks"), p.get("mismatches")))
