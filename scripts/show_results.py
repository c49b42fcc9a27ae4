"""Prints the essentials of gpurun_out bench / multi-GPU check logs."""
import json, sys, glob
for f in sorted(glob.glob("gpurun_out/mg_check*.log")):
    L = [l for l in open(f) if l.startswith("{")]
    if not L:
        print(f, "NO JSON:", open(f).read()[-600:]); continue
    d = json.loads(L[0])
    print(d["shape"], "world", d["world"], "ok", d["ok"], "ms_1gpu %.2f pcg %.2f" % (d["ms_1gpu"], d["ms_pcg_1gpu"]), d["pcg_1gpu"])
    for m in d["modes"]:
        print("  %-14s ok=%s rel=%.1e ms=%s pcg_ms=%s it=%s own/peer=%s" % (m["mode"], m["ok"], m["max_rel_cost_diff_vs_1gpu"],
              ["%.2f" % x for x in m["ms_sharded"]], ["%.2f" % x for x in m["ms_pcg"]], m["pcg_sharded"], [(i["n_own"], i["peer_exchange"]) for i in m["info"]]))
for f in sorted(glob.glob("gpurun_out/bench[0-9].log")):
    L = [l for l in open(f) if l.startswith("{")]
    if not L:
        print(f, "NO JSON:", open(f).read()[-600:]); continue
    d = json.loads(L[0])
    k = d["roofline"]["kernels"]
    print(f, "n=%d value=%.2f it/s ms/step=%.3f e2e=%.2f pcg/step=%.1f final=%.10g" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["pcg_iterations_per_step"], d["final_cost"]))
    print("   phase/step:", {a: round(b / d["steps"], 3) for a, b in d["phase_ms"].items()}, "launches", d["gpu_launches"], d.get("pcg_allreduce"))
    print("   kernels:", {a: (round(b["ms"] * 1e3, 1), round(b["frac"], 3)) for a, b in k.items()})
