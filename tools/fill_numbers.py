"""Fill the @PLACEHOLDER@ numbers of DESIGN.md / BASELINE.md / README.md from the final measurement files of a round:

    python tools/fill_numbers.py r02n

reads profiles/<tag>_bench_n{1,2,4,8}.json, profiles/<tag>_phases_n1.txt, profiles/<tag>_config_table.txt."""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


def main():
    tag = sys.argv[1]
    P = os.path.join(ROOT, "profiles")
    d1 = line(os.path.join(P, tag + "_bench_n1.json"))
    v = {}
    m = lambda x: "%.1f" % (x / 1e6)
    n1, n1ms = d1["value"], d1["ms_per_step"]
    B = 65536
    v["N1V"], v["N1MS"] = m(n1), "%.3f" % n1ms
    v["N1E"], v["N1EMS"] = m(d1["e2e"]["value"]), "%.3f" % (B / d1["e2e"]["value"] * 1e3)
    sus = d1.get("sustained") or {}
    v["N1S"], v["N1SMS"] = m(sus.get("value", 0)), "%.3f" % sus.get("ms_per_step", 0)
    uni = d1.get("uniform_ids") or {}
    v["UNIV"], v["UNIMS"] = m(uni.get("value", 0)), "%.3f" % uni.get("ms_per_step", 0)
    v["K1UF"] = "%.2f" % uni.get("k1_frac", 0)
    v["K1UGBS"] = "%.0f" % uni.get("k1_achieved_gbs", 0)
    f32, tfd = d1.get("fp32") or {}, d1.get("tf_dense") or {}
    v["FP32"], v["FP32MS"] = "%.2f" % (f32.get("value", 0) / 1e6), "%.2f" % f32.get("ms_per_step", 0)
    v["TFD"], v["TFDMS"] = "%.2f" % (tfd.get("value", 0) / 1e6), "%.2f" % tfd.get("ms_per_step", 0)
    cb, cl = d1["cpu_baseline"], d1.get("cpu_baseline_lazy") or {}
    v["CPUD"], v["CPUL"] = "%.1f" % (cb["value"] / 1e3), "%.1f" % (cl.get("value", 0) / 1e3)
    v["CPUDMS"], v["CPULMS"] = "%.0f" % (4096 / cb["value"] * 1e3), "%.0f" % (4096 / max(cl.get("value", 1), 1) * 1e3)
    v["RATIOL"] = "%.0f" % (n1 / max(cl.get("value", 1), 1))
    v["RATIOE"] = "%.0f" % (d1["e2e"]["value"] / max(cl.get("value", 1), 1))
    v["RATIOD"] = "%.0f" % (tfd.get("value", 0) / cb["value"])
    r = d1["roofline"]
    k1ms = d1["kernels"]["ctr_fm_gather_fwd"]["ms_per_call"]
    v["K1US"], v["K1GB"], v["K1GBS"], v["K1F"] = "%.0f" % (k1ms * 1e3), "%.2f" % (r["achieved"] / 1e3), "%.0f" % r["achieved"], "%.2f" % r["frac"]
    mlp = d1["mlp"]
    v["MLPMS"], v["MLPTF"], v["MLPF"] = "%.3f" % mlp["ms_per_step"], "%.0f" % mlp["achieved_tflops"], "%.2f" % mlp["frac"]
    for N in (2, 4, 8):
        p = os.path.join(P, "%s_bench_n%d.json" % (tag, N))
        if os.path.exists(p):
            d = line(p)
            v["N%dV" % N], v["N%dMS" % N] = m(d["value"]), "%.3f" % d["ms_per_step"]
            v["N%dX" % N], v["N%dE" % N] = "%.2f" % (d["value"] / n1), "%.2f" % (d["value"] / n1 / N)
            v["N%dEV" % N] = m(d["e2e"]["value"])
    ph = os.path.join(P, tag + "_phases_n1.txt")
    if os.path.exists(ph):
        names = {"load -> embed": "embed", "embed -> mlp_fwd": "MLP fwd", "mlp_fwd -> head": "head", "head -> mlp_bwd": "MLP bwd",
                 "wait_sort -> partials": "partials", "partials -> sparse_update": "row Adam", "dense_reduce -> update": "dense update"}
        out = []
        for ln in open(ph):
            mm = re.match(r"(.+?)\s+([0-9.]+) us", ln.strip())
            if mm and mm.group(1).strip() in names:
                out.append("%s %.0f" % (names[mm.group(1).strip()], float(mm.group(2))))
        v["PHASES"] = ", ".join(out) + " µs"
    ct = os.path.join(P, tag + "_config_table.txt")
    if os.path.exists(ct):
        v["CONFIGTABLE"] = "```\n" + "".join(l for l in open(ct) if l.startswith("c")).rstrip() + "\n```"
    for f in ("DESIGN.md", "BASELINE.md", "README.md"):
        path = os.path.join(ROOT, f)
        s = open(path).read()
        for k, val in v.items():
            s = s.replace("@%s@" % k, val)
        left = sorted(set(re.findall(r"@[A-Z0-9]+@", s)))
        if left:
            print(f, "unfilled:", left)
        open(path, "w").write(s)


if __name__ == "__main__":
    main()
