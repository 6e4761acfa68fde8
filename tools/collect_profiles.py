"""Copy one evidence run (tools/evidence.sh TAG, brought back in gpurun_out/) into profiles/ under the round's names
and print the digests profiles/rNN_summary.md quotes.

  python tools/collect_profiles.py TAG [ROUND]      e.g.  s28 r01
"""
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")


def raw_csv(rep, dst):
    with open(dst, "w") as fh:
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=fh, stderr=subprocess.DEVNULL, check=True)


def main():
    tag = sys.argv[1]
    rnd = sys.argv[2] if len(sys.argv) > 2 else "r01"
    g = lambda name: os.path.join(OUT, "%s_%s" % (tag, name))  # noqa: E731
    p = lambda name: os.path.join(PROF, "%s_%s" % (rnd, name))  # noqa: E731
    shutil.copy(g("bench.json"), p("bench_n1.json"))
    shutil.copy(g("ref.json"), p("bench_reference_arm.json"))
    shutil.copy(g("launches.csv"), p("launches_bench.csv"))
    if os.path.exists(g("launches_ukf_large.csv")):
        shutil.copy(g("launches_ukf_large.csv"), p("launches_ukf_large_step.csv"))
    for src, dst in (("large", "ncu_large_update.csv"), ("batched", "ncu_ekf_batched.csv"), ("laser", "ncu_laser_extract.csv"),
                     ("solve", "ncu_large_solve.csv"), ("dgemm", "ncu_dgemm_nt.csv"), ("chol", "ncu_chol_panel.csv"),
                     ("tf32", "ncu_tf32_gemm.csv"), ("ukfb", "ncu_ukf_batched.csv")):
        if os.path.exists(g(src + ".raw.csv")):
            shutil.copy(g(src + ".raw.csv"), p(dst))
        elif os.path.exists(g(src + ".ncu-rep")):
            raw_csv(g(src + ".ncu-rep"), p(dst))
        if os.path.exists(p(dst)) and (os.path.exists(g(src + ".raw.csv")) or os.path.exists(g(src + ".ncu-rep"))):
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_digest.py"), p(dst)], check=True)
    # dram traffic of the cfg4 streaming update (bench.py reports it as roofline.traffic)
    import csv
    import statistics
    rows = list(csv.reader(open(p("ncu_large_update.csv"))))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    units = rows[1]
    data = [r for r in rows[2:] if "large_update" in r[col["Kernel Name"]]]

    def med(key, scale_unit):
        mult = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3}[units[col[key]]]
        return statistics.median(float(r[col[key]].replace(",", "")) for r in data) * mult

    t = {"source": "profiles/%s_ncu_large_update.csv (ncu --set full, tools/prof_target.py large 64 8, evidence run %s)" % (rnd, tag),
         "kernel": data[0][col["Kernel Name"]].split("(")[0].split("::")[-1],
         "dram_bytes_read": int(med("dram__bytes_read.sum", "byte")),
         "dram_bytes_write": int(med("dram__bytes_write.sum", "byte")),
         "launch_us": round(med("gpu__time_duration.sum", "us"), 2)}
    with open(p("traffic.json"), "w") as fh:
        json.dump(t, fh, indent=1)
    print(json.dumps(t))
    # FP64-pipe utilisation of the cfg3 batched kernel (bench.py reports it beside the flop-convention fraction)
    if os.path.exists(p("ncu_ekf_batched.csv")):
        rows = list(csv.reader(open(p("ncu_ekf_batched.csv"))))
        col = {h: i for i, h in enumerate(rows[0])}
        key = "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
        data = [r for r in rows[2:] if "ekf_batched" in r[col["Kernel Name"]]]
        if data and key in col:
            b = {"source": "profiles/%s_ncu_ekf_batched.csv (ncu --set full, evidence run %s)" % (rnd, tag),
                 "kernel": data[0][col["Kernel Name"]].split("(")[0].split("::")[-1],
                 "pipe_fp64_active_pct": statistics.median(float(r[col[key]].replace(",", "")) for r in data)}
            with open(p("batched_pipe.json"), "w") as fh:
                json.dump(b, fh, indent=1)
            print(json.dumps(b))
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "launch_summary.py"), p("launches_bench.csv")], check=True)


if __name__ == "__main__":
    main()
