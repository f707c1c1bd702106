"""The round's ncu evidence in one go (run on the GPU box, one GPU):
  1. `tools/prof_step.py 3` alone (must exit 0),
  2. the launch list of the same command (`--metrics gpu__time_duration.sum --clock-control none`),
  3. `ncu --set full` of the LAST of the three steps (skip / count taken from the launch list: a
     step starts at its tree_sample_kernel), report kept in /tmp (too large to bring back),
  4. tools/ncu_summary.py and tools/ncu_traffic.py of that report.
python tools/final_capture.py <outdir> [commit]"""
import csv, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = sys.argv[1]
commit = sys.argv[2] if len(sys.argv) > 2 else ""
os.makedirs(out, exist_ok=True)
py = sys.executable
step = [py, os.path.join(ROOT, "tools", "prof_step.py"), "3"]


def run(cmd, log):
    with open(os.path.join(out, log), "w") as f:
        return subprocess.run(cmd, stdout=f, stderr=subprocess.STDOUT, cwd=ROOT).returncode


assert run(step, "plain.log") == 0, "prof_step failed without ncu"
run(["ncu", "--metrics", "gpu__time_duration.sum", "--clock-control", "none", "-c", "2000", "--csv",
     "--log-file", os.path.join(out, "launches.csv")] + step, "ncu1.log")
names = []
with open(os.path.join(out, "launches.csv")) as f:
    for row in csv.DictReader(l for l in f if not l.startswith("==")):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            names.append(row["Kernel Name"])
starts = [i for i, n in enumerate(names) if "tree_sample" in n]
skip, count = starts[-1], len(names) - starts[-1]
print(f"{len(names)} launches, last step = launches {skip}..{len(names) - 1} ({count})")
assert run(step, "plain2.log") == 0
rc = run(["ncu", "--set", "full", "--clock-control", "none", "-s", str(skip), "-c", str(count), "-f", "-o", "/tmp/step"] + step,
         "ncu2.log")
print("ncu --set full rc", rc)
with open(os.path.join(out, "ncu_full_step.txt"), "w") as f:
    subprocess.run([py, os.path.join(ROOT, "tools", "ncu_summary.py"), "/tmp/step.ncu-rep"], stdout=f, stderr=subprocess.STDOUT)
with open(os.path.join(out, "traffic.err"), "w") as f:
    subprocess.run([py, os.path.join(ROOT, "tools", "ncu_traffic.py"), "/tmp/step.ncu-rep",
                    os.path.join(out, "ncu_traffic.json"), commit], stdout=subprocess.DEVNULL, stderr=f)
