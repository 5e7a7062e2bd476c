"""On the GPU box: ncu launch list of the bench command, then one `--set full` capture of exactly one training iteration
(from the last k_project_fwd of the list to the launch before the next one would start).
  python tools/profile_iteration.py <tag>      -> gpurun_out/<tag>_launches.csv, gpurun_out/<tag>_full.ncu-rep"""
import csv
import subprocess
import sys

tag = sys.argv[1]
cmd = ["python", "bench.py", "--steps", "2", "--warmup", "3", "--no-cpu-baseline"]
subprocess.run(["ncu", "--metrics", "gpu__time_duration.sum", "--clock-control", "none", "-c", "600", "--csv", "--log-file",
                f"gpurun_out/{tag}_launches.csv"] + cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
names = []
for r in csv.reader(open(f"gpurun_out/{tag}_launches.csv")):
  if len(r) > 10 and r[0].isdigit():
    names.append(r[4])
starts = [i for i, n in enumerate(names) if "k_project_fwd" in n]
# the iteration before the last complete one (the list may end mid-iteration)
a, b = starts[-3], starts[-2]
print(f"{len(names)} launches, iteration = launches {a}..{b - 1} ({b - a} launches)")
subprocess.run(["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on", "--launch-skip", str(a),
                "--launch-count", str(b - a), "-f", "-o", f"gpurun_out/{tag}_full"] + cmd,
               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
