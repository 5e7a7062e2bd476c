"""On the GPU box: ncu launch list of the bench command, then one `--set full` capture of exactly one training iteration
(from a k_project_fwd of the list to the launch before the next one).
  python tools/profile_iteration.py <tag> [workload]  -> gpurun_out/<tag>_launches.csv, gpurun_out/<tag>_full.ncu-rep"""
import csv
import subprocess
import sys

tag = sys.argv[1]
workload = sys.argv[2] if len(sys.argv) > 2 else "c2"
cmd = ["python", "bench.py", "--workload", workload, "--steps", "2", "--warmup", "3", "--windows", "1", "--no-cpu-baseline"]
subprocess.run(["ncu", "--metrics", "gpu__time_duration.sum", "--clock-control", "none", "-c", "1500", "--csv", "--log-file",
                f"gpurun_out/{tag}_launches.csv"] + cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
names = []
for r in csv.reader(open(f"gpurun_out/{tag}_launches.csv")):
  if len(r) > 10 and r[0].isdigit():
    names.append(r[4])
starts = [i for i, n in enumerate(names) if "k_project_fwd" in n]
# an iteration of the timed window: the warm-up ones are followed by host-side overflow reads, so take the 3rd from the end
# of the value window (the list may end mid-iteration)
a, b = starts[-3], starts[-2]
print(f"{len(names)} launches, iteration = launches {a}..{b - 1} ({b - a} launches)")
subprocess.run(["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on", "--launch-skip", str(a),
                "--launch-count", str(b - a), "-f", "-o", f"gpurun_out/{tag}_full"] + cmd,
               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
