"""On the GPU box: one `ncu --set full --import-source on` capture of the named kernels of one bench iteration and its
per-line source page as csv (the .ncu-rep itself is removed: gpurun_out is size-limited).
  python tools/profile_source.py <tag> <workload> <kernel regex>  -> gpurun_out/<tag>_source_<i>.csv, <tag>_raw.csv"""
import subprocess
import sys

tag, workload, regex = sys.argv[1], sys.argv[2], sys.argv[3]
cmd = ["python", "bench.py", "--workload", workload, "--steps", "2", "--warmup", "3", "--windows", "1", "--no-cpu-baseline"]
rep = f"gpurun_out/{tag}_src.ncu-rep"
# skip the warm-up iterations' launches of these kernels (calibration steps run on other list sizes)
skip = int(sys.argv[4]) if len(sys.argv) > 4 else 8
count = int(sys.argv[5]) if len(sys.argv) > 5 else 4
subprocess.run(["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on", "-k", f"regex:{regex}",
                "--launch-skip", str(skip), "--launch-count", str(count), "-f", "-o", rep[:-8]] + cmd,
               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
with open(f"gpurun_out/{tag}_raw.csv", "w") as f:
  subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=False)
with open(f"gpurun_out/{tag}_source.csv", "w") as f:
  subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], stdout=f,
                 stderr=subprocess.DEVNULL, check=False)
subprocess.run(["rm", "-f", rep])
