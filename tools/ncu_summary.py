"""Reduce an `ncu --page raw --csv` dump of one training iteration to the two files kept under profiles/:
  <tag>_ncu_full_summary.csv  one line per launch with the columns the roofline discussion uses
  r2_traffic_<workload>.json  DRAM bytes per image summed per stage + per step, tagged with the build id of the CUDA
                              sources (bench.py's roofline.traffic reads it only when the tag matches the build that runs)

  ncu -i gpurun_out/<tag>_full.ncu-rep --page raw --csv > gpurun_out/<tag>_raw.csv
  python tools/ncu_summary.py gpurun_out/<tag>_raw.csv <tag> <workload>
"""
import csv
import json
import os
import re
import sys

COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct"]

# stage of the bench's stages_ms each kernel belongs to
STAGES = [("project_fwd", r"k_project_fwd"), ("tile_map", r"k_live_compact|k_os_|k_super_emit|k_seg_tables|k_expand|k_chunk_prefix"),
          ("raster_bwd", r"k_raster_bwd"), ("raster_fwd", r"k_raster_fwd|k_raster_records"), ("tile_order", r"k_tile_order"),
          ("loss", r"k_loss_l1_mse|k_ssim_"),
          ("reg", r"k_reg_"), ("project_bwd", r"k_project_bwd"), ("stats", r"k_accumulate_stats|k_pack_|k_fold_"),
          ("optim", r"k_optim_")]


def short(name):
  name = re.sub(r"<unnamed>::", "", name)
  return re.sub(r"\(.*$", "", name).strip()


def main(raw, tag, workload="c2"):
  sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
  import bench
  rows = list(csv.reader(open(raw)))
  head, units, body = rows[0], rows[1], rows[2:]
  ix = [head.index(c) for c in COLS]
  root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles")
  with open(os.path.join(root, f"{tag}_ncu_full_summary.csv"), "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(COLS)
    w.writerow([units[i] for i in ix])
    for r in body:
      w.writerow([short(r[ix[0]])] + [r[i] for i in ix[1:]])
  scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
  rd, wr, du = (head.index(c) for c in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
  traffic, time_us = {}, {}
  for r in body:
    stage = next((s for s, pat in STAGES if re.search(pat, r[ix[0]])), "other")
    b = float(r[rd]) * scale[units[rd]] + float(r[wr]) * scale[units[wr]]
    traffic[stage] = traffic.get(stage, 0) + int(round(b))
    time_us[stage] = round(time_us.get(stage, 0) + float(r[du]), 3)
  out = {"source": f"profiles/{tag}_ncu_full_summary.csv (ncu --set full, one training iteration of {workload}, per launch)",
         "build_id": bench.build_id(), "workload": workload, "launches": len(body), "dram_bytes_per_image": traffic,
         "dram_bytes_per_step": sum(traffic.values()), "serialised_us": time_us}
  json.dump(out, open(os.path.join(root, f"r2_traffic_{workload}.json"), "w"), indent=1)
  print(json.dumps(out, indent=1))


if __name__ == "__main__":
  main(*sys.argv[1:4])
