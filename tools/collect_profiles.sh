#!/bin/bash
# One gpurun call that produces everything profiles/ is built from (run from the repo root ON THE GPU BOX):
#   gpurun --timeout 1500 -- 'tools/collect_profiles.sh r01_final'
# then, back in the container:  tools/collect_profiles.sh --import r01_final
set -u
if [ "${1:-}" = "--import" ]; then
  tag=$2
  cp gpurun_out/${tag}_bench.json profiles/${tag}_bench.json
  cp gpurun_out/${tag}_bench_reference_arm.json profiles/${tag}_bench_reference_arm.json
  cp gpurun_out/${tag}_sweep_c3_c5.jsonl profiles/${tag}_sweep_c3_c5.jsonl
  cp gpurun_out/ref_gpu_path.json profiles/${tag}_ref_gpu_pytorch_path.json
  cp gpurun_out/${tag}_roi_trace.txt gpurun_out/${tag}_head_trace.txt gpurun_out/${tag}_gemm_trace.txt profiles/ 2>/dev/null
  python tools/ncu_summary.py launches gpurun_out/${tag}_launches.csv > profiles/${tag}_launches.txt
  python tools/ncu_summary.py report gpurun_out/${tag}_full.ncu-rep > profiles/${tag}_ncu_full.txt
  python - "$tag" <<'PY'
import csv, io, json, subprocess, sys
tag = sys.argv[1]
out = subprocess.run(["ncu", "-i", "gpurun_out/%s_full.ncu-rep" % tag, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[0]
traffic = {}
for r in rows[2:]:
    d = dict(zip(h, r))
    name = d["Kernel Name"]
    key = ("roi_align" if "roi_align_kernel" in name else "fc1" if "gemm_pair_kernel" in name or "gemm_bf16_kernel<256, 2>" in name
           else "fc2" if "gemm_bf16_kernel" in name else "head_decode" if "head_decode" in name else None)
    if key and key not in traffic:
        traffic[key] = (float(d["dram__bytes_read.sum"].replace(",", "")) + float(d["dram__bytes_write.sum"].replace(",", ""))) * 1e6
json.dump(dict(traffic, unit="bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, one ncu --set full capture, cold L2)",
               source="profiles/%s_ncu_full.txt" % tag), open("profiles/traffic.json", "w"), indent=1)
print(traffic)
PY
  exit 0
fi
tag=${1:-run}
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $o/${tag}_pytest_gpu.log
python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err
python bench.py --impl reference > $o/${tag}_bench_reference_arm.json 2> $o/${tag}_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"roi_align_kernel|gemm_bf16|gemm_pair|head_decode|roi_order" --launch-skip 20 --launch-count 5 -o $o/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_f.log 2>&1
python tools/sweep.py > $o/${tag}_sweep_c3_c5.jsonl 2> $o/${tag}_sweep.err
python tools/roi_trace.py > $o/${tag}_roi_trace.txt 2>&1
python tools/head_trace.py > $o/${tag}_head_trace.txt 2>&1
python tools/gemm_trace.py > $o/${tag}_gemm_trace.txt 2>&1
python __graft_entry__.py smoke 2>&1 | tail -1
cat $o/${tag}_pytest_gpu.log
