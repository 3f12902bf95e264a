#!/bin/bash
# One gpurun call that produces everything profiles/ is built from (run from the repo root ON THE GPU BOX):
#   gpurun --timeout 1800 -- 'tools/collect_profiles.sh r02_final'
# then, back in the container:  tools/collect_profiles.sh --import r02_final
set -u
if [ "${1:-}" = "--import" ]; then
  tag=$2
  for f in bench.json bench_reference_arm.json sweep_configs.jsonl roi_trace.txt tail_trace.txt gemm_trace_pipeline.txt \
           cublas_bar.json bench_channels_last.json sweep_configs_channels_last.jsonl roi_trace_channels_last.txt; do
    [ -f gpurun_out/${tag}_$f ] && cp gpurun_out/${tag}_$f profiles/${tag}_$f
  done
  [ -f gpurun_out/ref_gpu_path.json ] && cp gpurun_out/ref_gpu_path.json profiles/${tag}_ref_gpu_pytorch_path.json
  for f in perf_gdino_post perf_gdino_flow perf_box_head; do
    [ -f gpurun_out/$f.json ] && cp gpurun_out/$f.json profiles/${tag}_$f.json
  done
  python tools/ncu_summary.py launches gpurun_out/${tag}_launches.csv > profiles/${tag}_launches.txt
  python tools/ncu_summary.py report gpurun_out/${tag}_full.ncu-rep > profiles/${tag}_ncu_full.txt
  [ -f gpurun_out/${tag}_full_channels_last.ncu-rep ] && python tools/ncu_summary.py report gpurun_out/${tag}_full_channels_last.ncu-rep > profiles/${tag}_ncu_full_channels_last.txt
  python tools/sass_summary.py > profiles/sass_summary.txt
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
           else "fc_tail" if "fc_tail_kernel" in name else "roi_plan" if "roi_plan_kernel" in name else None)
    if key and key not in traffic:
        traffic[key] = (float(d["dram__bytes_read.sum"].replace(",", "")) + float(d["dram__bytes_write.sum"].replace(",", ""))) * 1e6
try:
    out = subprocess.run(["ncu", "-i", "gpurun_out/%s_full_channels_last.ncu-rep" % tag, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    for r in rows[2:]:
        d = dict(zip(rows[0], r))
        if "roi_align_kernel" in d["Kernel Name"]:
            traffic["roi_align_channels_last"] = (float(d["dram__bytes_read.sum"].replace(",", "")) + float(d["dram__bytes_write.sum"].replace(",", ""))) * 1e6
            break
except Exception as ex:
    print("no channels-last capture:", ex)
json.dump(dict(traffic, unit="bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, one ncu --set full capture, cold L2)",
               source="profiles/%s_ncu_full.txt" % tag), open("profiles/traffic.json", "w"), indent=1)
print(traffic)
PY
  exit 0
fi
tag=${1:-run}
o=gpurun_out
mkdir -p $o
python -m pytest tests -m gpu -q 2>&1 | tail -3 > $o/${tag}_pytest_gpu.log
python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err
python bench.py --impl reference > $o/${tag}_bench_reference_arm.json 2> $o/${tag}_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"roi_align_kernel|gemm_pair|fc_tail|roi_plan" --launch-skip 16 --launch-count 4 -o $o/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_f.log 2>&1
: > $o/${tag}_sweep_configs.jsonl
for c in C1 C3-1x2000 C3-8x300 C3-8x2000 C4 C5-sfp4 C5-dla C5-c128 C5-c512; do
  python bench.py --config $c --steps 20 --no-cpu-baseline >> $o/${tag}_sweep_configs.jsonl 2>> $o/${tag}_sweep.err
done
python bench.py --layout channels_last --no-cpu-baseline > $o/${tag}_bench_channels_last.json 2>> $o/${tag}_bench.err
ncu --set full --clock-control none --import-source on -k regex:"roi_align_kernel" --launch-skip 4 --launch-count 1 -o $o/${tag}_full_channels_last -f python bench.py --layout channels_last --steps 2 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_fc.log 2>&1
: > $o/${tag}_sweep_configs_channels_last.jsonl
for c in C3-1x2000 C3-8x2000 C5-sfp4 C5-dla C5-c128 C5-c512; do
  python bench.py --config $c --layout channels_last --steps 20 --no-cpu-baseline >> $o/${tag}_sweep_configs_channels_last.jsonl 2>> $o/${tag}_sweep.err
done
python tools/roi_trace.py > $o/${tag}_roi_trace.txt 2>&1
OVM_TRACE_LAYOUT=channels_last python tools/roi_trace.py > $o/${tag}_roi_trace_channels_last.txt 2>&1
python tools/head_trace.py > $o/${tag}_tail_trace.txt 2>&1
python tools/gemm_trace.py pipeline > $o/${tag}_gemm_trace_pipeline.txt 2>&1
python tools/cublas_bar.py --out $o/${tag}_cublas_bar.json > /dev/null 2>&1
python __graft_entry__.py smoke 2>&1 | tail -1
cat $o/${tag}_pytest_gpu.log
