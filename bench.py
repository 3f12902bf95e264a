"""bench.py -- lifted 3-D boxes / second of the 2D->3D lift (ROIAlign + cube head + decode) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1], SURVEY.md section 8d "C2"): OVMono3D_dinov2_SFP lift, one batch of 16 images x
100 boxes per rank and step, DINOv2 simple-feature-pyramid [16,256,128,128] / [16,256,64,64] / [16,256,32,32] fp32
NCHW (896-px canvas, strides 7/14/28), synthetic K, class-agnostic head 12544 -> 1024 -> 1024 -> 13 with bf16 GEMM
operands and fp32 accumulation.  Seeded synthetic inputs, random-init weights (ovmono3d_b200/synth.py).

A step is one pass of the path over one batch.  `value` times `ovm_lift_forward` with every input already in HBM
(per-step CUDA events on the launching stream; L2 is flushed between steps outside the event pairs and the steps
rotate over distinct input batches).  `e2e` times the plugin call `ROIHeads3DB200._forward_cube` with HOST inputs:
every step copies that step's feature pyramid, boxes, classes, scores and cameras from pinned host memory and reads
the detection records back to the host.  `e2e_resident_features` is the deployment case -- the backbone leaves the
pyramid in HBM, as in the reference (rcnn3d.py:97-109) -- with only boxes/classes/scores/cameras coming from the host.

`--impl reference` times the reference's own CPU path for the same workload: the reference is pure Python on
detectron2/pytorch3d, which cannot be installed here, so the arm runs `oracle/` -- the op-for-op
CPU restatement pinned to vectors produced by the reference's own code (tests/golden) -- on all host cores.

Multi-GPU: images shard across ranks (one 16-image batch per rank and step, no data-path collective); NCCL is used
only to all-gather the detection records once after the last step, as the reference gathers its detections
(omni3d_evaluation.py:717-720); that gather is inside the timed total.  `scaling` is "weak".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from typing import Dict, List, Optional

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "lifted_3d_boxes_per_sec"
UNIT = "boxes/s"
NUM_IMAGES = 16
BOXES_PER_IMAGE = 100
STRIDES = (7, 14, 28)
CANVAS = 896
CHANNELS = 256
FC_DIM = 1024
POOL = 7
L2_FLUSH_BYTES = 256 << 20   # > 126 MB L2
WORKLOAD = ("C2: OVMono3D_dinov2_SFP lift, 16 images x 100 boxes per rank-step, SFP p2-p4 256ch "
            "@128/64/32 (896px, strides 7/14/28) fp32 NCHW, head 12544-1024-1024-13 bf16 operands / fp32 accumulate")
_C2_WORKLOAD = WORKLOAD
CONFIG = "C2"
SEED = 1235
LAYOUT = "nchw"     # --layout: memory format of the feature pyramid ("nchw" = what the reference's backbone emits)
STREAM = False      # C4: the steps are consecutive batches of one stream (no L2 flush between them, one event pair)


def set_workload(name: str) -> None:
    """Select one of BASELINE.json's configs (ovmono3d_b200/workloads.py) for every arm of this file.  The default, C2,
    is the configuration the metric is quoted on; C4 is the C2 batch streamed 20 000 images deep."""
    global NUM_IMAGES, BOXES_PER_IMAGE, STRIDES, CANVAS, CHANNELS, FC_DIM, WORKLOAD, CONFIG, SEED, STREAM
    from ovmono3d_b200 import workloads

    STREAM = name == "C4"
    w = workloads.get("C2" if STREAM else name)
    NUM_IMAGES, BOXES_PER_IMAGE, STRIDES, CANVAS, CHANNELS, FC_DIM = (w.num_images, w.boxes_per_image, tuple(w.strides),
                                                                     w.canvas, w.channels, w.fc_dim)
    CONFIG, SEED = name, w.seed
    if STREAM:
        WORKLOAD = ("C4: Omni3D-test-scale synthetic stream, 20 000 images x 100 boxes in batches of 16 per rank-step "
                    "(C2 shapes), images sharded contiguously by rank, detections exchanged once at the end")
    elif name != "C2":
        WORKLOAD = w.describe()
    else:
        WORKLOAD = _C2_WORKLOAD


def config_dict(n: int) -> Dict[str, object]:
    """`config` of the JSON line: the workload only, identical in both arms (what is arm-specific goes to `method`)."""
    w = WORKLOAD if LAYOUT == "nchw" else WORKLOAD.replace("fp32 NCHW", "fp32 channels-last (torch.channels_last, NHWC)")
    return {"workload": w, "name": CONFIG, "images_per_step_per_gpu": NUM_IMAGES, "boxes_per_step_per_gpu": n,
            "layout": LAYOUT}


def _peaks() -> Dict[str, float]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows: List[List[str]] = []
        self.proc: Optional[subprocess.Popen] = None
        self.thread: Optional[threading.Thread] = None

    def start(self) -> None:
        # NVML in-process (a query is tens of microseconds); an `nvidia-smi -lms` child holds the driver lock for
        # ~20 ms per sample, which showed up as 20 ms stalls of synchronising calls inside the timed region
        try:
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(pynvml))
            bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                    ("sw_power_cap", 0x4))
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            self._stop = threading.Event()

            def poll():
                while not self._stop.is_set():
                    try:
                        sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        try:
                            r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                        except Exception:
                            r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append([str(sm), str(mx)] + ["Active" if r & b else "Not Active" for _, b in bits])
                    except Exception:
                        pass
                    self._stop.wait(0.02)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            self.nvml = True
            return
        except Exception:
            self.nvml = False
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return

        def pump():
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])

        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def _physical_index(self, pynvml) -> int:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if self.index < len(ids) and ids[self.index].isdigit():
                return int(ids[self.index])
        return self.index

    def stop(self) -> Dict[str, object]:
        if getattr(self, "nvml", False):
            self._stop.set()
            self.thread.join(timeout=2)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            if self.thread:
                self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for name, v in zip(self.NAMES, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- workload
def make_batch(seed: int):
    from ovmono3d_b200 import synth

    content = 512 if CONFIG == "C1" else None
    return synth.make_inputs(seed, NUM_IMAGES, BOXES_PER_IMAGE, strides=STRIDES, canvas=CANVAS, content=content,
                             channels=CHANNELS)


def make_weights():
    from ovmono3d_b200 import synth

    return synth.make_head_weights(77, in_dim=CHANNELS * POOL * POOL, fc_dim=FC_DIM)


def build_head(device: torch.device):
    """The plugin, built the way the reference builds its head: cfg -> ROI_HEADS_REGISTRY (roi_heads.py:31-36)."""
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.config import get_cfg
    from ovmono3d_b200.roi_heads import build_roi_heads

    names = ["p%d" % (2 + i) for i in range(len(STRIDES))]
    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = names
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False      # shipped config (configs/Base.yaml:81)
    shape = {f: ShapeSpec(channels=CHANNELS, stride=s) for f, s in zip(names, STRIDES)}
    head = build_roi_heads(c, shape)
    head.cube_head.load_state_dict(make_weights())
    return head.to(device).eval(), names


def roi_align_algorithmic_bytes(boxes: np.ndarray, levels: np.ndarray, out_elt: int) -> float:
    """per box, the fp32 feature footprint of the box at its level (clipped to the map) plus
    the pooled output C*49*elt."""
    total = 0.0
    for l, s in enumerate(STRIDES):
        m = levels == l
        if not m.any():
            continue
        b = boxes[m].astype(np.float64)
        size = CANVAS // s
        x_lo, x_hi = b[:, 0] / s - 0.5, b[:, 2] / s - 0.5
        y_lo, y_hi = b[:, 1] / s - 0.5, b[:, 3] / s - 0.5
        w = np.clip(np.floor(x_hi) - np.floor(x_lo) + 2, 1, size)
        h = np.clip(np.floor(y_hi) - np.floor(y_lo) + 2, 1, size)
        total += float((w * h).sum()) * CHANNELS * 4
    total += boxes.shape[0] * CHANNELS * POOL * POOL * out_elt
    return total


class DeviceBatch:
    """One batch with every input of the lift resident in HBM."""

    def __init__(self, inp, device, layout: Optional[str] = None):
        n_per = [len(b) for b in inp.boxes]
        self.n = sum(n_per)
        self.features = [f.to(device) for f in inp.features]
        if (layout or LAYOUT) == "channels_last":
            self.features = [f.contiguous(memory_format=torch.channels_last) for f in self.features]
        self.boxes = torch.cat(inp.boxes).to(device).contiguous()
        self.classes = torch.cat(inp.classes).to(device).contiguous()
        self.scores = torch.cat(inp.scores).to(device).contiguous()
        self.batch_idx = torch.from_numpy(np.repeat(np.arange(len(n_per), dtype=np.int32), n_per)).to(device)
        cams = torch.zeros(len(n_per), 12)
        for i in range(len(n_per)):
            cams[i, :9] = inp.Ks[i].reshape(9)
            cams[i, 9] = inp.im_scales_ratio[i]
            cams[i, 10] = inp.im_dims[i][0]
        self.cams = cams.to(device)


class HostBatch:
    """One batch in pinned host memory, shaped the way the reference's caller holds it (per-image lists)."""

    def __init__(self, inp):
        self.features = [(f.contiguous(memory_format=torch.channels_last) if LAYOUT == "channels_last" else f).pin_memory()
                         for f in inp.features]
        self.boxes = [b.pin_memory() for b in inp.boxes]
        self.classes = [c.pin_memory() for c in inp.classes]
        self.scores = [s.pin_memory() for s in inp.scores]
        self.Ks, self.ratios, self.im_dims = inp.Ks, list(inp.im_scales_ratio), [tuple(d) for d in inp.im_dims]
        self.n = sum(len(b) for b in inp.boxes)
        self.feature_bytes = sum(f.numel() * 4 for f in self.features)
        self.small_bytes = sum(b.numel() * 4 for b in self.boxes) + sum(c.numel() * 8 for c in self.classes) + \
            sum(s.numel() * 4 for s in self.scores) + len(self.Ks) * 12 * 4 + self.n * 4


# --------------------------------------------------------------------------------------------- our arm
STAGES_UNFUSED = ("roi_order", "roi_align", "fc1", "fc2", "head_decode")
STAGES_FUSED = ("roi_order", "roi_align", "fc1", "fc_tail")      # fc_tail = fc2 + heads + decode in one launch


def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def _max_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_ours(args) -> Dict[str, object]:
    import torch.distributed as dist

    from ovmono3d_b200 import ops
    from ovmono3d_b200.compat import Boxes, Instances

    rank, world, local = _dist_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the lift has no CPU path"
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N>1 with torch.distributed.run" % (args.gpus, world))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    K, W = args.steps, max(args.warmup, 3)
    ops.load_library()
    peaks = _peaks()

    head, names = build_head(device)
    num_sets = 3   # 3 x 352 MB of pyramids: the steps rotate over inputs larger than L2
    inputs = [make_batch(SEED + 1000 * rank + i) for i in range(num_sets)]
    dbs = [DeviceBatch(x, device) for x in inputs]
    n = dbs[0].n
    eng = head._get_engine()
    eng.prepare(device, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=device)
    records_all = torch.empty((K, n, ops.RECORD_WIDTH), device=device, dtype=torch.float32)

    def lift(db, out=None, peers=None):
        return eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams, out=out,
                           peer_ptrs=peers)[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident steps, one event pair per step, L2 flushed outside the pairs
    for i in range(W):
        flush.zero_()
        lift(dbs[i % num_sets])
    exchange, transport = None, "single GPU"
    if world > 1:   # the communicator and the exchange's buffers are set up before the timed region
        from ovmono3d_b200 import shard

        transport = os.environ.get("OVM_GATHER", "peer")
        try:
            exchange = shard.RecordExchange(K * n, device, transport=transport)
        except Exception as ex:       # symmetric memory not available on this box: the one-collective NCCL path
            sys.stderr.write("RecordExchange(%s) unavailable (%s); using nccl\n" % (transport, ex))
            transport = "nccl"
            exchange = shard.RecordExchange(K * n, device, transport=transport)
        records_all = exchange.local.view(K, n, ops.RECORD_WIDTH)    # the lift writes straight into the send block
        for i in range(min(W, K)):
            lift(dbs[i % num_sets], out=records_all[i], peers=exchange.peer_ptrs(i * n))
        exchange.finish(K * n)
    barrier()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(local)
    sampler.start()
    # untimed warm-up steps until the clock sampler has been running for 0.15 s: the timed loop starts on a GPU that is
    # under load (an idle 0.15 s sleep here let the clocks drop: the first timed step took 0.28 ms instead of 0.176)
    t_warm = time.perf_counter()
    i = 0
    while time.perf_counter() - t_warm < 0.15:
        flush.zero_()
        lift(dbs[i % num_sets])
        i += 1
        if i % 8 == 0:
            torch.cuda.synchronize()
    barrier()
    launches0 = ops.launch_count()
    t_wall0 = time.perf_counter()
    for i in range(K):
        if not STREAM:
            flush.zero_()
            starts[i].record()
        elif i == 0:
            starts[0].record()
        lift(dbs[i % num_sets], out=records_all[i], peers=exchange.peer_ptrs(i * n) if exchange is not None else None)
        if not STREAM:
            stops[i].record()
        elif i == K - 1:
            stops[0].record()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gathered = None
    if world > 1:
        # the L2 flushes between the steps are outside the step event pairs and let the ranks drift apart; line them
        # up again so that the gather's events time the exchange, not the wait for the slowest rank's flushes
        barrier()
    g0.record()
    if world > 1:   # the path's one exchange: detections gathered once at the end (omni3d_evaluation.py:717-720)
        gathered, counts = exchange.finish(K * n)
    g1.record()
    if world > 1:
        assert counts.cpu().tolist() == [K * n] * world and tuple(gathered.shape) == (world, K * n, ops.RECORD_WIDTH)
        assert torch.isfinite(gathered[:, :, :47]).all(), "non-finite gathered records"
    barrier()
    loop_wall_ms = (time.perf_counter() - t_wall0) * 1e3
    launches = ops.launch_count() - launches0
    if STREAM:      # one event pair around the whole stream
        loop_ms = starts[0].elapsed_time(stops[0])
        step_ms = [loop_ms / K] * K
    else:
        step_ms = [s.elapsed_time(e) for s, e in zip(starts, stops)]
    gather_ms = g0.elapsed_time(g1) if world > 1 else 0.0
    total_ms = _max_over_ranks(sum(step_ms) + gather_ms, world, device)
    time.sleep(0.15)
    clocks = sampler.stop()
    ms_per_step = total_ms / K
    value = world * n * K / (total_ms * 1e-3)
    assert torch.isfinite(records_all).all(), "non-finite detection records"

    out: Dict[str, object] = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": config_dict(n),
        "method": {"sharding": ("images sharded by rank, no data-path collective; detection records exchanged inside the "
                                "timed total: " + ("the decode stores every record into every rank's symmetric-memory "
                                                   "block over NVLink (ovm_lift_forward_peers), one signal-pad barrier after "
                                                   "the last step" if transport == "peer" else
                                                   "one NCCL all-gather of fixed-capacity blocks after the last step"))
                   if world > 1 else "single GPU",
                   "l2": ("no flush: consecutive batches of one stream, rotating over 3 input batches larger than L2"
                          if STREAM else "256 MiB memset between steps (outside the per-step event pairs) and steps rotate "
                          "over 3 input batches of %d MB each" % (sum(f.numel() * 4 for f in dbs[0].features) // 1000000)),
                   "timing": ("one CUDA-event pair around the whole stream (+ exchange), max over ranks" if STREAM else
                              "sum of per-step CUDA-event pairs on the launching stream (+ exchange), max over ranks")},
        "clocks": clocks, "gpu_launches": int(launches),
        "step_ms_min_median_max": [min(step_ms), float(np.median(step_ms)), max(step_ms)],
        "gather_ms_total": gather_ms, "gather_transport": transport, "loop_wall_ms_incl_flush": loop_wall_ms,
    }
    if rank == 0:
        out.update(kernel_breakdown(eng, dbs, flush, lift, K, peaks))
    if rank == 0 and world == 1 and LAYOUT == "nchw" and not STREAM:
        # the same steps on a torch.channels_last copy of the same pyramids (what a channels-last backbone hands over
        # without a copy): reported beside the headline, which stays on the reference's own NCHW layout
        cl = [DeviceBatch(x, device, layout="channels_last") for x in inputs]
        for i in range(W):
            flush.zero_()
            lift(cl[i % num_sets])
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for i in range(K):
            flush.zero_()
            ev[i][0].record()
            got = lift(cl[i % num_sets])
            ev[i][1].record()
        torch.cuda.synchronize()
        cl_ms = float(np.mean([a_.elapsed_time(b_) for a_, b_ in ev]))
        want = lift(dbs[(K - 1) % num_sets])
        dev_rel = float(((got - want)[:, :47].abs().max() / want[:, :47].abs().max()).item())
        kb = kernel_breakdown(eng, cl, flush, lift, K, peaks)
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        traffic = None
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f).get("roi_align_channels_last")
        out["channels_last_pyramid"] = {
            "ms_per_step": cl_ms, "value": n / (cl_ms * 1e-3), "unit": UNIT,
            "roi_align": dict(kb["kernels"]["roi_align"], traffic=traffic),
            "stages_ms": {k: v["ms"] for k, v in kb["kernels"].items()},
            "max_rel_diff_vs_nchw_records": dev_rel,
            "what": "same batches with the pyramids in torch.channels_last (no copy inside the step): roi_align_kernel<.., "
                    "channels-last> reads [row][column][64 channels] tiles; `python bench.py --layout channels_last` runs "
                    "every arm of this file on that layout"}
        del cl
    barrier()
    out.update(run_e2e(head, names, inputs, device, K, W, world, lift))
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(inputs[0], budget_s=12.0)
    if world > 1:
        dist.destroy_process_group()
    return out


def kernel_breakdown(eng, dbs, flush, lift, K: int, peaks) -> Dict[str, object]:
    """Per-stage device time inside the real step (ovm_set_stage_events) and the roofline of the dominant kernel."""
    from ovmono3d_b200 import ops

    STAGES = STAGES_FUSED if eng.tail_fused else STAGES_UNFUSED
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(len(STAGES) + 1)] for _ in range(K)]
    for row in evs:
        for e in row:
            e.record()      # creates the CUDA handle
    torch.cuda.synchronize()
    for i in range(K):
        flush.zero_()
        ops.set_stage_events(evs[i])
        lift(dbs[i % len(dbs)])
    ops.set_stage_events(None)
    torch.cuda.synchronize()
    ms = {s: float(np.mean([evs[i][j].elapsed_time(evs[i][j + 1]) for i in range(K)])) for j, s in enumerate(STAGES)}
    # what a stage interval holds besides the kernel: two event records back to back, nothing between
    pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(64)]
    for a_, b_ in pairs:
        a_.record()
        b_.record()
    torch.cuda.synchronize()
    event_pair_us = float(np.median([a_.elapsed_time(b_) for a_, b_ in pairs])) * 1e3

    n = dbs[0].n
    levels = eng.last_levels[:n].cpu().numpy()      # of the last batch run
    last = dbs[(K - 1) % len(dbs)]
    roi_bytes = roi_align_algorithmic_bytes(last.boxes.cpu().numpy(), levels, out_elt=2)
    k_in = CHANNELS * POOL * POOL
    fc1_flops = 2.0 * n * k_in * FC_DIM
    fc2_flops = 2.0 * n * FC_DIM * FC_DIM
    kernels = {
        "roi_align": {"ms": ms["roi_align"], "bound": "hbm", "achieved": roi_bytes / (ms["roi_align"] * 1e-3) / 1e9,
                      "peak": peaks["hbm"], "unit": "GB/s", "algorithmic_bytes": roi_bytes,
                      "note": "algorithmic bytes = the sum of every box's own footprint + its output (SURVEY 8d); boxes of an "
                              "image overlap and the L2 serves the re-reads, so DRAM traffic (`traffic`, ncu) can be lower and "
                              "achieved can exceed the DRAM peak when many boxes share an image"},
        "fc1": {"ms": ms["fc1"], "bound": "tensor", "achieved": fc1_flops / (ms["fc1"] * 1e-3) / 1e12,
                "peak": peaks["tf_burst"], "unit": "TFLOP/s", "algorithmic_flops": fc1_flops,
                "note": "split-K tcgen05 GEMM; the CTA that completes a tile applies bias/ReLU (one launch)"},
        "roi_order": {"ms": ms["roi_order"], "note": "work-item planner of the ROIAlign kernel (one CTA)"},
    }
    if eng.tail_fused:
        tf32 = eng.tail_precision == "tf32"
        heads_flops = 2.0 * n * FC_DIM * 13
        kernels["fc_tail"] = {
            "ms": ms["fc_tail"], "bound": "tensor", "achieved": (fc2_flops + heads_flops) / (ms["fc_tail"] * 1e-3) / 1e12,
            "peak": peaks["tf_burst"] * (0.5 if tf32 else 1.0), "unit": "TFLOP/s",
            "algorithmic_flops": fc2_flops + heads_flops,
            "note": "fc2 (%s operands) + 13 head rows + decode in one launch; peak = %s" % (
                eng.tail_precision, "half the measured bf16 peak (kind::tf32 rate)" if tf32 else "measured bf16 peak")}
    else:
        kernels["fc2"] = {"ms": ms["fc2"], "bound": "tensor", "achieved": fc2_flops / (ms["fc2"] * 1e-3) / 1e12,
                          "peak": peaks["tf_burst"], "unit": "TFLOP/s", "algorithmic_flops": fc2_flops}
        kernels["head_decode"] = {"ms": ms["head_decode"]}
    for k in kernels.values():
        if "peak" in k:
            k["frac"] = k["achieved"] / k["peak"]
            # the same with the measured cost of an empty event-to-event interval taken off the stage: closer to the
            # kernel's own duration (compare the ncu launch list in profiles/); `frac` stays the conservative figure
            net = max(k["ms"] - event_pair_us * 1e-3, 1e-6)
            k["frac_net_of_event_overhead"] = k["frac"] * k["ms"] / net
    top = max(("roi_align", "fc1"), key=lambda s: ms[s])
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu capture
    if os.path.exists(tp):
        with open(tp) as f:
            cl_batch = not dbs[0].features[0].is_contiguous()
            traffic = json.load(f).get("roi_align_channels_last" if top == "roi_align" and cl_batch else top)
    t = kernels[top]
    roofline = {"kernel": top, "bound": t["bound"], "achieved": t["achieved"], "peak": t["peak"], "unit": t["unit"],
                "frac": t["frac"], "traffic": traffic, "peak_source": peaks["source"] + " (burst: kernel timed alone "
                "between its own events)"}
    return {"roofline": roofline, "kernels": kernels,
            "stage_timing": {"event_pair_overhead_us": event_pair_us, "stage_sum_us": sum(ms.values()) * 1e3,
                             "note": "kernels[*].ms are event-to-event intervals around each launch inside the step "
                                     "(launch latency and the event records included, nothing subtracted); an empty "
                                     "interval measures event_pair_overhead_us"}}


def run_e2e(head, names, inputs, device, K: int, W: int, world: int, lift) -> Dict[str, object]:
    """The metric through the plugin call a user of the reference makes (`_forward_cube`, roi_heads.py:329) with
    HOST inputs: per step H2D of that step's inputs from pinned memory, the lift, D2H of the detection records."""
    from ovmono3d_b200 import ops
    from ovmono3d_b200.compat import Boxes, Instances

    hbs = [HostBatch(x) for x in inputs]
    n = hbs[0].n
    dev_feats = [torch.empty_like(f, device=device) for f in hbs[0].features]
    rec_host = torch.empty((n, ops.RECORD_WIDTH), dtype=torch.float32).pin_memory()

    def instances_of(hb):
        """The caller's per-image containers (the 2-D detector's output in the reference, roi_heads_gdino.py:157-167),
        holding the pinned host tensors.  Built once per host batch: `_forward_cube` mutates them in place (it replaces
        `scores` and adds the pred_* fields, roi_heads.py:817-843), so a step only puts the input scores back."""
        out = []
        for i in range(len(hb.boxes)):
            inst = Instances(hb.im_dims[i])
            inst.pred_boxes = Boxes(hb.boxes[i])
            inst.pred_classes = hb.classes[i]
            inst.scores = hb.scores[i]
            out.append(inst)
        return out

    held = {id(hb): instances_of(hb) for hb in hbs}
    feats = dict(zip(names, dev_feats))
    stream = torch.cuda.current_stream()

    def step(hb, copy_features: bool):
        if copy_features:
            for d, h in zip(dev_feats, hb.features):
                d.copy_(h, non_blocking=True)
        insts = held[id(hb)]
        for inst, sc in zip(insts, hb.scores):
            inst._fields["scores"] = sc
        head._forward_cube(feats, insts, hb.Ks, hb.im_dims, hb.ratios)
        rec_host.copy_(head.last_records, non_blocking=True)
        stream.synchronize()                          # the caller reads the detections now
        return rec_host

    def timed(copy_features: bool) -> float:
        for i in range(W):
            step(hbs[i % len(hbs)], copy_features)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(K):
            step(hbs[i % len(hbs)], copy_features)
        e1.record()
        torch.cuda.synchronize()
        return _max_over_ranks(e0.elapsed_time(e1), world, device)

    ms_strict = timed(True)
    # deployment case: the pyramid of each step is already in HBM (written by the backbone); to keep the steps on
    # distinct data the pre-copied device batches are not reused here -- the last copied pyramid stays in place and
    # only the boxes/cameras change, which is what a detector-in-the-loop would see.
    ms_resident = timed(False)
    assert torch.isfinite(rec_host).all()
    d2h = n * ops.RECORD_WIDTH * 4
    return {
        "e2e": {"value": world * n * K / (ms_strict * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(hbs[0].feature_bytes + hbs[0].small_bytes), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": ms_strict / K,
                "what": "ROIHeads3DB200._forward_cube with the feature pyramid, boxes, classes, scores and cameras "
                        "copied from pinned host memory and the [n,48] records read back, every step"},
        "e2e_resident_features": {"value": world * n * K / (ms_resident * 1e-3), "unit": UNIT,
                                  "h2d_bytes_per_step": int(hbs[0].small_bytes), "d2h_bytes_per_step": int(d2h),
                                  "ms_per_step": ms_resident / K,
                                  "what": "same call with the pyramid left in HBM by the backbone (as in the "
                                          "reference, rcnn3d.py:97-109); boxes/classes/scores/cameras from the host"},
    }


# --------------------------------------------------------------------------------------------- CPU legs (oracle)
def _oracle_step(inp, weights, cfg):
    from oracle import lift_oracle as lo      # the checker, used here only as the timed CPU baseline

    with torch.no_grad():
        return lo.forward_cube(inp.features, inp.boxes, inp.classes, inp.scores, inp.Ks, inp.im_dims,
                               inp.im_scales_ratio, weights, cfg)


def _oracle_setup():
    from oracle import lift_oracle as lo

    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    torch.set_num_threads(cores)
    cfg = lo.LiftConfig(scales=tuple(1.0 / s for s in STRIDES), dims_priors_enabled=False)
    return cores, cfg, make_weights()


def cpu_baseline(inp, budget_s: float) -> Dict[str, object]:
    """The oracle (CPU restatement of the reference's fp32 path: torchvision roi_align + F.linear + eager decode)
    on the box's host cores, on a bounded sample: whole C2 batches until `budget_s` of CPU work is spent."""
    cores, cfg, weights = _oracle_setup()
    n = sum(len(b) for b in inp.boxes)
    _oracle_step(inp, weights, cfg)
    times = []
    t_end = time.perf_counter() + budget_s
    while len(times) < 3 or (time.perf_counter() < t_end and len(times) < 50):
        t0 = time.perf_counter()
        _oracle_step(inp, weights, cfg)
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {"value": n / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d whole %s batches (%d images x %d boxes, fp32) after 1 warm-up, best batch %.1f ms, "
                      "mean %.1f ms" % (len(times), CONFIG, NUM_IMAGES, BOXES_PER_IMAGE, best * 1e3,
                                        float(np.mean(times)) * 1e3)}


def run_reference(args) -> Optional[Dict[str, object]]:
    """The reference's own CPU implementation of the path (MODEL.DEVICE=cpu, fp32, all host threads) on our arm's
    workload.  Under torchrun only rank 0 works."""
    rank, world, _ = _dist_env()
    if rank != 0:
        return None
    K, W = args.steps, max(args.warmup, 1)
    cores, cfg, weights = _oracle_setup()
    inputs = [make_batch(SEED + i) for i in range(2)]
    n = sum(len(b) for b in inputs[0].boxes)
    for i in range(W):
        _oracle_step(inputs[i % 2], weights, cfg)
    t0 = time.perf_counter()
    for i in range(K):
        _oracle_step(inputs[i % 2], weights, cfg)
    dt = time.perf_counter() - t0
    value = n * K / dt
    sample = "each step = one whole %s batch (%d images x %d boxes) on %d host threads" % (
        CONFIG, NUM_IMAGES, BOXES_PER_IMAGE, cores)
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K,
        "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_dict(n),
        "method": {"device": "cpu",
                   "note": "the reference (pure Python on detectron2 + pytorch3d, both un-vendored) cannot be installed "
                           "here; this arm runs oracle/, its op-for-op CPU restatement pinned to vectors written by the "
                           "reference's own code (tests/golden/make_golden.py)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layout", choices=["nchw", "channels_last"], default="nchw",
                    help="memory format of the feature pyramid (default: the reference's NCHW)")
    ap.add_argument("--config", default="C2",
                    help="BASELINE.json config to run (default C2, the one the metric is quoted on): C1, C2, C3-1x2000, "
                         "C3-8x2000, C3-8x300, C4 (the C2 batch streamed, no L2 flush), C5-sfp4, C5-dla, C5-c128, C5-c512")
    args = ap.parse_args()
    global LAYOUT
    LAYOUT = args.layout
    set_workload(args.config)
    # exactly one JSON line on stdout: libraries that print there (NCCL's version banner under NCCL_DEBUG) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if STREAM and args.steps == 50 and args.impl != "reference":
        args.steps = max(20000 // (NUM_IMAGES * max(args.gpus, 1)), 1)     # the whole 20 000-image stream
    if args.impl == "reference":
        if args.steps == 50 and args.warmup == 5:
            args.steps, args.warmup = 10, 2
        out = run_reference(args)
    else:
        out = run_ours(args)
    rank = int(os.environ.get("RANK", "0"))
    sys.stdout.flush()
    if rank == 0 and out is not None:
        os.write(real_stdout, (json.dumps(out) + "\n").encode())


if __name__ == "__main__":
    main()
