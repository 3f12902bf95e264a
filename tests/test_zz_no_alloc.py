"""`ovm_roi_align_static`: ROIAlign with the planner's scratch in a file-scope device array of the library, for callers
that cannot allocate device memory.  No counterpart in the reference.

Collected last on purpose (the file name sorts after every other test file): a
failure here must not hide any other test.  Passes alone in a fresh process and after the other tests.

The new entry point is checked against the existing one (`ovm_roi_align` with a caller-allocated `order_ws`), through
the C ABI, on two calls in a row -- the second call finds the first one's items and headers in the shared array -- with
different boxes and a different box count, on both pyramid layouts.  The two entry points run the same kernels on the
same plan, so the results are compared exactly; the oracle tolerance of tests/test_lift_gpu.py (1e-3) is asserted on
top against torchvision's ROIAlign rules through the oracle.
"""
import ctypes as C
import gc

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-3      # tests/test_lift_gpu.py


def _inputs(seed, images, boxes_per_image, channels, dev):
    from ovmono3d_b200 import synth

    strides = (7, 14, 28)
    inp = synth.make_inputs(seed, images, boxes_per_image, strides=strides, canvas=448, channels=channels)
    counts = [len(b) for b in inp.boxes]
    boxes = torch.cat(inp.boxes).to(dev).contiguous()
    bidx = torch.from_numpy(np.repeat(np.arange(len(counts), dtype=np.int32), counts)).to(dev)
    return inp, boxes, bidx, [1.0 / s for s in strides]


@pytest.mark.parametrize("layout", ["nchw", "channels_last"])
def test_static_scratch_matches_allocating_entry_point_twice(layout):
    from oracle import thirdparty as tp
    from ovmono3d_b200 import ops

    dev = torch.device("cuda:0")
    lib = ops.load_library()
    for call, (seed, per_image) in enumerate([(11, 60), (12, 23)]):        # two calls in a row, second one smaller
        inp, boxes, bidx, scales = _inputs(seed, 2, per_image, 128, dev)
        feats = [f.to(dev) for f in inp.features]
        if layout == "channels_last":
            feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
        n = boxes.shape[0]
        want, want_lv = ops.roi_align(feats, scales, boxes, bidx, 7, 0, torch.float32)       # existing entry point
        # everything the new entry point touches is allocated here, by the caller, before the call
        pooler, keep = ops.make_pooler(feats, scales, 7, 0)
        out = torch.full((n, 128, 7, 7), float("nan"), device=dev)
        lv = torch.full((n,), -7, dtype=torch.int32, device=dev)
        gc.collect()                      # (tensors of earlier tests may be freed at any moment: count allocations, not bytes)
        torch.cuda.synchronize()
        allocs0 = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
        segs0 = torch.cuda.memory_stats(dev)["segment.all.allocated"]
        free0, _ = torch.cuda.mem_get_info(dev)
        ops.roi_align_static(pooler, boxes, bidx, out, lv)
        torch.cuda.synchronize()
        st = torch.cuda.memory_stats(dev)
        assert st["allocation.all.allocated"] == allocs0 and st["segment.all.allocated"] == segs0, \
            "torch allocated device memory during the call"
        if call == 1:     # (the first call may load kernel images lazily; by the second one everything is resident)
            assert torch.cuda.mem_get_info(dev)[0] >= free0, "device memory was allocated during the call"
        assert torch.equal(lv, want_lv), "call %d: levels differ" % call
        assert torch.equal(out, want), "call %d: pooled features differ from ovm_roi_align" % call
        ref = tp.ROIPoolerOracle(7, scales, 0, "ROIAlignV2")(inp.features, inp.boxes)
        err = float((out.cpu() - ref).abs().max() / ref.abs().max())
        assert err <= FP32_TOL, "call %d: %.3g against the oracle" % (call, err)
        del keep


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    from ovmono3d_b200 import ops

    dev = torch.device("cuda:0")
    lib = ops.load_library()
    feats = [torch.zeros(1, 8, 16, 16, device=dev)]
    pooler, keep = ops.make_pooler(feats, [1 / 16], 7, 0)
    n = ops.ROI_STATIC_MAX_BOXES + 1
    boxes = torch.zeros(n, 4, device=dev)
    bidx = torch.zeros(n, dtype=torch.int32, device=dev)
    out = torch.empty(n, 8, 7, 7, device=dev)
    with pytest.raises(ops.LiftError, match="at most %d" % ops.ROI_STATIC_MAX_BOXES):
        ops.roi_align_static(pooler, boxes, bidx, out)
    # the limit itself is accepted, straight through the C ABI (no Python wrapper in between)
    rc = lib.ovm_roi_align_static(C.byref(pooler), boxes.data_ptr(), bidx.data_ptr(), ops.ROI_STATIC_MAX_BOXES,
                                  out.data_ptr(), ops.F32, 8 * 49, None, ops._stream())
    assert rc == 0, lib.ovm_last_error()
    torch.cuda.synchronize()
    assert float(out[: ops.ROI_STATIC_MAX_BOXES].abs().max()) == 0.0
    del keep
