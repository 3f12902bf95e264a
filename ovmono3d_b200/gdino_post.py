"""GroundingDINO post-processing on the GPU: raw `pred_logits` / `pred_boxes` -> the lift's prompt boxes.

Host-side mirror of the reference's `get_grounding_output` / `grounding_dino_inference_detector`
(cubercnn/modeling/roi_heads/roi_heads_gdino.py:174-202, 262-281) for the `token_spans is None` mode the reference
uses (:147).  The string work (caption, tokenizer, phrase token columns :300-311) stays on the host and is cached per
category list; everything numeric -- sigmoid, phrase sums, threshold, max / argmax, cxcywh -> xyxy pixels, NMS 0.5,
gather -- is one launch of `ovm_gdino_post`, and the surviving boxes / classes / scores stay on the device for
`_forward_cube` (the reference bounces through `.cpu()` three times, :199-202,230).  GroundingDINO itself is not part
of this package: `model` is any callable with its interface (`model(image[None], captions=[caption])` ->
{"pred_logits" [1,nq,T], "pred_boxes" [1,nq,4]}, and a `.tokenizer`).
"""
from __future__ import annotations

from typing import Any, Dict, List, Sequence, Tuple

import torch

from . import ops


def build_caption(cap_list: Sequence[str]) -> str:
    """roi_heads_gdino.py:176-181."""
    caption = " . ".join(cap_list).lower().strip()
    if not caption.endswith("."):
        caption = caption + " ."
    return caption


def phrase_token_ranges(tokenizer: Any, cap_list: Sequence[str], caption: str) -> torch.Tensor:
    """[P,2] int32 (begin, end) token columns of every phrase in the tokenised caption (roi_heads_gdino.py:300-311);
    raises AssertionError like the reference when a phrase does not tokenise to its slice of the caption."""
    tokenized = tokenizer(caption)
    tokenized_phrases = tokenizer(list(cap_list), add_special_tokens=False)["input_ids"]
    begin_id = 1
    ranges: List[Tuple[int, int]] = []
    for phrase_tokens in tokenized_phrases:
        end_id = begin_id + len(phrase_tokens)
        assert list(phrase_tokens) == list(tokenized["input_ids"][begin_id:end_id]), "assert error!!!"
        ranges.append((begin_id, end_id))
        begin_id = end_id + 1
    return torch.tensor(ranges, dtype=torch.int32).reshape(-1, 2)


def first_occurrence_lut(cap_list: Sequence[str]):
    """None when every phrase is unique, else an int64 table phrase index -> index of its first occurrence: the class
    the reference assigns through `filtered_texts.index([class_name])` (roi_heads_gdino.py:162)."""
    first = [cap_list.index(c) for c in cap_list]
    if first == list(range(len(cap_list))):
        return None
    return torch.tensor(first, dtype=torch.int64)


class GroundingDinoPostB200:
    """Callable with the signature `ROIHeads3DGDINOB200` expects from its `detector`:
    `detector(image_bgr [3,H,W], text_prompt [[cat], ...]) -> {"labels", "bboxes", "scores", "pred_classes"}`.
    "bboxes" / "scores" / "pred_classes" are CUDA tensors in the reference's order (decreasing score)."""

    def __init__(self, model: Any, box_threshold: float = 0.001, nms_threshold: float = 0.5):
        self.model = model
        self.box_threshold = float(box_threshold)      # roi_heads_gdino.py:145
        self.nms_threshold = float(nms_threshold)      # :272
        self._ranges: Dict[Tuple[str, ...], Tuple[str, torch.Tensor]] = {}

    def ranges_for(self, cap_list: Sequence[str], device: torch.device) -> Tuple[str, torch.Tensor]:
        key = tuple(cap_list)
        hit = self._ranges.get(key)
        if hit is None or hit[1].device != device:
            caption = build_caption(cap_list)
            hit = (caption, phrase_token_ranges(self.model.tokenizer, cap_list, caption).to(device))
            self._ranges[key] = hit
        return hit

    def postprocess(self, pred_logits: torch.Tensor, pred_boxes: torch.Tensor, phrase_ranges: torch.Tensor,
                    image_hw: Tuple[int, int]):
        """[B,nq,T], [B,nq,4] -> per image (boxes [k,4], scores [k], classes [k] int64, query [k] int32)."""
        boxes, scores, classes, query, count = ops.gdino_post(pred_logits, pred_boxes, phrase_ranges, image_hw,
                                                              self.box_threshold, self.nms_threshold)
        counts = count.tolist()                         # the one host read: the lift's launch sizes need it
        return [(boxes[b, :k], scores[b, :k], classes[b, :k], query[b, :k]) for b, k in enumerate(counts)]

    def __call__(self, image: torch.Tensor, text_prompt: List[List[str]]) -> Dict[str, Any]:
        if not image.is_cuda:
            raise ops.LiftError("GroundingDinoPostB200 needs the image on a CUDA device (no CPU path)")
        cap_list = [c[0] for c in text_prompt]
        caption, ranges = self.ranges_for(cap_list, image.device)
        with torch.no_grad():
            outputs = self.model(image[None], captions=[caption])
        h, w = image.shape[1:]
        (boxes, scores, classes, _), = self.postprocess(outputs["pred_logits"][:1].float(),
                                                        outputs["pred_boxes"][:1].float(), ranges, (h, w))
        lut = first_occurrence_lut(cap_list)
        if lut is not None:      # a category listed twice: the reference's `filtered_texts.index([name])` (:162) finds the first
            classes = lut.to(classes.device)[classes]
        return {"bboxes": boxes, "scores": scores, "pred_classes": classes,
                "labels": _LazyLabels(cap_list, classes)}


class _LazyLabels:
    """`ov_pred_instances["labels"]` (phrase per detection): built from the class indices only if somebody reads it."""

    def __init__(self, cap_list: Sequence[str], classes: torch.Tensor):
        self._caps, self._classes, self._list = list(cap_list), classes, None

    def _get(self) -> List[str]:
        if self._list is None:
            self._list = [self._caps[i] for i in self._classes.tolist()]
        return self._list

    def __iter__(self):
        return iter(self._get())

    def __len__(self) -> int:
        return int(self._classes.numel())

    def __getitem__(self, i):
        return self._get()[i]
