"""ORACLE (test infrastructure only) -- CPU restatement of the reference's GroundingDINO post-processing, the step
right before the lift, with the same torch / torchvision calls in the same order:

  get_grounding_output             cubercnn/modeling/roi_heads/roi_heads_gdino.py:174-202 (token_spans is None)
  get_phrase_logits_from_token_logits  :296-317
  grounding_dino_inference_detector    :262-281
  box_cxcywh_to_xyxy                   :283-287

and of `torchvision.ops.nms` / `batched_nms` (CPU kernels of the installed torchvision, the library the reference
calls at :6,272 and, through d2's `batched_nms`, at fast_rcnn.py:105).

PARITY: pinned to `tests/golden/gdino_post_*.npz`, written by the reference's own
`grounding_dino_inference_detector` driven with a stand-in model / tokenizer (tests/golden/make_golden.py).
Only `tests/`, `__graft_entry__.smoke()` and the CPU legs of `bench.py` may import this module.
"""
from typing import Dict, List, Sequence, Tuple

import torch
from torchvision.ops import batched_nms as tv_batched_nms
from torchvision.ops import nms as tv_nms


def build_caption(cap_list: Sequence[str]) -> str:
    """roi_heads_gdino.py:176-181."""
    caption = " . ".join(cap_list)
    caption = caption.lower()
    caption = caption.strip()
    if not caption.endswith("."):
        caption = caption + " ."
    return caption


def phrase_token_ids(tokenized_caption_ids: List[int], tokenized_phrases: List[List[int]]) -> List[List[int]]:
    """roi_heads_gdino.py:300-311: token columns of each phrase inside the tokenised caption."""
    begin_id = 1
    ids = list(range(len(tokenized_caption_ids)))
    phrases_ids = []
    for phrase_tokens in tokenized_phrases:
        end_id = begin_id + len(phrase_tokens)
        assert phrase_tokens == tokenized_caption_ids[begin_id:end_id], "assert error!!!"
        phrases_ids.append(ids[begin_id:end_id])
        begin_id = end_id + 1
    return phrases_ids


def phrase_logits_from_token_logits(token_logits: torch.Tensor, phrases_ids: List[List[int]]) -> torch.Tensor:
    """roi_heads_gdino.py:312-317."""
    phrase_logits = []
    for phrase_ids in phrases_ids:
        phrase_logits.append(token_logits[:, phrase_ids].sum(dim=-1))
    return torch.stack(phrase_logits, dim=1)


def gdino_post(pred_logits: torch.Tensor, pred_boxes: torch.Tensor, phrases_ids: List[List[int]],
               image_hw: Tuple[int, int], box_threshold: float = 0.001, nms_threshold: float = 0.5
               ) -> Dict[str, torch.Tensor]:
    """One image: pred_logits [nq, T] raw, pred_boxes [nq, 4] cxcywh normalised -> boxes / scores / classes / query
    of the detections the reference hands to `_forward_cube`, in its order."""
    logits = pred_logits.sigmoid()                                  # :186
    boxes = pred_boxes
    phrases_logits = phrase_logits_from_token_logits(logits, phrases_ids)
    filt_mask = phrases_logits.max(dim=1)[0] > box_threshold        # :195
    im_logits_filt = phrases_logits[filt_mask]
    boxes_filt = boxes[filt_mask].cpu()
    query = torch.nonzero(filt_mask).flatten()
    im_pred_scores, im_pred_classes = im_logits_filt.max(dim=-1)    # :199
    all_logits = im_pred_scores.cpu()
    h, w = image_hw
    boxes_filt = boxes_filt * torch.tensor([w, h, w, h])            # :270
    x_c, y_c, bw, bh = boxes_filt.unbind(-1)
    boxes_filt = torch.stack([(x_c - 0.5 * bw), (y_c - 0.5 * bh), (x_c + 0.5 * bw), (y_c + 0.5 * bh)], dim=-1)
    nms_idx = tv_nms(boxes_filt, all_logits, nms_threshold)         # :272
    return {"boxes": boxes_filt[nms_idx], "scores": all_logits[nms_idx], "classes": im_pred_classes[nms_idx],
            "query": query[nms_idx]}


def nms(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float) -> torch.Tensor:
    return tv_nms(boxes.float().cpu(), scores.cpu(), iou_threshold)


def batched_nms(boxes: torch.Tensor, scores: torch.Tensor, idxs: torch.Tensor, iou_threshold: float) -> torch.Tensor:
    """d2 layers/nms.py `batched_nms`: `box_ops.batched_nms(boxes.float(), scores, idxs, iou_threshold)`."""
    return tv_batched_nms(boxes.float().cpu(), scores.cpu(), idxs.cpu(), iou_threshold)
