"""Seeded GroundingDINO-output cases for the post-processing path, shared by `make_golden.py`
(which runs the reference's own `grounding_dino_inference_detector` on them) and the tests.

GroundingDINO and its BERT tokenizer are not available offline; the post-processing only needs (a) a tokenizer whose
caption tokenisation contains each phrase's tokens as a contiguous slice and (b) `pred_logits` / `pred_boxes` tensors.
`StandInTokenizer` splits on blanks, makes "." its own token and cuts long words into word pieces, which gives the
same structure ([CLS] phrase-tokens . phrase-tokens . [SEP]) as BERT's.
"""
from __future__ import annotations

import hashlib
import os
from dataclasses import dataclass
from typing import Dict, List, Tuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


class StandInTokenizer:
    CLS, SEP = 101, 102

    def __init__(self):
        self.vocab: Dict[str, int] = {}

    def _pieces(self, text: str) -> List[int]:
        out = []
        for word in text.replace(".", " . ").split():
            parts = [word] if len(word) <= 6 else [word[:4], "##" + word[4:]]
            for p in parts:
                out.append(self.vocab.setdefault(p, 1000 + len(self.vocab)))
        return out

    def __call__(self, text, add_special_tokens: bool = True):
        if isinstance(text, (list, tuple)):
            return {"input_ids": [self(t, add_special_tokens)["input_ids"] for t in text]}
        ids = self._pieces(text)
        return {"input_ids": [self.CLS] + ids + [self.SEP] if add_special_tokens else ids}


@dataclass(frozen=True)
class GdinoCase:
    name: str
    seed: int
    num_queries: int
    num_tokens: int
    categories: Tuple[str, ...]
    image_hw: Tuple[int, int]
    box_threshold: float = 0.001        # roi_heads_gdino.py:145
    nms_threshold: float = 0.5          # :272
    logit_shift: float = -4.0
    clusters: int = 60
    duplicate_rows: int = 8             # exact score ties (stable order)


OMNI = ("chair", "table", "cabinet", "car", "lamp", "books", "sofa", "pedestrian", "picture", "window", "pillow",
        "truck", "door", "blinds", "sink", "shelves", "television", "shoes", "cup", "bottle", "bookcase", "laptop",
        "desk", "cereal box", "floor mat", "traffic cone", "mirror", "barrier", "counter", "camera", "bicycle",
        "toilet", "bus", "bed", "refrigerator", "trailer", "box", "oven", "clothes", "van", "towel", "motorcycle",
        "night stand", "stove", "machine", "stationery", "bathtub", "cyclist", "curtain", "bin")

CASES = [
    GdinoCase("gdino_post_900", 11, 900, 256, OMNI, (896, 1184)),
    GdinoCase("gdino_post_small", 12, 37, 64, ("dining table", "cat", "refrigerator"), (512, 512), clusters=5,
              duplicate_rows=3),
    GdinoCase("gdino_post_none_pass", 13, 300, 256, OMNI[:7], (896, 896), logit_shift=-30.0),
    GdinoCase("gdino_post_all_pass_1024", 14, 1024, 256, OMNI[:20], (640, 1024), logit_shift=0.0, clusters=200),
    GdinoCase("gdino_post_threshold_0p3", 15, 900, 256, OMNI[:12], (1024, 1024), box_threshold=0.3,
              nms_threshold=0.3, logit_shift=-1.0, clusters=30),
]
CASE_BY_NAME = {c.name: c for c in CASES}


def golden_path(case: GdinoCase) -> str:
    return os.path.join(HERE, case.name + ".npz")


def build_case(case: GdinoCase):
    """-> (pred_logits [1,nq,T], pred_boxes [1,nq,4], tokenizer, caption, text_prompt)."""
    g = torch.Generator().manual_seed(case.seed)
    tok = StandInTokenizer()
    cap_list = list(case.categories)
    caption = " . ".join(cap_list).lower().strip()
    if not caption.endswith("."):
        caption += " ."
    n_tok = len(tok(caption)["input_ids"])
    assert n_tok <= case.num_tokens, (n_tok, case.num_tokens)
    nq, T = case.num_queries, case.num_tokens
    logits = torch.randn(nq, T, generator=g) * 2.0 + case.logit_shift
    hot = torch.randint(1, n_tok - 1, (nq,), generator=g)                  # every query likes one token
    logits[torch.arange(nq), hot] += torch.rand(nq, generator=g) * 6.0
    logits[:, n_tok:] = float("-inf")                                      # GroundingDINO masks the padding columns
    centres = torch.rand(case.clusters, 2, generator=g) * 0.8 + 0.1
    sizes = torch.exp(torch.rand(case.clusters, 2, generator=g) * (np.log(0.6) - np.log(0.03)) + np.log(0.03))
    which = torch.randint(0, case.clusters, (nq,), generator=g)
    boxes = torch.cat([centres[which] + (torch.rand(nq, 2, generator=g) - 0.5) * 0.04,
                       sizes[which] * torch.exp((torch.rand(nq, 2, generator=g) - 0.5) * 0.3)], dim=1)
    for k in range(min(case.duplicate_rows, nq // 2)):                     # exact ties: same logits, same box
        src, dst = 2 * k, nq - 1 - 3 * k
        logits[dst] = logits[src]
        boxes[dst] = boxes[src]
    return logits[None].contiguous(), boxes[None].contiguous().float(), tok, caption, [[c] for c in cap_list]


def input_checksum(logits: torch.Tensor, boxes: torch.Tensor) -> str:
    h = hashlib.sha256()
    h.update(logits.numpy().tobytes())
    h.update(boxes.numpy().tobytes())
    return h.hexdigest()
