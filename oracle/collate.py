"""TEST INFRASTRUCTURE — the reference's batch collation for the device-collate parity tests.

``reference_functions()`` extracts ``_collate_frames`` (fairseq/data/audio/speech_to_text_dataset.py:267-285) and
``collate_tokens`` (fairseq/data/data_utils.py:33-67) from the reference's files BY SOURCE (ast) and executes exactly
those function bodies — the modules themselves cannot be imported here (they pull in the whole fairseq package).  Only
available where /root/reference is mounted; ``collate_frames`` / ``collater_order`` below are the restatements used
everywhere else, pinned to the extracted functions by tests/golden/collate.npz (oracle/make_golden_collate.py).
"""
from __future__ import annotations

import ast
import os
from typing import Dict, List

import torch

REF = os.environ.get("S2T_REFERENCE", "/root/reference")


def reference_functions() -> Dict[str, object]:
    out = {}
    for rel, name in (("fairseq/data/audio/speech_to_text_dataset.py", "_collate_frames"),
                      ("fairseq/data/data_utils.py", "collate_tokens")):
        src = open(os.path.join(REF, rel)).read()
        fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == name)
        ns = {"torch": torch, "List": List}
        exec(compile(ast.Module(body=[fn], type_ignores=[]), rel, "exec"), ns)
        out[name] = ns[name]
    return out


def collate_frames(frames: List[torch.Tensor], is_audio_input: bool = False) -> torch.Tensor:
    """speech_to_text_dataset.py:267-285: zero-padded [B, max_len(, F)]."""
    max_len = max(f.size(0) for f in frames)
    shape = (len(frames), max_len) if is_audio_input else (len(frames), max_len, frames[0].size(1))
    out = frames[0].new_zeros(shape)
    for i, v in enumerate(frames):
        out[i, : v.size(0)] = v
    return out


def collater_order(frames: List[torch.Tensor]):
    """speech_to_text_dataset.py:420-423: (sorted n_frames, order) of ``n_frames.sort(descending=True)``."""
    n_frames = torch.tensor([f.size(0) for f in frames], dtype=torch.long)
    return n_frames.sort(descending=True)
