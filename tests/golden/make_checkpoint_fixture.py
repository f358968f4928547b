"""Writes tests/golden/ref_blocks.pt: `torch.save`d state_dicts of the reference's OWN classes
(/root/reference/src/replicann/arch/transformer.py:87-201), so that the checkpoint-interchange test
(tests/test_checkpoint.py) can load a genuine reference `.pt` where /root/reference is not available.

    PYTHONPATH=/root/reference/src python tests/golden/make_checkpoint_fixture.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, "/root/reference/src")
from replicann.arch.transformer import TransformerCrossDecoder, TransformerDecoder, TransformerEncoder  # noqa: E402

torch.manual_seed(20261107)
H, E, CTX = 2, 16, 12
blocks = {
    "encoder": TransformerEncoder(H, E),
    "decoder": TransformerDecoder(H, E, context_size=CTX),
    "cross_decoder": TransformerCrossDecoder(H, E, context_size=CTX, head_bias=True),
}
out = {"config": {"n_heads": H, "embedding_size": E, "context_size": CTX},
       **{k: m.state_dict() for k, m in blocks.items()}}
path = Path(__file__).resolve().parent / "ref_blocks.pt"
torch.save(out, path)
print(path, path.stat().st_size, "bytes;", {k: len(v) for k, v in out.items() if k != "config"})
