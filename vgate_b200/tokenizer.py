"""Tokenizers for the b200 backend.

No tokenizer.json or weights exist offline, so the default is a synthetic byte-level tokenizer:
one UTF-8 byte = one token id (0..255).  Generated ids can be any vocabulary id, so the
detokeniser is *reversible* -- each token becomes ``"<id> "`` -- which lets token-id parity be
checked end to end through the gateway, where only ``text`` survives
(vgate/batcher.py:433-438 drops ``token_ids``).  A mounted model directory with a
``tokenizer.json`` is used through the ``tokenizers`` package instead.
"""
import os
from typing import List


class ByteTokenizer:
    eos_token_id = -1          # synthetic checkpoints never stop early

    def encode(self, text: str) -> List[int]:
        ids = list(text.encode("utf-8"))
        return ids if ids else [0]

    def decode_piece(self, token_id: int) -> str:
        return f"{int(token_id)} "

    def decode(self, ids) -> str:
        return "".join(self.decode_piece(t) for t in ids)

    @staticmethod
    def parse(text: str) -> List[int]:
        """Inverse of :meth:`decode` (for tests that only see ``text``)."""
        return [int(t) for t in text.split()]


class HFTokenizer:
    def __init__(self, model_dir: str):
        from tokenizers import Tokenizer
        self._tok = Tokenizer.from_file(os.path.join(model_dir, "tokenizer.json"))
        self.eos_token_id = -1
        import json
        cfg = os.path.join(model_dir, "generation_config.json")
        if os.path.exists(cfg):
            eos = json.load(open(cfg)).get("eos_token_id", -1)
            self.eos_token_id = eos[0] if isinstance(eos, list) else eos

    def encode(self, text: str) -> List[int]:
        return self._tok.encode(text).ids

    def decode(self, ids) -> str:
        return self._tok.decode(list(ids), skip_special_tokens=True)


class IncrementalDetokenizer:
    """Turns a growing id list into text deltas whose concatenation equals decode(all ids)."""

    def __init__(self, tok):
        self.tok, self.ids, self.emitted = tok, [], ""

    def push(self, new_ids) -> str:
        self.ids.extend(int(t) for t in new_ids)
        if isinstance(self.tok, ByteTokenizer):
            return self.tok.decode(new_ids)
        text = self.tok.decode(self.ids)
        if text.endswith("�"):          # incomplete multi-byte character: wait for more ids
            return ""
        delta, self.emitted = text[len(self.emitted):], text
        return delta


def load_tokenizer(model_id: str):
    if model_id and os.path.isdir(model_id) and os.path.exists(os.path.join(model_id, "tokenizer.json")):
        return HFTokenizer(model_id)
    return ByteTokenizer()
