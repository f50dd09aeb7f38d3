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


_ROLES = ("System", "User", "Assistant", "Tool")


def qwen_chat_prompt(flat: str) -> str:
    """V-Gate flattens a chat into ``"Role: content\n...\nAssistant:"`` (main.py:269-271) and the reference passes
    that string to vLLM as a raw prompt (vllm_backend.py:113), which is what this backend does by default.  With a real
    Qwen2.5-Instruct checkpoint the model expects its chat template; ``VGATE_B200_CHAT_TEMPLATE=qwen`` rewrites the
    flat prompt into it (the ChatML form of Qwen2.5's ``tokenizer_config.json``):
    ``<|im_start|>role\ncontent<|im_end|>\n`` per message, then ``<|im_start|>assistant\n``.  A line starts a new
    message only when it begins with a capitalised role name followed by ": ", so message bodies keep their newlines;
    a prompt that does not end in the reference's ``"\nAssistant:"`` marker is returned unchanged."""
    tail = "\nAssistant:"
    if not flat.endswith(tail):
        return flat
    msgs = []
    for line in flat[:-len(tail)].split("\n"):
        head, sep, rest = line.partition(": ")
        if sep and head in _ROLES:
            msgs.append([head.lower(), rest])
        elif msgs:
            msgs[-1][1] += "\n" + line
        else:
            return flat                    # no role on the first line: not the reference's format
    if not msgs:
        return flat
    if msgs[0][0] != "system":             # Qwen2.5's template inserts its default system message
        msgs.insert(0, ["system", "You are Qwen, created by Alibaba Cloud. You are a helpful assistant."])
    return "".join(f"<|im_start|>{r}\n{c}<|im_end|>\n" for r, c in msgs) + "<|im_start|>assistant\n"


def load_tokenizer(model_id: str):
    if model_id and os.path.isdir(model_id) and os.path.exists(os.path.join(model_id, "tokenizer.json")):
        return HFTokenizer(model_id)
    return ByteTokenizer()
