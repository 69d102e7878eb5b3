"""TEST INFRASTRUCTURE — a byte-level stand-in for the Qwen tokenizer with the members ContextManager reaches
(ragen/llm_agent/ctx_manager.py:24-33, 269, 278, 334): name_or_path, encode, apply_chat_template, __call__ with left
padding, batch_decode.  The reference's own test uses the same device (tests/llm_agent/test_context_window.py:6-21).
Both the reference (tests/golden/make_golden.py) and ragen_b200 are driven with this one class, so token ids agree."""
import re
from types import SimpleNamespace

import torch

SPECIALS = {"<|endoftext|>": 151643, "<|im_start|>": 151644, "<|im_end|>": 151645}
_SPLIT = re.compile("(" + "|".join(re.escape(s) for s in SPECIALS) + ")")


class FakeQwenTokenizer:
    name_or_path = "qwen-fake-bytes"
    pad_token_id = SPECIALS["<|endoftext|>"]
    IM_START, IM_END = SPECIALS["<|im_start|>"], SPECIALS["<|im_end|>"]

    def encode(self, text):
        ids = []
        for part in _SPLIT.split(text):
            if part in SPECIALS:
                ids.append(SPECIALS[part])
            else:
                ids.extend(part.encode("utf-8"))
        return ids

    def apply_chat_template(self, messages, add_generation_prompt=False, tokenize=False):
        assert not tokenize
        text = "".join(f"<|im_start|>{m['role']}\n{m['content']}<|im_end|>\n" for m in messages)
        return text + ("<|im_start|>assistant\n" if add_generation_prompt else "")

    def __call__(self, texts, return_tensors="pt", padding=True, padding_side="left", truncation=False):
        assert return_tensors == "pt" and padding and padding_side == "left" and not truncation
        rows = [self.encode(t) for t in texts]
        L = max(len(r) for r in rows)
        ids = torch.full((len(rows), L), self.pad_token_id, dtype=torch.long)
        att = torch.zeros((len(rows), L), dtype=torch.long)
        for i, r in enumerate(rows):
            ids[i, L - len(r):] = torch.tensor(r, dtype=torch.long)
            att[i, L - len(r):] = 1
        return SimpleNamespace(input_ids=ids, attention_mask=att)

    def batch_decode(self, ids, skip_special_tokens=True):
        out = []
        inv = {v: k for k, v in SPECIALS.items()}
        for row in ids.tolist():
            parts, run = [], bytearray()
            for t in row:
                if t < 256:
                    run.append(t)
                    continue
                parts.append(run.decode("utf-8", errors="replace"))
                run = bytearray()
                if not skip_special_tokens:
                    parts.append(inv[t])
            parts.append(run.decode("utf-8", errors="replace"))
            out.append("".join(parts))
        return out


class ScriptedWg:
    """A deterministic stand-in for the actor worker group: the "policy" derives its answer from a CRC of the prompt
    text, so the reference managers and ragen_b200's managers get the same responses as long as their prompts agree.
    `proto` is the DataProto class of whichever side is being driven."""
    WORDS = ["Up", "Down", "Left", "Right", "Phoenix", "Dragon", "Jump", "up", "LEFT", "dragon"]

    def __init__(self, tokenizer, proto, salt=0):
        self.tokenizer, self.proto, self.salt = tokenizer, proto, salt
        self.calls = 0

    def generate_sequences(self, lm_inputs):
        import zlib

        import numpy as np

        texts = self.tokenizer.batch_decode(lm_inputs.batch["input_ids"], skip_special_tokens=False)
        out = []
        for t in texts:
            h = zlib.crc32(t.replace("<|endoftext|>", "").encode()) ^ self.salt ^ (self.calls * 2654435761 & 0xFFFFFFFF)
            rng = np.random.default_rng(h)
            shape = int(rng.integers(0, 12))
            if "Phoenix" in t or "Dragon" in t:
                acts = [self.WORDS[int(rng.integers(4, 10))]]
            else:
                acts = [self.WORDS[int(rng.integers(0, 10))] if rng.random() < 0.85 else self.WORDS[int(rng.integers(0, 4))]
                        for _ in range(int(rng.integers(0, 8)))]
            body = " || ".join(acts)
            if shape == 0:
                out.append("I forgot the tags " + body)
            elif shape == 1:
                out.append(f"plan</think> junk <answer>{body}</answer>")
            elif shape == 2:
                out.append(f" plan <|im_end|></think>\n<answer> {body} </answer> trailing")
            else:
                out.append(f"plan {shape}</think><answer>{body}</answer>")
        self.calls += 1
        res = self.proto()
        res.batch = None
        res.non_tensor_batch = {"response_texts": out, "env_ids": lm_inputs.non_tensor_batch["env_ids"],
                                "group_ids": lm_inputs.non_tensor_batch["group_ids"]}
        res.meta_info = lm_inputs.meta_info
        return res


def make_bpe_tokenizer(vocab_size=600, seed=0):
    """A REAL byte-level BPE tokenizer (the `tokenizers` library) with Qwen2's pre-tokenizer regex and ChatML special tokens,
    trained here on a synthetic corpus of prompts: its merges cross glyph boundaries ("##", "#_", "__#\\n" ...) the way a
    production tokenizer's do, which the byte-level stand-in above cannot show.  Same duck type as FakeQwenTokenizer plus the
    batched call with add_special_tokens=False that Hugging Face tokenizers offer."""
    import numpy as np
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers

    qwen_split = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
    tk = Tokenizer(models.BPE())
    tk.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(qwen_split), behavior="isolated"),
                                                pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tk.decoder = decoders.ByteLevel()
    rng = np.random.default_rng(seed)
    glyphs = "#_OX√PS"
    corpus = []
    for _ in range(400):
        W = int(rng.integers(4, 11))
        rows = ["".join(glyphs[int(c)] for c in rng.choice(7, W, p=[0.45, 0.35, 0.04, 0.04, 0.04, 0.04, 0.04])) for _ in range(W)]
        corpus.append("State:\n" + "\n".join(rows) + "\nYou have 7 actions left. Always output: <think> [Your thoughts] </think> "
                      "<answer> [your answer] </answer> with no extra text. Strictly follow this format.")
        corpus.append("You are solving the Sokoban puzzle. Reward:\n-0.1\n\nTurn 3:\n<think>plan</think><answer>Up || Down || Left</answer>")
    trainer = trainers.BpeTrainer(vocab_size=vocab_size, special_tokens=list(SPECIALS), initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                  show_progress=False)
    tk.train_from_iterator(corpus, trainer)
    return BpeChatTokenizer(tk)


class BpeChatTokenizer:
    name_or_path = "qwen-bpe-test"

    def __init__(self, tk):
        self.tk = tk
        self.pad_token_id = tk.token_to_id("<|endoftext|>")
        self.IM_START, self.IM_END = tk.token_to_id("<|im_start|>"), tk.token_to_id("<|im_end|>")

    def encode(self, text):
        return self.tk.encode(text, add_special_tokens=False).ids

    apply_chat_template = FakeQwenTokenizer.apply_chat_template

    def __call__(self, texts, return_tensors=None, padding=False, padding_side="left", truncation=False, add_special_tokens=True):
        rows = [e.ids for e in self.tk.encode_batch(list(texts), add_special_tokens=add_special_tokens)]
        if return_tensors != "pt":
            return {"input_ids": rows}
        L = max(len(r) for r in rows)
        ids = torch.full((len(rows), L), self.pad_token_id, dtype=torch.long)
        att = torch.zeros((len(rows), L), dtype=torch.long)
        for i, r in enumerate(rows):
            ids[i, L - len(r):] = torch.tensor(r, dtype=torch.long)
            att[i, L - len(r):] = 1
        return SimpleNamespace(input_ids=ids, attention_mask=att)

    def batch_decode(self, ids, skip_special_tokens=True):
        return self.tk.decode_batch([list(map(int, r)) for r in ids.tolist()], skip_special_tokens=skip_special_tokens)
