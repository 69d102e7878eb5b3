"""Question-answering text envs on the GPU: MetaMathQAEnv (ragen/env/metamathqa/env.py:10-69, config.py:4-10) and StaticEnv
(ragen/env/static/env.py:10-66, config.py:4-10, utils.py) behind the reference's env API, with the answer check as one
launch of ragen_text_score for a whole batch of answers (TextScoreBatch).

The questions come from Hugging Face datasets in the reference (`datasets.load_dataset`, a download); here the same call
is made when the `datasets` package and the data are available, and an in-memory list of rows can be passed instead
(`config.data`), which is what the tests do.  Host-side text handling that is not the scoring (picking the question,
extracting the label from the dataset's worked solution) is the reference's own few lines of `re`.
"""
import random
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _lib
from .base import BaseLanguageBasedEnv

CORRECT, VALID, TEXT_MATCH, UNSUPPORTED = 1, 2, 4, 128


@dataclass
class MetaMathQAEnvConfig:
    """ragen/env/metamathqa/config.py:4-10 + `data`."""
    dataset_path: str = "meta-math/MetaMathQA"
    cache_dir: str = "./data"
    split: str = "train"
    data: Optional[Dict[str, List[dict]]] = field(default=None, repr=False)  # {split: [{'query':..., 'response':...}]}


@dataclass
class StaticEnvConfig:
    """ragen/env/static/config.py:4-10 + `data`."""
    dataset_name: str = "metamathqa"  # metamathqa, gsm8k, mmlu
    cache_dir: str = "./data"
    split: Optional[str] = None
    data: Optional[Dict[str, List[dict]]] = field(default=None, repr=False)


class TextScoreBatch:
    """Scores a batch of (answer, label) pairs on the device; 1-step episodes own no device state."""

    def __init__(self, n=1, config=None, device=0):
        self.ctx = _lib.Context.get(device)
        self.device = torch.device("cuda", int(device))
        self.n, self.config = int(n), config

    def score(self, answers, labels, mode):
        """-> uint8 numpy [k] of result bits (ops.text_score)."""
        from .. import ops

        return ops.text_score(list(answers), list(labels), mode, self.device).cpu().numpy()


def _seeded_index(seed, n):
    """`with all_seed(seed): random.randint(0, n - 1)` (ragen/utils.py:7-18): the draw must not disturb the caller's streams."""
    if seed is None:
        return random.randint(0, n - 1)
    return random.Random(seed).randint(0, n - 1)  # random.seed(seed); random.randint(...) on a private generator


def _load(path_kwargs, cache_dir, data):
    if data is not None:
        return data
    try:
        from datasets import load_dataset
    except ImportError as e:  # the reference has the same hard dependency
        raise RuntimeError("this env needs the `datasets` package and its data; pass config.data instead") from e
    return load_dataset(**path_kwargs, cache_dir=cache_dir)


class MetaMathQAEnv(BaseLanguageBasedEnv):
    def __init__(self, config: Optional[MetaMathQAEnvConfig] = None, batch: Optional[TextScoreBatch] = None, slot: int = 0, **kwargs):
        BaseLanguageBasedEnv.__init__(self)
        self.config = config if config is not None else MetaMathQAEnvConfig()
        self.dataset = _load({"path": self.config.dataset_path}, self.config.cache_dir, self.config.data)
        self.current_question_idx = self.current_question = self.correct_answer = self.step_num = self.render_cache = None
        self.batch = batch if batch is not None else TextScoreBatch(1, self.config, device=kwargs.get("device", 0))
        self.slot = slot

    def _extract_answer(self, response):
        match = re.search(r"The answer is: (.*?)$", response, re.DOTALL)
        return match.group(1).strip() if match else None

    def reset(self, seed=None, mode=None):
        dataset = self.dataset[self.config.split]
        self.current_question_idx = _seeded_index(seed, len(dataset))
        question_data = dataset[self.current_question_idx]
        self.current_question = question_data["query"]
        self.correct_answer = self._extract_answer(question_data["response"])
        self.step_num = 0
        self.render_cache = self.current_question
        return self.render_cache

    def _check_answer(self, user_answer):
        bits = int(self.batch.score([user_answer], [self.correct_answer or ""], "metamathqa")[0])
        return bool(bits & CORRECT), bool(bits & VALID)

    def finish_step(self, is_correct, is_valid):
        """env.py:42-53 from the device result."""
        reward = 1.0 / (2 ** self.step_num) if is_correct else 0.0
        observation, done = ("Correct!", True) if is_correct else ("Incorrect. Please think again.", False)
        self.step_num += 1
        self.render_cache = observation
        return observation, reward, done, {"action_is_valid": is_valid, "success": is_correct}

    def step(self, action):
        return self.finish_step(*self._check_answer(action))

    def render(self, mode=None):
        return self.render_cache

    def close(self):
        pass


def _extract_answer_from_text(text):
    """ragen/env/static/utils.py:14-31."""
    for pattern in (r"The answer is:?\s*(.*?)(?:\n|$)", r"Answer:?\s*(.*?)(?:\n|$)", r"Final answer:?\s*(.*?)(?:\n|$)",
                    r"Therefore,\s*(.*?)(?:\n|$)", r"Thus,\s*(.*?)(?:\n|$)"):
        match = re.search(pattern, text, re.DOTALL)
        if match:
            return match.group(1).strip()
    return text.strip().split("\n")[-1].strip()


def _process_mmlu(item):
    choices = [item["choices"][i] for i in range(len(item["choices"]))]
    question = item["question"] + "\n" + "\n".join(f"{chr(65 + i)}. {choice}" for i, choice in enumerate(choices))
    return question, chr(65 + item["answer"])


# ragen/env/static/utils.py:134-180: dataset -> (load_dataset arguments, processor, scorer)
REGISTERD_STATIC_ENV = {
    "metamathqa": {"config": {"path": "meta-math/MetaMathQA"},
                   "processor": lambda item: (item["query"], _extract_answer_from_text(item["response"])), "compute_score": "exact_match"},
    "gsm8k": {"config": {"path": "openai/gsm8k", "name": "main"},
              "processor": lambda item: (item["question"], item["answer"].split("####")[1].strip().lower()), "compute_score": "numeric"},
    "mmlu": {"config": {"path": "cais/mmlu", "name": "abstract_algebra"}, "processor": _process_mmlu, "compute_score": "multiple_choice"},
}


class StaticEnv(BaseLanguageBasedEnv):
    def __init__(self, config: Optional[StaticEnvConfig] = None, batch: Optional[TextScoreBatch] = None, slot: int = 0, **kwargs):
        BaseLanguageBasedEnv.__init__(self)
        self.config = config if config is not None else StaticEnvConfig()
        entry = REGISTERD_STATIC_ENV[self.config.dataset_name]
        dataset_config = getattr(self.config, "dataset_config", None) or entry["config"]
        self.dataset = _load(dataset_config, self.config.cache_dir, self.config.data)
        self.split = list(self.dataset.keys())[0] if self.config.split is None else self.config.split
        self.current_question_idx = self.current_question = self.correct_answer = self.step_num = None
        self.processor, self.score_mode = entry["processor"], entry["compute_score"]
        self.batch = batch if batch is not None else TextScoreBatch(1, self.config, device=kwargs.get("device", 0))
        self.slot = slot

    def reset(self, seed=None, mode=None):
        dataset_split = self.dataset[self.split]
        self.current_question_idx = _seeded_index(seed, len(dataset_split))
        self.current_question, self.correct_answer = self.processor(dataset_split[self.current_question_idx])
        self.step_num = 0
        return self.current_question

    def compute_score(self, action, label):
        bits = int(self.batch.score([action], [label], self.score_mode)[0])
        return {"is_correct": bool(bits & CORRECT), "is_valid": bool(bits & VALID)}

    def finish_step(self, is_correct, is_valid):
        """env.py:46-66 from the device result."""
        reward = 1.0 / (2 ** self.step_num) if is_correct else 0.0
        observation, done = ("Correct!", True) if is_correct else ("Incorrect. Please think again.", False)
        self.step_num += 1
        return observation, reward, done, {"success": is_correct, "is_valid": is_valid}

    def step(self, action):
        r = self.compute_score(action, self.correct_answer)
        return self.finish_step(r["is_correct"], r["is_valid"])

    def render(self, mode=None):
        return self.current_question

    def close(self):
        pass
